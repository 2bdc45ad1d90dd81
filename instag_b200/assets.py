"""On-disk formats of trained InsTaG heads, so that real assets can be fed to
the rasterizer (tests, bench.py --cloud) instead of synthetic clouds.  Host-side only.

* `.ply` point clouds as written by `GaussianModel.save_ply` / `save_deformed_ply`
  (scene/gaussian_model.py:429-484): binary little-endian, one `vertex` element of float32
  properties in the order of `construct_list_of_attributes` (:429-441):
  `x y z nx ny nz f_dc_0.. f_rest_0.. opacity scale_0.. rot_0..`; `f_dc`/`f_rest` are the
  [P,K,3] feature tensors transposed to [P,3,K] and flattened (:448-449), `load_ply` undoes that
  (:486-527).  The reference goes through the `plyfile` package; the format is restated here.
* checkpoints: the tuple `GaussianModel.capture()` returns (scene/gaussian_model.py:115-132), saved
  by torch.save inside `chkpnt_*.pth` (train_face.py) / `chkpnt_fuse_latest.pth`
  (synthesize_fuse.py:108).
"""
import numpy as np

CAPTURE_FIELDS = ("active_sh_degree", "xyz", "features_dc", "features_rest", "identity", "scaling", "rotation",
                  "opacity", "max_radii2D", "xyz_gradient_accum", "denom", "optimizer_state", "spatial_lr_scale",
                  "neural_renderer_state", "neural_motion_grid_state")  # scene/gaussian_model.py:115-132


def attribute_names(n_dc, n_rest, n_scale=3, n_rot=4):
    """construct_list_of_attributes (scene/gaussian_model.py:429-441)."""
    l = ["x", "y", "z", "nx", "ny", "nz"]
    l += [f"f_dc_{i}" for i in range(n_dc)]
    l += [f"f_rest_{i}" for i in range(n_rest)]
    l.append("opacity")
    l += [f"scale_{i}" for i in range(n_scale)]
    l += [f"rot_{i}" for i in range(n_rot)]
    return l


def _np(t):
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    return np.asarray(t, dtype=np.float32)


def write_gaussian_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation):
    """save_ply (scene/gaussian_model.py:443-461).  features_dc [P,1,3], features_rest [P,K,3] (the
    GaussianModel layout), opacity [P,1], scaling [P,3] (raw), rotation [P,4] (raw)."""
    xyz = _np(xyz)
    P = xyz.shape[0]
    f_dc = np.transpose(_np(features_dc), (0, 2, 1)).reshape(P, -1)      # :448
    f_rest = np.transpose(_np(features_rest), (0, 2, 1)).reshape(P, -1)  # :449
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, _np(opacity).reshape(P, 1), _np(scaling), _np(rotation)]
    table = np.ascontiguousarray(np.concatenate(cols, axis=1), dtype="<f4")                # :458
    names = attribute_names(f_dc.shape[1], f_rest.shape[1], cols[5].shape[1], cols[6].shape[1])
    assert table.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(table.tobytes())


_PLY_TYPES = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
              "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4"}


def read_ply_vertices(path):
    """Structured array of the first element of a PLY file (binary little/big endian or ascii; scalar
    properties only, which is all the reference writes)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, in_first, seen = None, None, [], False, 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: header without end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                seen += 1
                in_first = seen == 1
                if in_first:
                    count = int(tok[2])
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list properties are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or count is None:
            raise ValueError(f"{path}: incomplete header")
        if fmt == "ascii":
            rows = [f.readline().split() for _ in range(count)]
            out = np.empty(count, dtype=[(n, t) for n, t in props])
            for i, (n, _) in enumerate(props):
                out[n] = [r[i] for r in rows]
            return out
        end = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, end + t) for n, t in props])
        data = np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)
        return data


def read_gaussian_ply(path, max_sh_degree=None):
    """load_ply (scene/gaussian_model.py:486-527).  Returns float32 arrays in the GaussianModel
    layout: xyz [P,3], features_dc [P,1,3], features_rest [P,K,3], opacity [P,1], scaling [P,3],
    rotation [P,4], and the SH degree the file holds."""
    v = read_ply_vertices(path)
    names = v.dtype.names
    xyz = np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32)
    P = xyz.shape[0]
    numbered = lambda prefix: sorted((n for n in names if n.startswith(prefix)), key=lambda n: int(n.split("_")[-1]))
    f_dc = np.stack([v[n] for n in numbered("f_dc_")], axis=1).astype(np.float32).reshape(P, 3, -1)
    rest = numbered("f_rest_")
    degree = int(round(np.sqrt((len(rest) + 3) / 3))) - 1
    if max_sh_degree is not None:
        assert len(rest) == 3 * (max_sh_degree + 1) ** 2 - 3, "f_rest count does not match max_sh_degree"  # :502
        degree = max_sh_degree
    f_rest = (np.stack([v[n] for n in rest], axis=1).astype(np.float32) if rest else np.zeros((P, 0), np.float32))
    f_rest = f_rest.reshape(P, 3, -1)                                                                       # :506
    return dict(xyz=xyz,
                features_dc=np.ascontiguousarray(np.transpose(f_dc, (0, 2, 1))),                            # :520
                features_rest=np.ascontiguousarray(np.transpose(f_rest, (0, 2, 1))),                        # :521
                opacity=np.asarray(v["opacity"], np.float32)[:, None],
                scaling=np.stack([v[n] for n in numbered("scale_")], axis=1).astype(np.float32),
                rotation=np.stack([v[n] for n in numbered("rot")], axis=1).astype(np.float32),
                sh_degree=degree)


def unpack_capture(model_args):
    """Name the entries of a `GaussianModel.capture()` tuple (scene/gaussian_model.py:115-132)."""
    if len(model_args) != len(CAPTURE_FIELDS):
        raise ValueError(f"capture() tuple has {len(model_args)} entries, expected {len(CAPTURE_FIELDS)}")
    return dict(zip(CAPTURE_FIELDS, model_args))


def load_fuse_checkpoint(path, map_location="cpu"):
    """`chkpnt_fuse_latest.pth` (synthesize_fuse.py:108): (face model tuple, face motion-net state,
    mouth model tuple, mouth motion-net state).  Returns (face dict, mouth dict) of named capture()
    entries; the motion networks are outside this library's scope."""
    import torch
    face, _, mouth, _ = torch.load(path, map_location=map_location, weights_only=False)
    return unpack_capture(face), unpack_capture(mouth)


def rasterizer_inputs(g, device="cuda", activated=True):
    """Tensors for `GaussianRasterizer(...)` from a dict of raw GaussianModel parameters (what
    read_gaussian_ply / unpack_capture return): means3D, shs = cat(features_dc, features_rest)
    (scene/gaussian_model.py get_features), and with activated=True the activations of
    gaussian_renderer/__init__.py:227-235 through the fused kernel (training_ops.activate_gaussians)."""
    import torch
    t = lambda x: (x.detach() if hasattr(x, "detach") else torch.from_numpy(np.ascontiguousarray(x))).to(device).float()
    xyz, scaling, rotation, opacity = t(g["xyz"]), t(g["scaling"]), t(g["rotation"]), t(g["opacity"]).reshape(-1, 1)
    shs = torch.cat([t(g["features_dc"]), t(g["features_rest"])], dim=1).contiguous()
    if activated:
        from .training_ops import activate_gaussians
        means3D, scales, rotations, opac = activate_gaussians(xyz, scaling, rotation, opacity)
    else:
        means3D, scales, rotations, opac = xyz, scaling, rotation, opacity
    return dict(means3D=means3D, shs=shs, scales=scales, rotations=rotations, opacities=opac,
                sh_degree=int(g.get("sh_degree", g.get("active_sh_degree", 0))))
