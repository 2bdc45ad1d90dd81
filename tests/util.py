"""Shared helpers of the test-suite: scene construction, running the three
implementations (product CUDA library, compiled reference, CPU oracle) and
picking their intermediates apart."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from instag_b200 import synthetic as S  # noqa: E402


def make_scene(P=2000, W=64, H=64, sh_degree=None, M=None, seed=0, ED=1, scale_mult=1.0, bg=(0.0, 1.0, 0.0),
               cam_kwargs=None, extent=0.1):
    """numpy scene dict: cloud + camera + bg (+ image size)."""
    cloud = S.make_cloud(P, sh_degree=sh_degree, M=M, seed=seed, extra_dims=ED, scale_mult=scale_mult, extent=extent)
    cam = S.make_camera(W, H, **(cam_kwargs or {}))
    sc = dict(cloud)
    sc.update(cam)
    sc["bg"] = np.asarray(bg, np.float32)
    return sc


def oracle_scene(sc):
    from oracle import cpu_oracle as O
    return O.Scene(W=sc["W"], H=sc["H"], tanfovx=sc["tanfovx"], tanfovy=sc["tanfovy"], bg=sc["bg"],
                   viewmatrix=sc["viewmatrix"], projmatrix=sc["projmatrix"], campos=sc["campos"],
                   means3D=sc["means3D"], opacities=sc["opacities"], scales=sc["scales"],
                   rotations=sc["rotations"], sh_degree=sc["sh_degree"], shs=sc["shs"],
                   colors_precomp=sc["colors_precomp"], extra_attrs=sc["extra_attrs"],
                   norm3Ds_precomp=sc.get("norm3Ds_precomp"), cov3Ds_precomp=sc.get("cov3Ds_precomp"),
                   scale_modifier=sc.get("scale_modifier", 1.0))


def to_torch(sc, device="cuda"):
    import torch
    t = {}
    for k, v in sc.items():
        t[k] = torch.from_numpy(np.ascontiguousarray(v)).to(device) if isinstance(v, np.ndarray) else v
    return t


def settings_from(tsc, debug=False, cls=None):
    if cls is None:
        from diff_gauss import GaussianRasterizationSettings as cls
    return cls(image_height=tsc["H"], image_width=tsc["W"], tanfovx=tsc["tanfovx"], tanfovy=tsc["tanfovy"],
               bg=tsc["bg"], scale_modifier=tsc.get("scale_modifier", 1.0), viewmatrix=tsc["viewmatrix"].view(4, 4),
               projmatrix=tsc["projmatrix"].view(4, 4), sh_degree=tsc["sh_degree"], campos=tsc["campos"],
               prefiltered=False, debug=debug)


def raw_args_fwd(tsc, debug=False):
    """Positional args of _C.rasterize_gaussians (DGR diff_gauss/__init__.py:71-94)."""
    import torch
    e = torch.Tensor([])
    ex = tsc["extra_attrs"] if tsc.get("extra_attrs") is not None else e
    return (tsc["bg"], tsc["means3D"], tsc["colors_precomp"] if tsc["colors_precomp"] is not None else e,
            tsc["opacities"], tsc["scales"], tsc["rotations"], tsc.get("scale_modifier", 1.0),
            tsc["cov3Ds_precomp"] if tsc.get("cov3Ds_precomp") is not None else e,
            tsc["norm3Ds_precomp"] if tsc.get("norm3Ds_precomp") is not None else e, ex,
            ex.shape[1] if ex.numel() else 0, tsc["viewmatrix"], tsc["projmatrix"], tsc["tanfovx"], tsc["tanfovy"],
            tsc["H"], tsc["W"], tsc["shs"] if tsc["shs"] is not None else e, tsc["sh_degree"], tsc["campos"], False,
            debug)


def igs_intermediates(P, H, W, num_rendered, geom, binb, img):
    """Views into libigs_raster's workspaces, named like the reference's state."""
    import torch
    from instag_b200 import _native as N
    gl, il, bl = N.geom_layout(P), N.img_layout(W, H), N.bin_layout(1)
    splat = geom[gl["splat"]:gl["splat"] + P * 64].view(torch.float32).view(P, 16)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out = dict(
        conic_opacity=splat[:, 0:4], means2D=splat[:, 4:6], depths=splat[:, 6], rgb=splat[:, 8:11],
        norm3D=splat[:, 11:14],
        tiles_touched=geom[gl["tiles_touched"]:gl["tiles_touched"] + 4 * P].view(torch.int32),
        point_offsets=geom[gl["point_offsets"]:gl["point_offsets"] + 4 * P].view(torch.int32),
        clamped=geom[gl["clamped"]:gl["clamped"] + P],
        ranges=img[il["ranges"]:il["ranges"] + 8 * T].view(torch.int32).view(T, 2),
        n_contrib=img[il["n_contrib"]:il["n_contrib"] + 4 * H * W].view(torch.int32),
        point_list=binb[bl["point_list"]:bl["point_list"] + 4 * num_rendered].view(torch.int32))
    # keys start after the capacity-sized point_list: recover capacity from the buffer size
    cap_guess = None
    for cap in range(max(num_rendered, 1), max(num_rendered, 1) + 1):
        cap_guess = cap
    out["_bin"] = binb
    return out


def igs_sorted_keys(binb, capacity, num_rendered):
    import torch
    from instag_b200 import _native as N
    bl = N.bin_layout(capacity)
    return binb[bl["keys"]:bl["keys"] + 8 * num_rendered].view(torch.int64)


def bin_capacity_from_bytes(nbytes):
    """Invert igs_bin_bytes (12 B per instance, two 256 B-aligned arrays)."""
    from instag_b200 import _native as N
    lib = N.lib()
    cap = max((nbytes - 512) // 12, 1)
    while lib.igs_bin_bytes(cap) < nbytes:
        cap += 1
    while cap > 1 and lib.igs_bin_bytes(cap - 1) >= nbytes:
        cap -= 1
    return cap
