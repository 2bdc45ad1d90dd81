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
    out["_bin"] = binb
    return out


def igs_sorted_keys(binb, capacity, num_rendered):
    import torch
    from instag_b200 import _native as N
    bl = N.bin_layout(capacity)
    return binb[bl["keys"]:bl["keys"] + 8 * num_rendered].view(torch.int64)


def bin_capacity_from_bytes(nbytes):
    """Invert igs_bin_bytes (monotone in the capacity): largest capacity whose workspace fits."""
    from instag_b200 import _native as N
    lib = N.lib()
    lo, hi = 1, max(nbytes // 4, 2)
    while lo < hi:
        mid = (lo + hi + 1) // 2
        if lib.igs_bin_bytes(mid) <= nbytes:
            lo = mid
        else:
            hi = mid - 1
    return lo


# ---------------------------------------------------------------------------
# running the three implementations on one scene
# ---------------------------------------------------------------------------
GRAD_NAMES = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dnorm3D", "dL_dsh",
              "dL_dscales", "dL_drotations", "dL_dextra")


def _bwd_args(tsc, rad, tg, geom, n, binb, img, alpha, debug=False):
    import torch
    e = torch.Tensor([])
    ex = tsc["extra_attrs"] if tsc.get("extra_attrs") is not None else e
    return (tsc["bg"], tsc["means3D"], rad, tsc["colors_precomp"] if tsc["colors_precomp"] is not None else e,
            tsc["scales"], tsc["rotations"], ex, tsc.get("scale_modifier", 1.0),
            tsc["cov3Ds_precomp"] if tsc.get("cov3Ds_precomp") is not None else e,
            tsc["norm3Ds_precomp"] if tsc.get("norm3Ds_precomp") is not None else e,
            tsc["viewmatrix"], tsc["projmatrix"], tsc["tanfovx"], tsc["tanfovy"], tg["color"], tg["depth"], tg["norm"],
            tg["alpha"], tg["extra"] if ex.numel() else e, tsc["shs"] if tsc["shs"] is not None else e,
            tsc["sh_degree"], tsc["campos"], geom, n, binb, img, alpha, debug)


def run_igs(sc, grads=None, debug=False):
    """Product path through the C ABI (instag_b200.rasterizer._forward_native /
    _backward_native = the ctypes calls into libigs_raster.so).  Returns numpy."""
    import torch
    from instag_b200 import rasterizer as R
    tsc = to_torch(sc)
    P, H, W = tsc["means3D"].shape[0], sc["H"], sc["W"]
    n, color, depth, norm, alpha, radii, extra, geom, binb, img = R._forward_native(*raw_args_fwd(tsc, debug),
                                                                                   write_keys=True)
    torch.cuda.synchronize()
    out = dict(num_rendered=n, color=color, depth=depth, norm=norm, alpha=alpha, radii=radii, extra=extra)
    if P > 0:
        # the reference's point_offsets is not on this library's path: produced on request
        from instag_b200 import _native as N
        N.check(N.lib().igs_point_offsets(P, geom.data_ptr(), torch.cuda.current_stream().cuda_stream), "point_offsets")
        torch.cuda.synchronize()
        inter = igs_intermediates(P, H, W, n, geom, binb, img)
        inter.pop("_bin")
        cap = bin_capacity_from_bytes(binb.numel())
        inter["keys"] = igs_sorted_keys(binb, cap, n)
        cl = inter.pop("clamped")
        inter["clamped"] = torch.stack([(cl >> c) & 1 for c in range(3)], dim=1)
        out.update(inter)
    if grads is not None:
        tg = {k: torch.from_numpy(v).cuda() for k, v in grads.items()}
        res = R._backward_native(*_bwd_args(tsc, radii, tg, geom, n, binb, img, alpha, debug))
        torch.cuda.synchronize()
        out.update(dict(zip(GRAD_NAMES, res)))
    return {k: (v.detach().cpu().numpy() if hasattr(v, "detach") else v) for k, v in out.items()}


def run_ref(sc, grads=None):
    """The unmodified reference (oracle/_ref/libdgr_ref.so) on the same inputs."""
    import torch
    from oracle import ref_cuda as RC
    tsc = to_torch(sc)
    P, H, W = tsc["means3D"].shape[0], sc["H"], sc["W"]
    n, color, depth, norm, alpha, radii, extra, geom, binb, img = RC.forward_raw(*raw_args_fwd(tsc))
    torch.cuda.synchronize()
    out = dict(num_rendered=n, color=color, depth=depth, norm=norm, alpha=alpha, radii=radii, extra=extra)
    if P > 0:
        out.update(RC.intermediates(P, H, W, n, geom, binb, img))
    if grads is not None:
        tg = {k: torch.from_numpy(v).cuda() for k, v in grads.items()}
        res = RC.backward_raw(*_bwd_args(tsc, radii, tg, geom, n, binb, img, alpha))
        torch.cuda.synchronize()
        out.update(dict(zip(GRAD_NAMES, res)))
    return {k: (v.detach().cpu().numpy() if hasattr(v, "detach") else v) for k, v in out.items()}


def run_oracle(sc, grads=None):
    """CPU restatement (oracle/), same dictionary keys."""
    from oracle import cpu_oracle as O
    s = oracle_scene(sc)
    f = O.forward(s)
    out = dict(f)
    out["means2D"] = f["means2D"]
    out["clamped"] = f["clamped"]
    if grads is not None:
        b = O.backward(s, f, grads["color"], grads["depth"], grads["norm"], grads["alpha"], grads.get("extra"))
        out.update(b)
    return out


# ---------------------------------------------------------------------------
# comparison helpers (the tolerances of BASELINE.json, written out)
# ---------------------------------------------------------------------------
IMG_ATOL = 1e-5    # color / depth / alpha / norm / extra: max-abs
GRAD_RTOL = 1e-4   # gradients: max-abs error relative to the tensor's max-abs


def rel_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-20))


def assert_ints_equal(a, b, names, vis=None):
    for k in names:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape, f"{k}: shape {x.shape} vs {y.shape}"
        assert np.array_equal(x.astype(np.int64), y.astype(np.int64)), \
            f"{k}: {int((x.astype(np.int64) != y.astype(np.int64)).sum())} of {x.size} differ"


def assert_forward_close(a, b, atol=IMG_ATOL, mask=None, heads=("color", "depth", "norm", "alpha", "extra")):
    """mask: optional [H*W] bool of pixels to EXCLUDE (oracle borderline pixels)."""
    for k in heads:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape, f"{k}: shape {x.shape} vs {y.shape}"
        if x.size == 0:
            continue
        d = np.abs(x - y).reshape(x.shape[0], -1)
        if mask is not None:
            d = d[:, ~mask]
        assert d.size == 0 or d.max() <= atol, f"{k}: max-abs {d.max():.3e} > {atol}"


def assert_grads_close(a, b, rtol=GRAD_RTOL, names=GRAD_NAMES):
    bad = []
    for k in names:
        if k not in a or k not in b:
            continue
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape, f"{k}: shape {x.shape} vs {y.shape}"
        r = rel_err(x, y)
        if not r <= rtol:
            bad.append(f"{k}: rel {r:.3e}")
    assert not bad, "; ".join(bad)
