"""Drop-in API behaviour on the GPU: the keyword call gaussian_renderer makes,
autograd plumbing, edge cases (P=0, everything culled, capacity regrowth,
non-contiguous inputs, markVisible), error texts of the reference."""
import numpy as np
import pytest
import torch

import util as U
from instag_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def _call(sc, requires_grad=True, use_ref=False, **over):
    from diff_gauss import GaussianRasterizationSettings, GaussianRasterizer
    t = U.to_torch(sc)
    t.update(over)
    rs = U.settings_from(t)
    if use_ref:
        from oracle.ref_cuda import RefGaussianRasterizer as GaussianRasterizer  # noqa: F811
    leaves = {}
    for k in ("means3D", "opacities", "scales", "rotations", "shs", "colors_precomp", "extra_attrs"):
        v = t.get(k)
        if v is not None and requires_grad:
            v = v.clone().requires_grad_(True)
        leaves[k] = v
    means2D = torch.zeros_like(t["means3D"], requires_grad=requires_grad)
    out = GaussianRasterizer(raster_settings=rs)(
        means3D=leaves["means3D"], means2D=means2D, shs=leaves["shs"], colors_precomp=leaves["colors_precomp"],
        opacities=leaves["opacities"], scales=leaves["scales"], rotations=leaves["rotations"], cov3Ds_precomp=None,
        extra_attrs=leaves["extra_attrs"])
    leaves["means2D"] = means2D
    return out, leaves


def _loss(out, H, W, ED=1):
    g = S.upstream_grads(H, W, extra_dims=ED, seed=5)
    color, depth, norm, alpha, radii, extra = out
    l = (color * torch.from_numpy(g["color"]).cuda()).sum() + (depth * torch.from_numpy(g["depth"]).cuda()).sum()
    l = l + (norm * torch.from_numpy(g["norm"]).cuda()).sum() + (alpha * torch.from_numpy(g["alpha"]).cuda()).sum()
    if ED:
        l = l + (extra * torch.from_numpy(g["extra"]).cuda()).sum()
    return l


def test_module_call_and_autograd_match_reference():
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("reference library not built")
    sc = U.make_scene(P=3000, W=96, H=64, sh_degree=1, M=4, scale_mult=2.0, seed=21)
    out, lv = _call(sc)
    rout, rlv = _call(sc, use_ref=True)
    assert [tuple(o.shape) for o in out] == [(3, 64, 96), (1, 64, 96), (3, 64, 96), (1, 64, 96), (3000,), (1, 64, 96)]
    assert out[4].dtype == torch.int32 and torch.equal(out[4], rout[4])
    for a, b in zip(out[:4] + out[5:], rout[:4] + rout[5:]):
        assert (a - b).abs().max().item() <= 1e-5
    _loss(out, 64, 96).backward()
    _loss(rout, 64, 96).backward()
    assert lv["means2D"].grad.shape == (3000, 3) and (lv["means2D"].grad[:, 2] == 0).all()
    for k in ("means3D", "means2D", "opacities", "scales", "rotations", "shs", "extra_attrs"):
        assert U.rel_err(lv[k].grad.cpu().numpy(), rlv[k].grad.cpu().numpy()) <= 2e-4, k


def test_no_grad_and_detached_inputs():
    sc = U.make_scene(P=500, W=48, H=48, seed=22)
    with torch.no_grad():
        out, _ = _call(sc, requires_grad=False)
    assert not out[0].requires_grad
    out, lv = _call(sc)
    out[0].sum().backward()  # only the colour head receives gradient: others are None -> zeros
    assert lv["means3D"].grad is not None and torch.isfinite(lv["means3D"].grad).all()


def test_empty_cloud_and_all_culled():
    sc = U.make_scene(P=0, W=40, H=24, seed=1)
    o = U.run_igs(sc)
    # reference rasterize_points.cu:93: with P == 0 nothing runs, outputs keep their zero fill (not bg)
    assert o["num_rendered"] == 0 and (o["alpha"] == 0).all() and (o["color"] == 0).all()
    sc = U.make_scene(P=777, W=40, H=24, seed=1)
    sc["means3D"] = sc["means3D"] + np.array([0, 0, -5.0], np.float32)
    g = S.upstream_grads(24, 40)
    o = U.run_igs(sc, grads=g)
    assert o["num_rendered"] == 0 and (o["radii"] == 0).all() and (o["alpha"] == 0).all()
    assert np.allclose(o["color"], np.asarray(sc["bg"]).reshape(3, 1, 1))
    for k in U.GRAD_NAMES:
        assert (o[k] == 0).all(), k


def test_capacity_regrowth_gives_identical_result():
    from instag_b200 import rasterizer as R
    sc = U.make_scene(P=3000, W=128, H=128, scale_mult=6.0, seed=23)
    a = U.run_igs(sc)
    dev = torch.device("cuda", torch.cuda.current_device())
    R._get_scratch(dev).capacity_hint[(3000, 128, 128)] = 1000   # far too small: forces IGS_ERR_CAPACITY
    b = U.run_igs(sc)
    assert a["num_rendered"] == b["num_rendered"] > 1000
    for k in ("keys", "point_list", "ranges", "n_contrib", "color", "alpha"):
        assert np.array_equal(a[k], b[k]), k


def test_non_contiguous_and_sentinel_inputs():
    sc = U.make_scene(P=800, W=48, H=48, seed=24)
    base, _ = _call(sc, requires_grad=False)
    t = U.to_torch(sc)
    wide = torch.zeros(800, 6, device="cuda")
    wide[:, ::2] = t["means3D"]
    out, _ = _call(sc, requires_grad=False, means3D=wide[:, ::2])
    for a, b in zip(base, out):
        assert torch.equal(a, b)


def test_misaligned_views_are_accepted_like_the_reference_does():
    """A contiguous view at an odd storage offset (data_ptr % 16 != 0): the host sides copy it, the C ABI
    itself refuses it cleanly (the kernels' 16-byte accesses would fault and poison the context)."""
    from instag_b200 import _native as N
    from instag_b200 import rasterizer as R
    sc = U.make_scene(P=801, W=48, H=48, seed=24)
    base, _ = _call(sc, requires_grad=False)
    t = U.to_torch(sc)
    odd = {}
    for k in ("means3D", "scales", "rotations", "colors_precomp", "opacities"):
        flat = torch.zeros(t[k].numel() + 1, device="cuda")
        flat[1:] = t[k].reshape(-1)
        odd[k] = flat[1:].view(t[k].shape)
        assert odd[k].data_ptr() % 16 != 0 and odd[k].is_contiguous()
    out, _ = _call(sc, requires_grad=False, **odd)
    for a, b in zip(base, out):
        assert torch.equal(a, b)
    # the C ABI with a misaligned pointer: IGS_ERR_INVALID, not a fault
    R._forward_native(*U.raw_args_fwd(t))
    torch.cuda.synchronize()
    dev = t["means3D"].device
    s = R._get_scratch(dev, torch.cuda.current_stream(dev).cuda_stream)
    keep = s.fwd_args.means3D
    s.fwd_args.means3D = odd["means3D"].data_ptr()
    try:
        assert N.lib().igs_forward(s.fwd_args_ref, torch.cuda.current_stream(dev).cuda_stream, s.n_out_ref) == N.IGS_ERR_INVALID
        assert "aligned" in N.last_error()
    finally:
        s.fwd_args.means3D = keep
    torch.cuda.synchronize()


def test_mark_visible():
    from diff_gauss import GaussianRasterizer
    sc = U.make_scene(P=4000, W=32, H=32, seed=3, extent=0.8)
    t = U.to_torch(sc)
    vis = GaussianRasterizer(U.settings_from(t)).markVisible(t["means3D"])
    from oracle import cpu_oracle as O
    assert vis.dtype == torch.bool
    assert np.array_equal(vis.cpu().numpy(), O.mark_visible(sc["means3D"], sc["viewmatrix"], sc["projmatrix"]))


def test_reference_error_texts():
    from diff_gauss import GaussianRasterizer
    sc = U.make_scene(P=10, W=16, H=16, seed=3)
    t = U.to_torch(sc)
    r = GaussianRasterizer(U.settings_from(t))
    m2 = torch.zeros_like(t["means3D"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], colors_precomp=t["colors_precomp"])
    with pytest.raises(ValueError, match="scales == None is not allowed"):
        r(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], colors_precomp=t["colors_precomp"],
          cov3Ds_precomp=torch.zeros(10, 6, device="cuda"))
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r(means3D=t["means3D"].cpu(), means2D=m2, opacities=t["opacities"], colors_precomp=t["colors_precomp"],
          scales=t["scales"], rotations=t["rotations"])


def test_deferred_count_through_the_c_abi():
    """IgsFwdArgs.defer_count + igs_forward_count: the frame is queued without waiting, the count
    (and IGS_ERR_CAPACITY, exactly as igs_forward would have returned it) comes from the second
    call; images equal those of the waiting call."""
    from instag_b200 import _native as N
    from instag_b200 import rasterizer as R
    sc = U.make_scene(P=3000, W=96, H=80, sh_degree=None, ED=1, scale_mult=2.0, seed=21)
    t = U.to_torch(sc)
    out = R._forward_native(*U.raw_args_fwd(t))
    n0, color0 = int(out[0]), out[1].clone()
    torch.cuda.synchronize()
    dev = t["means3D"].device
    stream = torch.cuda.current_stream(dev).cuda_stream
    s = R._get_scratch(dev, stream)
    a, lib = s.fwd_args, N.lib()
    assert n0 > 0 and a.capacity >= n0
    out[1].zero_()
    a.defer_count = 1
    try:
        assert lib.igs_forward(s.fwd_args_ref, stream, s.n_out_ref) == N.IGS_OK and s.n_out.value == -1
        assert lib.igs_forward_count(s.fwd_args_ref, stream, s.n_out_ref) == N.IGS_OK and s.n_out.value == n0
        torch.cuda.synchronize()
        assert torch.equal(out[1], color0)
        # a workspace that is too small: nothing faults, the count says so
        a.capacity = n0 - 1
        assert lib.igs_forward(s.fwd_args_ref, stream, s.n_out_ref) == N.IGS_OK
        assert lib.igs_forward_count(s.fwd_args_ref, stream, s.n_out_ref) == N.IGS_ERR_CAPACITY and s.n_out.value == n0
        torch.cuda.synchronize()
    finally:
        a.defer_count = 0
        s.dirty.add("ctrl")


def test_run_ahead_frames_of_the_python_host_side():
    """`_forward_native(defer=True)` + `_backward_native(num_rendered=-1)` + `check_deferred`: frames queued
    without the host reading their count give the images of the waiting calls bit for bit and their
    gradients to the run-to-run spread of the float atomics;
    a frame that outgrows its workspaces is not rendered, its backward stands down without touching
    memory it does not own, and `check_deferred` says so."""
    from instag_b200 import rasterizer as R
    sc = U.make_scene(P=3000, W=96, H=80, sh_degree=None, ED=1, scale_mult=2.0, seed=22)
    t = U.to_torch(sc)
    g = S.upstream_grads(sc["H"], sc["W"], extra_dims=1, seed=3)
    tg = {k: torch.from_numpy(v).cuda() for k, v in g.items()}
    dev = t["means3D"].device
    w = R._forward_native(*U.raw_args_fwd(t))                       # waits; leaves the capacity hint
    n0 = int(w[0])
    want = R._backward_native(*U._bwd_args(t, w[5], tg, w[7], n0, w[8], w[9], w[4]))
    torch.cuda.synchronize()
    assert n0 > 0
    for _ in range(3):                                               # three frames in flight, none read
        d = R._forward_native(*U.raw_args_fwd(t), defer=True)
        assert d[0] == -1
        got = R._backward_native(*U._bwd_args(t, d[5], tg, d[7], -1, d[8], d[9], d[4]))
    assert R.check_deferred(dev) == n0
    torch.cuda.synchronize()
    for a, b in zip(w[1:7], d[1:7]):
        assert torch.equal(a, b)
    for a, b in zip(want, got):   # (float atomics: two runs of one frame differ in the last bits)
        assert a.shape == b.shape and (a.numel() == 0 or U.rel_err(b.cpu().numpy(), a.cpu().numpy()) <= 1e-5)
    # a hint that is too small for the frame
    s = R._get_scratch(dev, torch.cuda.current_stream(dev).cuda_stream)
    key = (3000, sc["H"], sc["W"])
    assert key in s.capacity_hint
    s.capacity_hint[key] = n0 // 2
    d = R._forward_native(*U.raw_args_fwd(t), defer=True)
    R._backward_native(*U._bwd_args(t, d[5], tg, d[7], -1, d[8], d[9], d[4]))
    with pytest.raises(RuntimeError, match="not rendered"):
        R.check_deferred(dev)
    torch.cuda.synchronize()                                         # nothing faulted
    assert s.capacity_hint[key] >= n0
    again = R._forward_native(*U.raw_args_fwd(t), defer=True)
    R.check_deferred(dev)
    torch.cuda.synchronize()
    assert torch.equal(again[1], w[1])
