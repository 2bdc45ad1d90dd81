"""The two host sides of the autograd node -- csrc/igs_torch.cpp (default) and rasterizer.py
(IGS_HOST=python) -- run the same kernels: outputs must be bit-identical and gradients equal up to
the summation order of the backward's atomics (which differs run to run), and the compiled one must cover the same edge cases (empty cloud, capacity regrowth, heads without
gradient, the gradient-allocator hook, errors)."""
import numpy as np
import pytest
import torch

import util as U
from instag_b200 import rasterizer as R
from instag_b200 import synthetic as S

pytestmark = pytest.mark.gpu

ATOMIC_ORDER_RTOL = 1e-5  # same kernels, same inputs: only the order of the float atomics differs


def _same_grad(a, b):
    return a.shape == b.shape and U.rel_err(a.cpu().numpy(), b.cpu().numpy()) <= ATOMIC_ORDER_RTOL


def _run(sc, host, monkeypatch, heads=(0, 1, 2, 3, 5)):
    from diff_gauss import GaussianRasterizer
    monkeypatch.setenv("IGS_HOST", host)
    t = U.to_torch(sc)
    rs = U.settings_from(t)
    leaves = {}
    for k in ("means3D", "opacities", "scales", "rotations", "shs", "colors_precomp", "extra_attrs"):
        v = t.get(k)
        leaves[k] = v.clone().requires_grad_(True) if v is not None else None
    leaves["means2D"] = torch.zeros_like(t["means3D"], requires_grad=True)
    out = GaussianRasterizer(rs)(
        means3D=leaves["means3D"], means2D=leaves["means2D"], shs=leaves["shs"],
        colors_precomp=leaves["colors_precomp"], opacities=leaves["opacities"], scales=leaves["scales"],
        rotations=leaves["rotations"], extra_attrs=leaves["extra_attrs"])
    H, W = out[0].shape[1:]
    g = S.upstream_grads(H, W, extra_dims=out[5].shape[0] if out[5].numel() else 0, seed=5)
    names = ("color", "depth", "norm", "alpha", None, "extra")
    loss = None
    for i in heads:
        if out[i].numel():
            term = (out[i] * torch.from_numpy(g[names[i]]).cuda()).sum()
            loss = term if loss is None else loss + term
    if loss is not None and loss.requires_grad:
        loss.backward()
    return out, leaves


@pytest.mark.parametrize("kind", ["precomp_extra", "sh"])
def test_compiled_and_python_host_sides_agree(kind, monkeypatch):
    sc = (U.make_scene(P=4000, W=112, H=80, scale_mult=2.0, seed=31) if kind == "precomp_extra"
          else U.make_scene(P=2500, W=96, H=64, sh_degree=2, M=9, scale_mult=2.0, seed=32))
    a, la = _run(sc, "cxx", monkeypatch)
    b, lb = _run(sc, "python", monkeypatch)
    assert a[4].dtype == torch.int32 and not a[4].requires_grad
    for x, y in zip(a, b):
        assert x.shape == y.shape and torch.equal(x, y)
    for k in la:
        if la[k] is None:
            assert lb[k] is None
            continue
        assert la[k].grad is not None, k
        assert la[k].grad.shape == la[k].shape and _same_grad(la[k].grad, lb[k].grad), k


def test_compiled_host_side_partial_heads_and_no_grad(monkeypatch):
    sc = U.make_scene(P=1500, W=64, H=64, seed=33)
    a, la = _run(sc, "cxx", monkeypatch, heads=(0,))    # the other heads' gradients arrive undefined
    b, lb = _run(sc, "python", monkeypatch, heads=(0,))
    for k in ("means3D", "opacities", "scales", "rotations", "colors_precomp"):
        assert _same_grad(la[k].grad, lb[k].grad), k
    monkeypatch.setenv("IGS_HOST", "cxx")
    from diff_gauss import GaussianRasterizer
    t = U.to_torch(sc)
    with torch.no_grad():
        out = GaussianRasterizer(U.settings_from(t))(
            means3D=t["means3D"], means2D=torch.zeros_like(t["means3D"]), colors_precomp=t["colors_precomp"],
            opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"], extra_attrs=t.get("extra_attrs"))
    assert not out[0].requires_grad and torch.equal(out[0], a[0])


def test_compiled_host_side_empty_cloud(monkeypatch):
    sc = U.make_scene(P=0, W=40, H=24, seed=1)
    out, lv = _run(sc, "cxx", monkeypatch)
    # reference rasterize_points.cu:93: with P == 0 nothing runs and the outputs keep their zero fill
    assert out[4].shape == (0,) and (out[0] == 0).all() and (out[3] == 0).all()
    assert lv["means3D"].grad is None or lv["means3D"].grad.shape == (0, 3)


def test_compiled_host_side_regrows_capacity(monkeypatch):
    # first call of a (P, H, W) guesses max(8P, 65536) instances; this scene needs more
    sc = U.make_scene(P=1201, W=256, H=256, scale_mult=40.0, seed=34)
    ref = U.run_igs(sc)
    assert ref["num_rendered"] > 65536, ref["num_rendered"]
    out, _ = _run(sc, "cxx", monkeypatch)
    assert np.array_equal(out[0].detach().cpu().numpy(), ref["color"])
    assert np.array_equal(out[4].cpu().numpy(), ref["radii"])


def test_gradient_allocator_hook_reaches_the_compiled_host_side(monkeypatch):
    monkeypatch.setenv("IGS_HOST", "cxx")
    sc = U.make_scene(P=1000, W=48, H=48, seed=35)
    starts, widths, total, param_end = R.gradient_layout(1000, 0, 1)
    block = torch.empty(total, dtype=torch.float32, device="cuda")
    calls = []

    def alloc(n, device):
        calls.append((int(n), torch.device(device)))
        return block

    prev = R.set_gradient_allocator(alloc)
    try:
        _, lv = _run(sc, "cxx", monkeypatch)
    finally:
        R.set_gradient_allocator(prev)
    assert calls == [(total, block.device)]
    g = lv["means3D"].grad
    assert block.data_ptr() <= g.data_ptr() < block.data_ptr() + 4 * total
    assert torch.equal(block[starts[3]:starts[3] + 3000].view(1000, 3), g)
    _, lv2 = _run(sc, "cxx", monkeypatch)          # hook removed: an ordinary allocation again
    assert not (block.data_ptr() <= lv2["means3D"].grad.data_ptr() < block.data_ptr() + 4 * total)
    assert _same_grad(lv2["means3D"].grad, g)


def test_compiled_host_side_errors(monkeypatch):
    monkeypatch.setenv("IGS_HOST", "cxx")
    from diff_gauss import GaussianRasterizer
    sc = U.make_scene(P=10, W=16, H=16, seed=3)
    t = U.to_torch(sc)
    r = GaussianRasterizer(U.settings_from(t))
    with pytest.raises(RuntimeError, match="CUDA-only"):
        r(means3D=t["means3D"].cpu(), means2D=torch.zeros(10, 3), opacities=t["opacities"],
          colors_precomp=t["colors_precomp"], scales=t["scales"], rotations=t["rotations"])
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        r(means3D=t["means3D"][:, :2], means2D=torch.zeros(10, 3, device="cuda"), opacities=t["opacities"],
          colors_precomp=t["colors_precomp"], scales=t["scales"], rotations=t["rotations"])
