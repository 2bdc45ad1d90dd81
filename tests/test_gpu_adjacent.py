"""GPU: the SURVEY 8(f) f1/f2 kernels (through the C ABI / their host wrappers) against the numpy
oracle and the golden vectors made from the reference's torch expressions.  uint8 frames, the
un-clamped image, max_radii2D and denom bit-exact; transcendental outputs within 2e-6 relative."""
import os

import numpy as np
import pytest

import util as U  # noqa: F401
from oracle import adjacent as A

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def _t(x, dtype=None):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    return t if dtype is None else t.to(dtype)


@pytest.mark.parametrize("name", ["adjacent_compose_plain.npz", "adjacent_compose_dilate.npz"])
def test_compose_matches_golden_bit_exact(name):
    import torch
    from instag_b200 import synthesis as SY
    g = dict(np.load(os.path.join(G, name)))
    img = torch.empty(g["image"].shape, dtype=torch.float32, device="cuda")
    out = SY.compose_frame(_t(g["face"]), _t(g["face_alpha"]), _t(g["mouth"]), _t(g["mouth_alpha"]),
                           _t(g["background"]), dilate=int(g["dilate"]) > 0, ksize=13, out_image=img)
    assert np.array_equal(img.cpu().numpy(), g["image"])
    assert np.array_equal(out.cpu().numpy(), g["pred"])


@pytest.mark.parametrize("H,W,k", [(512, 512, 13), (75, 50, 0), (33, 130, 5), (1, 1, 13)])
def test_compose_matches_oracle_on_ragged_sizes(H, W, k):
    from instag_b200 import synthesis as SY
    rng = np.random.default_rng(H * 1000 + W)
    f = (rng.random((3, H, W), dtype=np.float32) * 1.4 - 0.2)
    m = rng.random((3, H, W), dtype=np.float32)
    fa, ma = rng.random((1, H, W), dtype=np.float32), rng.random((1, H, W), dtype=np.float32)
    bg = rng.integers(0, 256, (3, H, W)).astype(np.float32)
    want, _ = A.compose_frame(f, fa, m, ma, bg, k)
    got = SY.compose_frame(_t(f), _t(fa), _t(m), _t(ma), _t(bg), dilate=k > 0, ksize=k or 13)
    assert np.array_equal(got.cpu().numpy(), want)
    want0, _ = A.compose_frame(f, fa, m, ma, None, 0)
    assert np.array_equal(SY.compose_frame(_t(f), _t(fa), _t(m), _t(ma)).cpu().numpy(), want0)


def test_frame_ring_and_sharded_sequence_return_frames_in_order():
    import torch
    from instag_b200 import synthesis as SY
    H, W, n = 40, 56, 21
    rng = np.random.default_rng(5)
    data = [dict(f=rng.random((3, H, W), dtype=np.float32), fa=rng.random((1, H, W), dtype=np.float32),
                 m=rng.random((3, H, W), dtype=np.float32), ma=rng.random((1, H, W), dtype=np.float32))
            for _ in range(n)]

    class View:
        def __init__(self, i):
            self.i = i
            self.background = torch.full((3, H, W), float(i * 10 % 256))

    views = [View(i) for i in range(n)]
    face = lambda v: {"render": _t(data[v.i]["f"]), "alpha": _t(data[v.i]["fa"])}
    mouth = lambda v: {"render": _t(data[v.i]["m"]), "alpha": _t(data[v.i]["ma"])}
    for world in (1, 2):
        seen = {}
        for rank in range(world):
            idx, frames = SY.synthesize_sequence(views, face, mouth, dilate=True, rank=rank, world=world, ring_depth=3)
            assert idx == list(range(rank, n, world)) and frames.shape == (len(idx), H, W, 3)
            for i, fr in zip(idx, frames):
                seen[i] = fr
        assert sorted(seen) == list(range(n))
        for i in range(n):
            d = data[i]
            want, _ = A.compose_frame(d["f"], d["fa"], d["m"], d["ma"], np.full((3, H, W), float(i * 10 % 256), np.float32), 13)
            assert np.array_equal(seen[i], want), i


def test_activate_matches_golden_forward_and_backward():
    import torch
    from instag_b200 import training_ops as TO
    g = dict(np.load(os.path.join(G, "adjacent_activate.npz")))
    leaves = {k: _t(g[k]).requires_grad_(True) for k in ("xyz", "d_xyz", "scaling", "d_scale", "rotation", "d_rot", "opacity")}
    m, s, r, o = TO.activate_gaussians(leaves["xyz"], leaves["scaling"], leaves["rotation"], leaves["opacity"],
                                       d_xyz=leaves["d_xyz"], d_scale=leaves["d_scale"], d_rot=leaves["d_rot"])
    assert np.array_equal(m.detach().cpu().numpy(), g["means3D"])
    for a, b in ((s, "scales"), (r, "rotations"), (o, "opacities")):
        assert np.abs(a.detach().cpu().numpy() - g[b]).max() <= 2e-6 * max(1.0, np.abs(g[b]).max()), b
    loss = (m * _t(g["up_means3D"])).sum() + (s * _t(g["up_scales"])).sum() + (r * _t(g["up_rotations"])).sum() + \
        (o * _t(g["up_opacities"])).sum()
    loss.backward()
    for k in ("xyz", "d_xyz", "scaling", "d_scale", "rotation", "d_rot", "opacity"):
        a, b = leaves[k].grad.cpu().numpy(), g["g_" + k]
        assert np.abs(a - b).max() <= 1e-5 * max(1.0, np.abs(b).max()), k
    # without displacements (render(), gaussian_renderer/__init__.py:37-133)
    m2, s2, r2, o2 = TO.activate_gaussians(_t(g["xyz"]), _t(g["scaling"]), _t(g["rotation"]), _t(g["opacity"]))
    wm, ws, wr, wo = A.activate(g["xyz"], None, g["scaling"], None, g["rotation"], None, g["opacity"])
    assert np.array_equal(m2.cpu().numpy(), wm) and np.abs(s2.cpu().numpy() - ws).max() <= 2e-6 * np.abs(ws).max()
    assert np.abs(r2.cpu().numpy() - wr).max() <= 2e-6 and np.abs(o2.cpu().numpy() - wo).max() <= 2e-6


def test_densify_stats_matches_golden():
    from instag_b200 import training_ops as TO
    g = dict(np.load(os.path.join(G, "adjacent_densify.npz")))
    mr, acc, den = _t(g["max_radii2D"]), _t(g["xyz_gradient_accum"]), _t(g["denom"])
    TO.add_densification_stats(_t(g["radii"]), _t(g["grad"]), mr, acc, den)
    assert np.array_equal(mr.cpu().numpy(), g["out_max_radii2D"])
    assert np.abs(acc.cpu().numpy() - g["out_xyz_gradient_accum"]).max() <= 1e-6
    assert np.array_equal(den.cpu().numpy(), g["out_denom"])


def test_adjacent_entry_points_reject_cpu_tensors():
    import torch
    from instag_b200 import synthesis as SY, training_ops as TO
    with pytest.raises(RuntimeError, match="CUDA-only"):
        SY.compose_frame(torch.zeros(3, 4, 4), torch.zeros(1, 4, 4), torch.zeros(3, 4, 4), torch.zeros(1, 4, 4))
    with pytest.raises(RuntimeError, match="CUDA-only"):
        TO.activate_gaussians(torch.zeros(2, 3), torch.zeros(2, 3), torch.zeros(2, 4), torch.zeros(2, 1))
