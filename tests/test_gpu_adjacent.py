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
    for world, overlap in ((1, False), (2, False), (2, True)):
        seen = {}
        for rank in range(world):
            idx, frames = SY.synthesize_sequence(views, face, mouth, dilate=True, rank=rank, world=world, ring_depth=3,
                                                 overlap=overlap)
            assert idx == list(range(rank, n, world)) and frames.shape == (len(idx), H, W, 3)
            for i, fr in zip(idx, frames):
                seen[i] = fr
        assert sorted(seen) == list(range(n))
        for i in range(n):
            d = data[i]
            want, _ = A.compose_frame(d["f"], d["fa"], d["m"], d["ma"], np.full((3, H, W), float(i * 10 % 256), np.float32), 13)
            assert np.array_equal(seen[i], want), i


def test_two_stream_frame_equals_serial_frame():
    """PairOverlap: the mouth rasterization on a side stream under the face one (separate library
    scratch per stream) gives the frames of the serial loop, bit for bit."""
    import torch
    from diff_gauss import GaussianRasterizer
    from instag_b200 import synthesis as SY
    clouds = [U.to_torch(U.make_scene(P=6000, W=96, H=96, sh_degree=1, M=4, scale_mult=2.0, seed=s)) for s in (41, 42)]

    class View:
        def __init__(self, i):
            self.i = i

    def renderer(t):
        def fn(view):
            m = t["means3D"] + 0.002 * view.i
            out = GaussianRasterizer(U.settings_from(t))(
                means3D=m, means2D=torch.zeros_like(m), shs=t["shs"], opacities=t["opacities"], scales=t["scales"],
                rotations=t["rotations"], extra_attrs=t.get("extra_attrs"))
            return {"render": out[0], "alpha": out[3]}
        return fn

    views = [View(i) for i in range(12)]
    face, mouth = renderer(clouds[0]), renderer(clouds[1])
    i0, serial = SY.synthesize_sequence(views, face, mouth, ring_depth=4)
    i1, both = SY.synthesize_sequence(views, face, mouth, ring_depth=4, overlap=True)
    assert i0 == i1 == list(range(12)) and serial.any()
    assert np.array_equal(serial, both)


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


def test_ply_asset_renders_like_the_tensors_it_was_written_from(tmp_path):
    """f4: save_ply-format file -> read -> fused activations -> GaussianRasterizer == direct call."""
    import torch
    from diff_gauss import GaussianRasterizer
    from instag_b200 import assets as AS, synthetic as S, training_ops as TO
    rng = np.random.default_rng(3)
    P, deg = 3000, 2
    K = (deg + 1) ** 2
    raw = dict(xyz=rng.uniform(-0.1, 0.1, (P, 3)).astype(np.float32),
               features_dc=rng.normal(0, 0.5, (P, 1, 3)).astype(np.float32),
               features_rest=rng.normal(0, 0.1, (P, K - 1, 3)).astype(np.float32),
               opacity=rng.normal(0, 2, (P, 1)).astype(np.float32),
               scaling=rng.normal(-4.5, 0.4, (P, 3)).astype(np.float32),
               rotation=rng.normal(0, 1, (P, 4)).astype(np.float32))
    path = os.path.join(tmp_path, "point_cloud.ply")
    AS.write_gaussian_ply(path, **raw)
    g = AS.read_gaussian_ply(path, max_sh_degree=deg)
    ins = AS.rasterizer_inputs(g)
    cam = U.to_torch(dict(S.make_camera(96, 80), bg=np.zeros(3, np.float32)))
    cam["sh_degree"] = deg
    rs = U.settings_from(cam)
    m2 = torch.zeros_like(ins["means3D"])
    a = GaussianRasterizer(rs)(means3D=ins["means3D"], means2D=m2, shs=ins["shs"], opacities=ins["opacities"],
                               scales=ins["scales"], rotations=ins["rotations"])
    t = lambda k: _t(raw[k])
    m, s, r, o = TO.activate_gaussians(t("xyz"), t("scaling"), t("rotation"), t("opacity"))
    b = GaussianRasterizer(rs)(means3D=m, means2D=m2, shs=torch.cat([t("features_dc"), t("features_rest")], 1).contiguous(),
                               opacities=o, scales=s, rotations=r)
    assert int((a[4] > 0).sum()) > 100
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_l1_ssim_loss_matches_golden_and_oracle():
    """f3: loss within 1e-6, gradient within 1e-4 of its max (fp32 separable window vs torch conv2d)."""
    import torch
    from instag_b200 import losses as LS
    g = dict(np.load(os.path.join(G, "adjacent_loss.npz")))
    img = _t(g["image"]).requires_grad_(True)
    loss, l1, ss = LS.l1_ssim_loss(img, _t(g["gt"]), float(g["lam"]), return_terms=True)
    assert abs(float(loss.detach()) - float(g["loss"])) <= 1e-6 and abs(float(l1) - float(g["l1"])) <= 1e-6
    assert abs(float(ss) - float(g["ssim"])) <= 1e-6
    (3.0 * loss).backward()
    assert np.abs(img.grad.cpu().numpy() - g["grad_x3"]).max() <= 1e-4 * np.abs(g["grad_x3"]).max()
    # ragged size, identical images (ssim = 1, |x - y| = 0 -> sign 0), one channel
    rng = np.random.default_rng(2)
    for shape in ((3, 75, 50), (1, 16, 16), (3, 512, 512)):
        x = rng.random(shape, dtype=np.float32)
        y = np.clip(x + 0.1 * rng.standard_normal(shape).astype(np.float32), 0, 1)
        wl, w1, ws, wg = A.l1_ssim_loss(x, y, 0.2)
        tx = _t(x).requires_grad_(True)
        gl, g1, gs = LS.l1_ssim_loss(tx, _t(y), 0.2, return_terms=True)
        gl.backward()
        assert abs(float(gl.detach()) - wl) <= 2e-6 and abs(float(gs) - ws) <= 2e-6, shape
        assert np.abs(tx.grad.cpu().numpy() - wg).max() <= 1e-4 * np.abs(wg).max(), shape
    same = _t(rng.random((3, 40, 40), dtype=np.float32)).requires_grad_(True)
    l0 = LS.l1_ssim_loss(same, same.detach().clone(), 0.2)
    l0.backward()
    assert abs(float(l0.detach())) <= 1e-6 and float(same.grad.abs().max()) <= 1e-6


def test_l1_ssim_gradient_feeds_the_rasterizer_backward():
    """End of the chain f3 -> a12: loss on the rendered colour -> dL_dcolor -> per-Gaussian gradients,
    against the same chain with the reference's torch loss."""
    import sys
    import torch
    from diff_gauss import GaussianRasterizer
    from instag_b200 import losses as LS
    sc = U.make_scene(P=2500, W=80, H=64, sh_degree=None, ED=1, scale_mult=2.0, seed=31)
    t = U.to_torch(sc)
    gt = torch.rand((3, 64, 80), device="cuda", generator=torch.Generator("cuda").manual_seed(1))

    def run(fused):
        col = t["colors_precomp"].clone().requires_grad_(True)
        op = t["opacities"].clone().requires_grad_(True)
        out = GaussianRasterizer(U.settings_from(t))(
            means3D=t["means3D"], means2D=torch.zeros_like(t["means3D"]), colors_precomp=col, opacities=op,
            scales=t["scales"], rotations=t["rotations"], extra_attrs=t["extra_attrs"])
        if fused:
            loss = LS.l1_ssim_loss(out[0], gt, 0.2)
        else:
            loss = (out[0] - gt).abs().mean() + 0.2 * (1.0 - _torch_ssim(out[0], gt))
        loss.backward()
        return float(loss), col.grad, op.grad

    a, b = run(True), run(False)
    assert abs(a[0] - b[0]) <= 2e-6
    for x, y in zip(a[1:], b[1:]):
        assert float((x - y).abs().max()) <= 2e-4 * float(y.abs().max())


def _torch_ssim(img1, img2):
    """utils/loss_utils.py:42-73 restated with torch ops (the reference file itself is not on the GPU box)."""
    import torch
    import torch.nn.functional as F
    g = torch.tensor([np.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float32)
    g = (g / g.sum()).unsqueeze(1)
    C_ = img1.size(-3)
    w = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0).expand(C_, 1, 11, 11).contiguous().to(img1.device)
    mu1, mu2 = F.conv2d(img1, w, padding=5, groups=C_), F.conv2d(img2, w, padding=5, groups=C_)
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=C_) - mu1.pow(2)
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=C_) - mu2.pow(2)
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=C_) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))).mean()


def _torch_normalize(x):  # utils/loss_utils.py:17-20
    import torch
    return (x - torch.mean(x, dim=1, keepdim=True)) / (torch.std(x, dim=1, keepdim=True) + 1e-2 * torch.std(x.reshape(-1)))


def test_geometry_losses_match_golden_oracle_and_feed_the_rasterizer_backward():
    """f3, the rest: the normal / depth / alpha regularisers of train_face.py:469,491,540.  (a) the golden
    made from the reference's expressions, (b) the numpy oracle on a ragged size, (c) the whole chain
    rasterizer -> losses -> dL_ddepth / dL_dnorm / dL_dalpha -> per-Gaussian gradients against the same
    chain written with the reference's torch expressions."""
    import torch
    from diff_gauss import GaussianRasterizer
    from instag_b200 import losses as LS
    g = dict(np.load(os.path.join(G, "adjacent_geometry.npz")))
    leaves = [_t(g[k]).requires_grad_(True) for k in ("depth", "normal", "alpha")]
    total, terms = LS.geometry_losses(*leaves, _t(g["depth_mono"]), _t(g["normal_gt"]), _t(g["head_mask"]),
                                      _t(g["face_mask"]), _t(g["mouth_mask"]))
    (2.0 * total).backward()
    want = [float(g[k]) for k in ("l_normal", "l_depth", "l_alpha")]
    assert np.abs(terms.cpu().numpy() - want).max() <= 2e-7
    for leaf, key in zip(leaves, ("g_depth_x2", "g_normal_x2", "g_alpha_x2")):
        assert np.abs(leaf.grad.cpu().numpy() - g[key]).max() <= 3e-5 * np.abs(g[key]).max(), key
    # (b) ragged size, against the oracle; a switched-off term contributes nothing
    rng = np.random.default_rng(5)
    H, W = 37, 301
    d, a = rng.random((1, H, W), dtype=np.float32), rng.random((1, H, W), dtype=np.float32)
    n, ng = rng.standard_normal((3, H, W)).astype(np.float32), rng.standard_normal((3, H, W)).astype(np.float32)
    dm = rng.random((H, W), dtype=np.float32)
    hm, fm, mm = rng.random((H, W)) < 0.7, rng.random((H, W)) < 0.5, rng.random((H, W)) < 0.1
    (ln, ld, la), (gd, gn, ga) = A.geometry_losses(d, n, a, dm, ng, hm, fm, mm)
    lv = [_t(x).requires_grad_(True) for x in (d, n, a)]
    total, terms = LS.geometry_losses(*lv, _t(dm), _t(ng), _t(hm), _t(fm), _t(mm))
    total.backward()
    assert np.abs(terms.cpu().numpy() - [ln, ld, la]).max() <= 2e-7
    for leaf, w in zip(lv, (gd, gn, ga)):
        assert np.abs(leaf.grad.cpu().numpy() - w).max() <= 3e-5 * np.abs(w).max()
    _, off = LS.geometry_losses(*[x.detach() for x in lv], _t(dm), _t(ng), _t(hm), _t(fm), _t(mm), w_depth=0.0)
    assert float(off[1]) == 0.0 and abs(float(off[0]) - ln) <= 2e-7
    # (c) the chain through the rasterizer
    sc = U.make_scene(P=4000, W=96, H=80, sh_degree=None, ED=1, scale_mult=2.0, seed=33)
    t = U.to_torch(sc)
    gen = torch.Generator("cuda").manual_seed(2)
    depth_mono = torch.rand((80, 96), device="cuda", generator=gen)
    normal_gt = torch.nn.functional.normalize(torch.randn((3, 80, 96), device="cuda", generator=gen), dim=0)
    yy, xx = torch.meshgrid(torch.arange(80, device="cuda"), torch.arange(96, device="cuda"), indexing="ij")
    head_mask, face_mask = (yy - 40) ** 2 + (xx - 48) ** 2 < 38 ** 2, (yy - 40) ** 2 + (xx - 48) ** 2 < 25 ** 2
    mouth_mask = ((yy - 55) ** 2 * 4 + (xx - 48) ** 2) < 12 ** 2

    def run(fused):
        m3 = t["means3D"].clone().requires_grad_(True)
        op = t["opacities"].clone().requires_grad_(True)
        rot = t["rotations"].clone().requires_grad_(True)
        image, depth, normal, alpha, radii, extra = GaussianRasterizer(U.settings_from(t))(
            means3D=m3, means2D=torch.zeros_like(m3), colors_precomp=t["colors_precomp"], opacities=op,
            scales=t["scales"], rotations=rot, extra_attrs=t["extra_attrs"])
        if fused:
            loss, _ = LS.geometry_losses(depth, normal, alpha, depth_mono, normal_gt, head_mask, face_mask, mouth_mask)
        else:
            loss = 0.01 * (1 - normal_gt * normal).sum(0)[head_mask ^ mouth_mask].mean()                       # train_face.py:469
            loss = loss + 1e-2 * (_torch_normalize(depth[0])[face_mask ^ mouth_mask] -
                                  _torch_normalize(depth_mono)[face_mask ^ mouth_mask]).abs().mean()               # :491
            loss = loss + 1e-3 * (((1 - alpha) * head_mask).mean() + (alpha * ~head_mask).mean())                 # :540
        loss.backward()
        return float(loss), m3.grad, op.grad, rot.grad

    a_, b_ = run(True), run(False)
    assert abs(a_[0] - b_[0]) <= 2e-6
    for x, y in zip(a_[1:], b_[1:]):
        assert float((x - y).abs().max()) <= 3e-4 * float(y.abs().max())


def test_deferred_counts_give_the_same_frames_and_repair_a_frame_that_was_cut_short():
    """DeferredCounts: the no-grad calls of the sequence do not wait for num_rendered.  The splats
    grow 3x in the middle of the sequence, so the instance count of that frame exceeds the
    workspaces sized from the frames before (2x the largest count seen): the driver must notice
    when it harvests the frame and render it again.  Frames equal those of the waiting loop bit for bit."""
    import torch
    from diff_gauss import GaussianRasterizer
    from instag_b200 import rasterizer as R
    from instag_b200 import synthesis as SY
    if R.host_path() != "cxx":
        pytest.skip("deferred counts live in the compiled host side")
    clouds = [U.to_torch(U.make_scene(P=6000, W=96, H=96, sh_degree=1, M=4, scale_mult=1.0, seed=s)) for s in (51, 52)]

    class View:
        def __init__(self, i):
            self.i = i

    def renderer(t):
        def fn(view):
            m = t["means3D"] + 0.002 * view.i
            sc = t["scales"] * (3.0 if 5 <= view.i < 9 else 1.0)
            out = GaussianRasterizer(U.settings_from(t))(
                means3D=m, means2D=torch.zeros_like(m), shs=t["shs"], opacities=t["opacities"], scales=sc,
                rotations=t["rotations"], extra_attrs=t.get("extra_attrs"))
            return {"render": out[0], "alpha": out[3]}
        return fn

    views = [View(i) for i in range(14)]
    face, mouth = renderer(clouds[0]), renderer(clouds[1])
    b = R.binding()
    again = []
    real = b.deferred_overflowed

    s0 = b.deferred_mark()
    i0, waiting = SY.synthesize_sequence(views, face, mouth, ring_depth=4, deferred_counts=False)
    assert b.deferred_mark() == s0  # nothing deferred by the waiting loop

    class Spy:  # the binding with deferred_overflowed watched
        def __getattr__(self, k):
            return getattr(b, k)

        def deferred_overflowed(self, first, last):
            r = real(first, last)
            again.append(bool(r))
            return r

    dc_init = SY.DeferredCounts.__init__

    def spy_init(self):
        dc_init(self)
        self.b = Spy()
    SY.DeferredCounts.__init__ = spy_init
    try:
        for overlap in (False, True):
            again.clear()
            b.reset_capacity_hints()  # (the loops before have seen the 3x frames)
            s0 = b.deferred_mark()
            i1, deferred = SY.synthesize_sequence(views, face, mouth, ring_depth=4, overlap=overlap)
            assert i0 == i1 == list(range(14))
            assert b.deferred_mark() - s0 >= 2 * 12  # all but the first calls of each cloud (and the repairs) deferred
            assert any(again), "the 3x frame should have exceeded its workspaces"
            assert np.array_equal(waiting, deferred)
    finally:
        SY.DeferredCounts.__init__ = dc_init
    # outside the context the calls wait again
    s1 = b.deferred_mark()
    with torch.no_grad():
        face(views[0])
    assert b.deferred_mark() == s1


def test_recorded_frames_equal_the_eager_frames():
    """synthesis.RecordedFrames: a frame (face pass, mouth pass on a side stream, compose into the ring slot)
    recorded once per slot into CUDA graphs and replayed with the camera refreshed in place gives, frame for
    frame, the bytes of the eager loop."""
    import torch
    from diff_gauss import GaussianRasterizer
    from instag_b200 import rasterizer as R
    from instag_b200 import synthesis as SY
    from instag_b200 import synthetic as S
    if R.host_path() != "cxx":
        pytest.skip("the compiled host side records; IGS_HOST=python does not")
    H = W = 128
    dev = torch.device("cuda", torch.cuda.current_device())
    clouds = [U.to_torch(S.make_cloud(4000, sh_degree=1, M=4, seed=0, scale_mult=2.0), dev),
              U.to_torch(S.make_cloud(1500, sh_degree=1, M=4, seed=1, scale_mult=2.0), dev)]
    cams = S.camera_sequence(7, W, H, seed=5)
    # one block of fixed address for what changes per frame: view matrix | projection matrix | camera position
    block = torch.zeros(36, dtype=torch.float32, device=dev)
    packed = [torch.from_numpy(np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1),
                                               c["campos"].reshape(-1), np.zeros(1, np.float32)]).astype(np.float32)).to(dev)
              for c in cams]
    tsc = dict(U.to_torch(dict(cams[0], bg=np.zeros(3, np.float32)), dev))
    tsc.update(viewmatrix=block[0:16], projmatrix=block[16:32], campos=block[32:35])
    settings = [U.settings_from({**cl, **tsc}) for cl in clouds]
    m2 = [torch.zeros_like(cl["means3D"]) for cl in clouds]
    bgimg = torch.full((3, H, W), 127.0, device=dev)
    pair = SY.PairOverlap(dev)

    def render(k):
        cl = clouds[k]

        def fn(_view):
            o = GaussianRasterizer(settings[k])(means3D=cl["means3D"], means2D=m2[k], shs=cl["shs"], opacities=cl["opacities"],
                                                scales=cl["scales"], rotations=cl["rotations"], extra_attrs=cl.get("extra_attrs"))
            return {"render": o[0], "alpha": o[3]}
        return fn
    render_face, render_mouth = render(0), render(1)

    def frame(out):
        face, mouth = pair(render_face, render_mouth, None)
        SY.compose_frame(face["render"], face["alpha"], mouth["render"], mouth["alpha"], bgimg, dilate=True, out=out)

    want = []
    with torch.no_grad():
        for p in packed:
            block.copy_(p)
            out = torch.empty((H, W, 3), dtype=torch.uint8, device=dev)
            frame(out)
            want.append(out.cpu().numpy())
    ring = SY.FrameRing(H, W, depth=3, device=dev)
    block.copy_(packed[0])
    rec = SY.RecordedFrames(frame, ring)
    got = {}
    for i, p in enumerate(packed):
        if ring.full():
            got.update(dict(ring.collect(block=True)))
        block.copy_(p, non_blocking=True)
        rec.render(i)
    got.update(dict(ring.collect(block=True)))
    assert not rec.overflowed()
    assert sorted(got) == list(range(len(packed)))
    assert int((want[0] != 127).sum()) > 1000 and not np.array_equal(want[0], want[3])
    for i in range(len(packed)):
        assert np.array_equal(got[i], want[i]), i
