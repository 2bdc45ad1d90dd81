"""CPU restatement (numpy) of the steps either side of the rasterizer call that SURVEY.md 8(f)
ranks next.  TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg, never by the product path.

Pinned against the reference's own expressions: tests/golden/make_adjacent_golden.py evaluates
the literal torch lines of the reference on CPU (synthesize_fuse.py:19-22,64-76;
gaussian_renderer/__init__.py:227-235 with scene/gaussian_model.py:43,48,51;
train_face.py:675-682 + scene/gaussian_model.py:683-685) and commits inputs + outputs as
tests/golden/adjacent_*.npz; tests/test_adjacent_oracle.py holds this file to them.
"""
import numpy as np

F32 = np.float32


def dilate(alpha, ksize=13):
    """synthesize_fuse.py:19-22: F.max_pool2d(alpha, ksize, stride=1, padding=(ksize-1)//2).
    alpha [1,H,W] or [H,W]; padding never wins (max_pool2d pads with -inf)."""
    a = np.asarray(alpha, F32)
    sq = a.ndim == 3
    a2 = a[0] if sq else a
    pad = (ksize - 1) // 2
    H, W = a2.shape
    p = np.full((H + 2 * pad, W + 2 * pad), -np.inf, F32)
    p[pad:pad + H, pad:pad + W] = a2
    out = np.full((H, W), -np.inf, F32)
    for dy in range(ksize):
        for dx in range(ksize):
            np.maximum(out, p[dy:dy + H, dx:dx + W], out=out)
    return out[None] if sq else out


def compose_frame(face_rgb, face_alpha, mouth_rgb, mouth_alpha, background=None, dilate_ksize=0):
    """synthesize_fuse.py:64-76.  Returns (pred uint8 [H,W,3], image float32 [3,H,W])."""
    fr, fa = np.asarray(face_rgb, F32), np.asarray(face_alpha, F32)
    mr, am = np.asarray(mouth_rgb, F32), np.asarray(mouth_alpha, F32)
    if dilate_ksize:
        am = dilate(am, dilate_ksize)                                           # :66-68
    bg = np.zeros_like(fr) if background is None else (np.asarray(background, F32) / F32(255.0))
    mouth_image = (mr + (bg * (F32(1.0) - am)).astype(F32)).astype(F32)         # :70
    image = (fr + (mouth_image * (F32(1.0) - fa)).astype(F32)).astype(F32)      # :74
    with np.errstate(invalid="ignore"):
        clamped = np.where(np.isnan(image), F32(0), np.clip(image, F32(0), F32(1))).astype(F32)
        pred = (np.transpose(clamped, (1, 2, 0)) * F32(255)).astype(np.uint8)   # :76
    return pred, image


def softplus(x):
    """torch.nn.functional.softplus, beta=1, threshold=20 (scene/gaussian_model.py:43)."""
    x = np.asarray(x, F32)
    with np.errstate(over="ignore"):
        return np.where(x > 20, x, np.log1p(np.exp(x.astype(np.float64)))).astype(F32)


def sigmoid(x):
    x = np.asarray(x, np.float64)
    return (1.0 / (1.0 + np.exp(-x))).astype(F32)


def activate(xyz, d_xyz, scaling, d_scale, rotation, d_rot, opacity):
    """gaussian_renderer/__init__.py:227-235.  d_* may be None."""
    z = lambda d, like: np.zeros_like(like) if d is None else np.asarray(d, F32)
    means3D = (np.asarray(xyz, F32) + z(d_xyz, xyz)).astype(F32)
    scales = softplus((np.asarray(scaling, F32) + z(d_scale, scaling)).astype(F32))
    q = (np.asarray(rotation, F32) + z(d_rot, rotation)).astype(np.float64)
    n = np.maximum(np.sqrt((q * q).sum(-1, keepdims=True)), 1e-12)
    return means3D, scales, (q / n).astype(F32), sigmoid(opacity)


def activate_backward(scaling, d_scale, rotation, d_rot, opacity, g_means3D, g_scales, g_rotations, g_opacities):
    """Autograd of activate(): gradients w.r.t. xyz, _scaling, _rotation, _opacity (the d_* inputs
    receive the same gradients as their parameters)."""
    z = lambda d, like: np.zeros_like(like) if d is None else np.asarray(d, F32)
    x = (np.asarray(scaling, F32) + z(d_scale, scaling)).astype(np.float64)
    g_scaling = np.where(x > 20, g_scales, g_scales / (1.0 + np.exp(-x)))
    q = (np.asarray(rotation, F32) + z(d_rot, rotation)).astype(np.float64)
    n = np.sqrt((q * q).sum(-1, keepdims=True))
    h = q / np.maximum(n, 1e-12)
    g = np.asarray(g_rotations, np.float64)
    g_rot = np.where(n > 1e-12, (g - h * (h * g).sum(-1, keepdims=True)) / np.maximum(n, 1e-12), g * 1e12)
    s = 1.0 / (1.0 + np.exp(-np.asarray(opacity, np.float64)))
    g_op = np.asarray(g_opacities, np.float64) * s * (1 - s)
    return (np.asarray(g_means3D, F32), g_scaling.astype(F32), g_rot.astype(F32), g_op.astype(F32))


def densify_stats(radii, grad_means2D, max_radii2D, xyz_gradient_accum, denom):
    """train_face.py:675-682 + scene/gaussian_model.py:683-685; returns new copies."""
    vis = np.asarray(radii) > 0
    mr = np.array(max_radii2D, F32, copy=True)
    acc = np.array(xyz_gradient_accum, F32, copy=True)
    den = np.array(denom, F32, copy=True)
    mr[vis] = np.maximum(mr[vis], np.asarray(radii)[vis].astype(F32))
    g = np.asarray(grad_means2D, F32)[vis, :2]
    nrm = np.sqrt((g.astype(np.float64) ** 2).sum(-1)).astype(F32)
    acc.reshape(-1)[vis] += nrm
    den.reshape(-1)[vis] += F32(1)
    return mr, acc, den


def _ssim_window(window_size=11, sigma=1.5):
    """utils/loss_utils.py:42-50: fp32 1-D Gaussian, normalised, outer product."""
    g = np.array([np.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)], F32)
    g = (g / g.sum(dtype=F32)).astype(F32)
    return np.outer(g, g).astype(F32)


def _conv_same(img, win):
    """F.conv2d(img, window, padding=window_size//2, groups=channel) for [C,H,W], float64 accumulation."""
    C, H, W = img.shape
    k = win.shape[0]
    pad = k // 2
    p = np.zeros((C, H + 2 * pad, W + 2 * pad), np.float64)
    p[:, pad:pad + H, pad:pad + W] = img
    out = np.zeros((C, H, W), np.float64)
    for dy in range(k):
        for dx in range(k):
            out += float(win[dy, dx]) * p[:, dy:dy + H, dx:dx + W]
    return out


def l1_ssim_loss(image, gt, lambda_dssim=0.2):
    """train_face.py:421-422 with utils/loss_utils.py:28-29,52-73.  Returns (loss, Ll1, ssim, dloss/dimage)
    (float64 arithmetic; the gradient is the analytic adjoint of the same expression)."""
    x, y = np.asarray(image, np.float64), np.asarray(gt, np.float64)
    win = _ssim_window()
    mu1, mu2 = _conv_same(x, win), _conv_same(y, win)
    s1 = _conv_same(x * x, win) - mu1 * mu1
    s2 = _conv_same(y * y, win) - mu2 * mu2
    s12 = _conv_same(x * y, win) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    a1, a2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    b1, b2 = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    f = a1 * a2 / (b1 * b2)
    n = x.size
    ssim, l1 = f.mean(), np.abs(x - y).mean()
    loss = l1 + lambda_dssim * (1.0 - ssim)
    df_dmu1 = 2 * mu2 * a2 / (b1 * b2) - f * 2 * mu1 / b1
    df_ds1 = -f / b2
    df_ds12 = 2 * a1 / (b1 * b2)
    A = df_dmu1 - 2 * mu1 * df_ds1 - mu2 * df_ds12
    dssim = _conv_same(A, win) + 2 * x * _conv_same(df_ds1, win) + y * _conv_same(df_ds12, win)
    grad = (np.sign(x - y) - lambda_dssim * dssim) / n
    return float(loss), float(l1), float(ssim), grad.astype(F32)


def geometry_losses(depth, normal, alpha, depth_mono, normal_gt, head_mask, face_mask, mouth_mask):
    """The geometry regularisers of train_face.py on the rasterizer's depth / normal / alpha images:
        :469  0.01 * (1 - normal_gt * normal).sum(0)[head ^ mouth].mean()
        :491  1e-2 * (normalize(depth)[face ^ mouth] - normalize(depth_mono)[face ^ mouth]).abs().mean()
              normalize (utils/loss_utils.py:17-20): per ROW of the [H,W] map, (x - mean_row) /
              (std_row + 1e-2 * std(all)), unbiased standard deviations
        :540  1e-3 * (((1 - alpha) * head).mean() + (alpha * ~head).mean())
    Returns ((l_normal, l_depth, l_alpha), (dL/ddepth [1,H,W], dL/dnormal [3,H,W], dL/dalpha [1,H,W])) of their
    sum, float64 arithmetic, analytic adjoints."""
    d, m = np.asarray(depth, np.float64)[0], np.asarray(depth_mono, np.float64)
    n, ng, a = np.asarray(normal, np.float64), np.asarray(normal_gt, np.float64), np.asarray(alpha, np.float64)[0]
    head, face, mouth = (np.asarray(x).astype(bool) for x in (head_mask, face_mask, mouth_mask))
    H, W = d.shape
    m1, m2 = head ^ mouth, face ^ mouth
    l_normal = 0.01 * (3.0 - (ng * n).sum(0))[m1].mean()
    g_normal = -0.01 * ng * m1[None] / m1.sum()

    def stats(x):
        mu = x.mean(1, keepdims=True)
        sd = x.std(1, ddof=1, keepdims=True)
        return mu, sd, x.mean(), x.std(ddof=1)
    mu, sd, mu_g, sd_g = stats(d)
    s = sd + 1e-2 * sd_g
    z = (d - mu) / s
    mu_m, sd_m, _, sd_gm = stats(m)
    zm = (m - mu_m) / (sd_m + 1e-2 * sd_gm)
    diff = z - zm
    l_depth = 1e-2 * np.abs(diff[m2]).mean()
    u = 1e-2 * np.sign(diff) * m2 / m2.sum()                         # dL/dz
    A = u.sum(1, keepdims=True)
    B = (u * (d - mu)).sum(1, keepdims=True) / (s * s)               # = sum_i u_i (x_i - mu_r) / s_r^2
    g_depth = u / s - A / (W * s) - B * (d - mu) / ((W - 1) * sd) - 1e-2 * B.sum() * (d - mu_g) / ((H * W - 1) * sd_g)
    l_alpha = 1e-3 * (((1 - a) * head).mean() + (a * ~head).mean())
    g_alpha = 1e-3 * np.where(head, -1.0, 1.0) / (H * W)
    return (float(l_normal), float(l_depth), float(l_alpha)), \
        (g_depth[None].astype(F32), g_normal.astype(F32), g_alpha[None].astype(F32))
