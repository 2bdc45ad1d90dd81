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
