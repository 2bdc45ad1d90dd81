"""Synthetic Gaussian clouds and cameras of the shapes BASELINE.json names.

There are no datasets or checkpoints on the box, so the benchmark and the
tests use seeded synthetic inputs that follow the reference's own statistics
(SURVEY.md section 8d):

* cloud "head sized": means ~ U[-0.1, 0.1]^3 (scene/dataset_readers.py:353);
* post-activation scales log-normal, median 0.35*0.2/P^(1/3), sigma_log 0.4;
* rotations = normalised N(0,1)^4 (scene/gaussian_model.py:51);
* opacity = sigmoid(N(0, 2^2));
* SH: N(0, .5^2) for the DC term, N(0, .1^2) above; precomputed colours U[0,1];
* camera as scene/cameras.py:55-64 + utils/graphics_utils.py:38-96: znear .01,
  zfar 100, focal 1400*(W/512) px, centre 0.87 away from the origin,
  viewmatrix = W2C^T, projmatrix = viewmatrix @ P^T, campos = inv(view)[3,:3].

Everything is numpy (float32, C-contiguous); callers move it where they need.
"""
import math

import numpy as np

ZNEAR, ZFAR = 0.01, 100.0


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:73-93 getProjectionMatrix (untransposed)."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    bottom, left = -top, -right
    P = np.zeros((4, 4), np.float32)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W=512, H=512, yaw=0.0, pitch=0.0, shift=(0.0, 0.0, 0.0), dist=0.87, focal=None):
    """Pinhole camera looking at the origin from `dist` along +z, optionally
    jittered.  Returns a dict with the tensors GaussianRasterizationSettings wants."""
    focal = 1400.0 * (W / 512.0) if focal is None else focal
    fovx = 2 * math.atan(W / (2 * focal))
    fovy = 2 * math.atan(H / (2 * focal))
    cy, sy, cp, sp = math.cos(yaw), math.sin(yaw), math.cos(pitch), math.sin(pitch)
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    R_w2c = Rx @ Ry                       # world -> camera rotation
    t = np.array([shift[0], shift[1], dist + shift[2]])  # camera looks down +z at the origin
    w2c = np.eye(4)
    w2c[:3, :3] = R_w2c
    w2c[:3, 3] = t
    view = np.float32(w2c).T.copy()       # world_view_transform (transposed W2C)
    proj = projection_matrix(ZNEAR, ZFAR, fovx, fovy).T
    full = (view @ proj).astype(np.float32)
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return dict(W=W, H=H, tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5),
                viewmatrix=np.ascontiguousarray(view), projmatrix=np.ascontiguousarray(full),
                campos=np.ascontiguousarray(campos))


def camera_sequence(n, W=512, H=512, seed=0):
    """n jittered cameras: yaw/pitch +-0.1 rad, translation +-0.04."""
    rng = np.random.default_rng(seed + 1000)
    cams = []
    for _ in range(n):
        yaw, pitch = rng.uniform(-0.1, 0.1, 2)
        sh = rng.uniform(-0.04, 0.04, 3)
        cams.append(make_camera(W, H, yaw=float(yaw), pitch=float(pitch), shift=tuple(sh)))
    return cams


def make_cloud(P, sh_degree=None, M=None, seed=0, extra_dims=1, scale_mult=1.0, extent=0.1):
    """Seeded Gaussian cloud.  sh_degree None -> precomputed colours [P,3];
    otherwise shs [P,M,3] with M >= (sh_degree+1)^2."""
    rng = np.random.default_rng(seed)
    f = np.float32
    means = rng.uniform(-extent, extent, (P, 3)).astype(f)
    s0 = 0.35 * (2 * extent) / max(P, 1) ** (1.0 / 3.0) * scale_mult
    scales = (s0 * np.exp(rng.normal(0.0, 0.4, (P, 3)))).astype(f)
    q = rng.normal(0.0, 1.0, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True) + 1e-12
    opac = (1.0 / (1.0 + np.exp(-rng.normal(0.0, 2.0, (P, 1))))).astype(f)
    cloud = dict(means3D=means, scales=scales, rotations=q.astype(f), opacities=opac,
                 extra_attrs=np.ones((P, extra_dims), f) if extra_dims else None,
                 shs=None, colors_precomp=None, sh_degree=0)
    if sh_degree is None:
        cloud["colors_precomp"] = rng.uniform(0.0, 1.0, (P, 3)).astype(f)
    else:
        M = (sh_degree + 1) ** 2 if M is None else M
        shs = rng.normal(0.0, 0.1, (P, M, 3))
        shs[:, 0, :] = rng.normal(0.0, 0.5, (P, 3))
        cloud["shs"] = shs.astype(f)
        cloud["sh_degree"] = sh_degree
    return cloud


def upstream_grads(H, W, extra_dims=1, seed=1):
    """Fixed random dL/d(output) images so that every head receives gradient
    (loss = sum w*out, as the mix of losses in train_face.py:456-540 does)."""
    rng = np.random.default_rng(seed)
    f = np.float32
    return dict(color=rng.normal(0, 1, (3, H, W)).astype(f), depth=rng.normal(0, 1, (1, H, W)).astype(f),
                norm=rng.normal(0, 1, (3, H, W)).astype(f), alpha=rng.normal(0, 1, (1, H, W)).astype(f),
                extra=rng.normal(0, 1, (max(extra_dims, 1), H, W)).astype(f)[:extra_dims])


# BASELINE.json configs (SURVEY.md 8d "Per-config inputs")
CONFIGS = {
    "C1": dict(P=10_000, W=512, H=512, sh_degree=0, M=1, backward=False, bg=(0.0, 0.0, 0.0)),
    "C2": dict(P=100_000, W=512, H=512, sh_degree=None, M=0, backward=True, bg=(0.0, 1.0, 0.0)),
    "C3_mouth": dict(P=30_000, W=512, H=512, sh_degree=1, M=4, backward=True, bg=(0.0, 1.0, 0.0)),
    "C3_face": dict(P=100_000, W=512, H=512, sh_degree=1, M=4, backward=True, bg=(0.0, 1.0, 0.0)),
    "C5": dict(P=1_000_000, W=1024, H=1024, sh_degree=3, M=16, backward=True, bg=(0.0, 1.0, 0.0)),
}


def make_config(name, seed=0, P=None):
    cfg = dict(CONFIGS[name])
    if P is not None:
        cfg["P"] = P
    cloud = make_cloud(cfg["P"], sh_degree=cfg["sh_degree"], M=cfg["M"] or None, seed=seed)
    cam = make_camera(cfg["W"], cfg["H"])
    return cfg, cloud, cam
