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


def make_shell_cloud(P, sh_degree=None, M=None, seed=0, extra_dims=1, radii=(0.075, 0.1, 0.09), thickness=0.002,
                     opaque=True):
    """A trained head is not a filled cube but a thin surface shell: points on an ellipsoid of
    half-axes `radii` (a head-sized 15 x 20 x 18 cm) displaced along the normal by N(0, thickness),
    flat splats (one scale axis 5x smaller than the other two), and, with `opaque`, the bimodal
    opacity of a converged model (most splats near 1, a tail near 0).  Front and back surface overlap
    on screen, the silhouette is seen at grazing angles: longer lists per tile than the uniform cube,
    and pixels that saturate after a few dozen entries.  Same dictionary as make_cloud."""
    rng = np.random.default_rng(seed)
    f = np.float32
    d = rng.normal(0.0, 1.0, (P, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True) + 1e-12
    means = (d * np.asarray(radii) * (1.0 + rng.normal(0.0, thickness / min(radii), (P, 1)))).astype(f)
    area = 4.0 * math.pi * ((radii[0] * radii[1]) ** 1.6 / 3 + (radii[0] * radii[2]) ** 1.6 / 3 +
                            (radii[1] * radii[2]) ** 1.6 / 3) ** (1 / 1.6)
    s0 = 1.1 * math.sqrt(area / max(P, 1))            # splats about as large as their spacing on the surface
    scales = s0 * np.exp(rng.normal(0.0, 0.4, (P, 3)))
    scales[np.arange(P), rng.integers(0, 3, P)] *= 0.2
    q = rng.normal(0.0, 1.0, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True) + 1e-12
    logit = np.where(rng.uniform(size=(P, 1)) < 0.8, rng.normal(4.0, 1.5, (P, 1)), rng.normal(-3.0, 1.5, (P, 1))) \
        if opaque else rng.normal(0.0, 2.0, (P, 1))
    opac = (1.0 / (1.0 + np.exp(-logit))).astype(f)
    cloud = dict(means3D=means, scales=scales.astype(f), rotations=q.astype(f), opacities=opac,
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


def make_opacity_extreme_cloud(P, seed=0, **kw):
    """The uniform cloud with opacities pushed to the two ends the kernels special-case: a third
    exactly 1.0 (alpha clamps at 0.99 on every pixel near the centre), a third around 1/255 (the
    blend threshold decides pixel by pixel), a third ordinary."""
    cloud = make_cloud(P, seed=seed, **kw)
    rng = np.random.default_rng(seed + 77)
    which = rng.integers(0, 3, (P, 1))
    op = cloud["opacities"]
    op = np.where(which == 0, 1.0, np.where(which == 1, rng.uniform(0.5 / 255, 3.0 / 255, (P, 1)), op))
    cloud["opacities"] = op.astype(np.float32)
    return cloud


def make_concentrated_cloud(P, seed=0, extent=0.02, **kw):
    """Every instance in a few dozen tiles: the cloud of make_cloud squeezed into a cube of half-width
    `extent` (64 px across at 512^2: 4 x 4 tiles and their border, 64 tiles at most) with the splat sizes of the
    head-sized cloud, so that lists run to many thousands of entries per tile -- the per-tile sort's
    two-level path and the render kernels' long units."""
    cloud = make_cloud(P, seed=seed, **kw)
    cloud["means3D"] = (cloud["means3D"] * (extent / kw.get("extent", 0.1))).astype(np.float32)
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
    # not BASELINE configs: other input distributions at the C2 shape (VERDICT r01 "one input distribution")
    "shell": dict(P=100_000, W=512, H=512, sh_degree=None, M=0, backward=True, bg=(0.0, 1.0, 0.0), cloud="shell"),
    "concentrated": dict(P=100_000, W=512, H=512, sh_degree=None, M=0, backward=True, bg=(0.0, 1.0, 0.0),
                         cloud="concentrated"),
    "opacity_extreme": dict(P=100_000, W=512, H=512, sh_degree=None, M=0, backward=True, bg=(0.0, 1.0, 0.0),
                            cloud="opacity_extreme"),
}


def cloud_for(cfg, seed=0):
    """The cloud a CONFIGS entry names (default: the SURVEY 8d uniform cube)."""
    kind = cfg.get("cloud", "uniform")
    kw = dict(sh_degree=cfg["sh_degree"], M=cfg["M"] or None)
    if kind == "shell":
        return make_shell_cloud(cfg["P"], seed=seed, **kw)
    if kind == "concentrated":
        return make_concentrated_cloud(cfg["P"], seed=seed, **kw)
    if kind == "opacity_extreme":
        return make_opacity_extreme_cloud(cfg["P"], seed=seed, **kw)
    return make_cloud(cfg["P"], seed=seed, **kw)


def make_config(name, seed=0, P=None):
    cfg = dict(CONFIGS[name])
    if P is not None:
        cfg["P"] = P
    cloud = make_cloud(cfg["P"], sh_degree=cfg["sh_degree"], M=cfg["M"] or None, seed=seed)
    cam = make_camera(cfg["W"], cfg["H"])
    return cfg, cloud, cam
