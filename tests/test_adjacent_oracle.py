"""CPU: the numpy oracle of the SURVEY 8(f) rows f1/f2 (oracle/adjacent.py) against golden vectors
made from the reference's own torch expressions (tests/golden/make_adjacent_golden.py)."""
import os

import numpy as np

import util as U  # noqa: F401  (sys.path)
from oracle import adjacent as A

G = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    return dict(np.load(os.path.join(G, name)))


def test_compose_plain_is_bit_exact():
    g = _load("adjacent_compose_plain.npz")
    pred, image = A.compose_frame(g["face"], g["face_alpha"], g["mouth"], g["mouth_alpha"], g["background"], 0)
    assert np.array_equal(image, g["image"])
    assert np.array_equal(pred, g["pred"])


def test_compose_with_dilation_is_bit_exact():
    g = _load("adjacent_compose_dilate.npz")
    pred, image = A.compose_frame(g["face"], g["face_alpha"], g["mouth"], g["mouth_alpha"], g["background"],
                                  int(g["dilate"]))
    assert np.array_equal(image, g["image"])
    assert np.array_equal(pred, g["pred"])


def test_compose_without_background_and_nan():
    rng = np.random.default_rng(0)
    f, m = rng.random((3, 5, 7), dtype=np.float32), rng.random((3, 5, 7), dtype=np.float32)
    fa, ma = rng.random((1, 5, 7), dtype=np.float32), rng.random((1, 5, 7), dtype=np.float32)
    f[0, 0, 0] = np.nan
    pred, image = A.compose_frame(f, fa, m, ma)
    assert pred.shape == (5, 7, 3) and pred.dtype == np.uint8 and pred[0, 0, 0] == 0
    assert np.allclose(image[1], f[1] + m[1] * (1 - fa[0]))


def test_activate_forward_and_backward():
    g = _load("adjacent_activate.npz")
    m, s, r, o = A.activate(g["xyz"], g["d_xyz"], g["scaling"], g["d_scale"], g["rotation"], g["d_rot"], g["opacity"])
    assert np.array_equal(m, g["means3D"])
    for a, b in ((s, g["scales"]), (r, g["rotations"]), (o, g["opacities"])):
        assert np.abs(a - b).max() <= 2e-6 * max(1.0, np.abs(b).max())
    gx, gs, gr, go = A.activate_backward(g["scaling"], g["d_scale"], g["rotation"], g["d_rot"], g["opacity"],
                                         g["up_means3D"], g["up_scales"], g["up_rotations"], g["up_opacities"])
    for a, b in ((gx, g["g_xyz"]), (gx, g["g_d_xyz"]), (gs, g["g_scaling"]), (gs, g["g_d_scale"]),
                 (gr, g["g_rotation"]), (gr, g["g_d_rot"]), (go, g["g_opacity"])):
        assert np.abs(a - b).max() <= 1e-5 * max(1.0, np.abs(b).max())


def test_densify_stats():
    g = _load("adjacent_densify.npz")
    mr, acc, den = A.densify_stats(g["radii"], g["grad"], g["max_radii2D"], g["xyz_gradient_accum"], g["denom"])
    assert np.array_equal(mr, g["out_max_radii2D"])
    assert np.abs(acc - g["out_xyz_gradient_accum"]).max() <= 1e-6
    assert np.array_equal(den, g["out_denom"])
    inv = g["radii"] <= 0
    assert np.array_equal(acc[inv], g["xyz_gradient_accum"][inv]) and np.array_equal(mr[inv], g["max_radii2D"][inv])


def test_l1_ssim_loss_matches_the_references_loss_utils():
    g = _load("adjacent_loss.npz")  # made by importing /root/reference/utils/loss_utils.py
    loss, l1, ss, grad = A.l1_ssim_loss(g["image"], g["gt"], float(g["lam"]))
    assert abs(loss - float(g["loss"])) <= 1e-6 and abs(l1 - float(g["l1"])) <= 1e-6 and abs(ss - float(g["ssim"])) <= 1e-6
    assert np.abs(3 * grad - g["grad_x3"]).max() <= 1e-5 * np.abs(g["grad_x3"]).max()


def test_geometry_losses_match_train_face_expressions():
    g = _load("adjacent_geometry.npz")  # train_face.py:469,491,540 with the reference's own normalize()
    (ln, ld, la), (gd, gn, ga) = A.geometry_losses(g["depth"], g["normal"], g["alpha"], g["depth_mono"], g["normal_gt"],
                                                   g["head_mask"], g["face_mask"], g["mouth_mask"])
    assert abs(ln - float(g["l_normal"])) <= 1e-7 and abs(ld - float(g["l_depth"])) <= 1e-7 and abs(la - float(g["l_alpha"])) <= 1e-8
    for a, b in ((2 * gd, g["g_depth_x2"]), (2 * gn, g["g_normal_x2"]), (2 * ga, g["g_alpha_x2"])):
        assert np.abs(a - b).max() <= 2e-5 * np.abs(b).max()
