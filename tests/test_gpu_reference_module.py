"""GPU: this repo's `diff_gauss` against the reference's OWN extension -- its Python module, its
pybind / torch binding and its kernels, built unmodified into baseline/_ref by baseline/build_ref.sh
-- through the call surface InsTaG uses (gaussian_renderer/__init__.py):

* the module API on identical tensors, outputs and autograd gradients;
* the train_face.py step shape: one SH render + one colors_precomp attention render on detached
  geometry, both sharing ONE means2D tensor, one backward (gaussian_renderer/__init__.py:241-268);
* the reference's gaussian_renderer/__init__.py itself, UNMODIFIED (copied beside the extension by the
  build recipe), with stub model / camera / motion-network objects: `render` and `render_motion` run
  once on top of each rasterizer.

Skipped where baseline/_ref has not been built (it travels to the GPU box with the repo snapshot)."""
import importlib.util
import math
import os
import sys
import types

import numpy as np
import pytest

import util as U

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ref():
    from baseline import reference as BR
    if not BR.available():
        pytest.skip("baseline/_ref not built (baseline/build_ref.sh)")
    return BR.load()


def _leaves(t, names):
    return {k: t[k].clone().requires_grad_(True) for k in names if t.get(k) is not None}


def _run_module(mod, t, weights, colors_key):
    import torch
    names = ["means3D", "opacities", "scales", "rotations", colors_key]
    lv = _leaves(t, names)
    m2 = torch.zeros_like(lv["means3D"], requires_grad=True)
    rs = U.settings_from(t, cls=mod.GaussianRasterizationSettings)
    out = mod.GaussianRasterizer(rs)(
        means3D=lv["means3D"], means2D=m2, shs=lv.get("shs"), colors_precomp=lv.get("colors_precomp"),
        opacities=lv["opacities"], scales=lv["scales"], rotations=lv["rotations"], extra_attrs=t["extra_attrs"])
    color, depth, norm, alpha, radii, extra = out
    loss = sum((o * w).sum() for o, w in zip((color, depth, norm, alpha, extra), weights))
    loss.backward()
    torch.cuda.synchronize()
    grads = {k: v.grad.detach().cpu().numpy() for k, v in lv.items()}
    grads["means2D"] = m2.grad.detach().cpu().numpy()
    return [o.detach().cpu().numpy() for o in out], grads


@pytest.mark.parametrize("kind", ["precomp", "sh1_m4", "sh3"])
def test_module_api_matches_reference_extension(kind):
    import torch
    ref = _ref()
    import diff_gauss as ours
    kw = dict(precomp=dict(sh_degree=None), sh1_m4=dict(sh_degree=1, M=4), sh3=dict(sh_degree=3, M=16))[kind]
    sc = U.make_scene(P=20000, W=160, H=128, seed=41, scale_mult=1.5, **kw)
    t = U.to_torch(sc)
    g = U.S.upstream_grads(128, 160, seed=5)
    weights = [torch.from_numpy(g[k]).cuda() for k in ("color", "depth", "norm", "alpha", "extra")]
    key = "colors_precomp" if kind == "precomp" else "shs"
    a_out, a_g = _run_module(ours, t, weights, key)
    b_out, b_g = _run_module(ref, t, weights, key)
    assert np.array_equal(a_out[4], b_out[4])  # radii
    for i, name in enumerate(("color", "depth", "norm", "alpha")):
        assert np.abs(a_out[i] - b_out[i]).max() <= U.IMG_ATOL, name
    assert np.abs(a_out[5] - b_out[5]).max() <= U.IMG_ATOL
    for k in b_g:
        assert U.rel_err(a_g[k], b_g[k]) <= U.GRAD_RTOL, (k, U.rel_err(a_g[k], b_g[k]))


def _pair(mod, t, attn, w_main, w_attn):
    """gaussian_renderer/__init__.py:241-268 with the motion networks left out."""
    import torch
    lv = _leaves(t, ["means3D", "opacities", "scales", "rotations", "shs"])
    attn_leaf = attn.clone().requires_grad_(True)
    m2 = torch.zeros_like(lv["means3D"], requires_grad=True) + 0
    m2.retain_grad()
    rast = mod.GaussianRasterizer(U.settings_from(t, cls=mod.GaussianRasterizationSettings))
    image, depth, norm, alpha, radii, extra = rast(
        means3D=lv["means3D"], means2D=m2, shs=lv["shs"], colors_precomp=None, opacities=lv["opacities"],
        scales=lv["scales"], rotations=lv["rotations"], cov3Ds_precomp=None, extra_attrs=torch.ones_like(lv["opacities"]))
    attn_img, _, _, _, _, _ = rast(
        means3D=lv["means3D"].detach(), means2D=m2, shs=None, colors_precomp=attn_leaf,
        opacities=lv["opacities"].detach(), scales=lv["scales"].detach(), rotations=lv["rotations"].detach(),
        cov3Ds_precomp=None, extra_attrs=torch.ones_like(lv["opacities"]))
    loss = sum((o * w).sum() for o, w in zip((image, depth, norm, alpha, extra), w_main)) + (attn_img * w_attn).sum()
    loss.backward()
    torch.cuda.synchronize()
    g = {k: v.grad.detach().cpu().numpy() for k, v in lv.items()}
    g["attn"] = attn_leaf.grad.detach().cpu().numpy()
    g["means2D"] = m2.grad.detach().cpu().numpy()
    return image.detach().cpu().numpy(), attn_img.detach().cpu().numpy(), g


def test_train_face_pair_shares_means2D_and_detaches_geometry():
    import torch
    ref = _ref()
    import diff_gauss as ours
    sc = U.make_scene(P=30000, W=192, H=192, sh_degree=1, M=4, seed=43, scale_mult=1.4)
    t = U.to_torch(sc)
    rng = np.random.default_rng(3)
    attn = torch.from_numpy(np.concatenate([rng.uniform(0, 1, (30000, 2)), np.zeros((30000, 1))], 1).astype(np.float32)).cuda()
    g = U.S.upstream_grads(192, 192, seed=6)
    w_main = [torch.from_numpy(g[k]).cuda() for k in ("color", "depth", "norm", "alpha", "extra")]
    w_attn = torch.from_numpy(U.S.upstream_grads(192, 192, seed=7)["color"]).cuda()
    a_img, a_attn, a_g = _pair(ours, t, attn, w_main, w_attn)
    b_img, b_attn, b_g = _pair(ref, t, attn, w_main, w_attn)
    assert np.abs(a_img - b_img).max() <= U.IMG_ATOL and np.abs(a_attn - b_attn).max() <= U.IMG_ATOL
    for k in b_g:  # means2D receives the sum of both renders' gradients; the geometry only the first's
        assert U.rel_err(a_g[k], b_g[k]) <= U.GRAD_RTOL, (k, U.rel_err(a_g[k], b_g[k]))
    assert np.abs(b_g["attn"][:, 2]).max() > 0  # (the zero third channel still receives gradient)


# ---- the reference's gaussian_renderer/__init__.py, unmodified, on top of either rasterizer ----
CALLER = os.path.join(ROOT, "baseline", "_ref", "instag_caller", "gaussian_renderer", "__init__.py")


def _load_caller(raster_module, tag):
    """Import the reference's gaussian_renderer/__init__.py with `diff_gauss` bound to `raster_module`
    and its other imports (model / motion networks / SH helper: not on this path) stubbed."""
    stubs = {"scene": types.ModuleType("scene"), "scene.gaussian_model": types.ModuleType("scene.gaussian_model"),
             "scene.motion_net": types.ModuleType("scene.motion_net"), "utils": types.ModuleType("utils"),
             "utils.sh_utils": types.ModuleType("utils.sh_utils")}
    stubs["scene.gaussian_model"].GaussianModel = object
    stubs["scene.motion_net"].MotionNetwork = object
    stubs["scene.motion_net"].MouthMotionNetwork = object
    stubs["utils.sh_utils"].eval_sh = None
    saved = {k: sys.modules.get(k) for k in list(stubs) + ["diff_gauss"]}
    try:
        sys.modules.update(stubs)
        sys.modules["diff_gauss"] = raster_module
        spec = importlib.util.spec_from_file_location("gaussian_renderer_" + tag, CALLER)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


class _Cam:
    def __init__(self, t):
        import torch
        self.FoVx, self.FoVy = 2 * math.atan(t["tanfovx"]), 2 * math.atan(t["tanfovy"])
        self.image_height, self.image_width = t["H"], t["W"]
        self.world_view_transform, self.full_proj_transform = t["viewmatrix"].view(4, 4), t["projmatrix"].view(4, 4)
        self.camera_center = t["campos"]
        self.talking_dict = {"auds": torch.zeros(1), "au_exp": torch.zeros(1)}


class _Pipe:
    debug, compute_cov3D_python, convert_SHs_python = False, False, False


class _Model:
    """The attributes of scene/gaussian_model.py::GaussianModel that render / render_motion read."""

    def __init__(self, t):
        import torch
        f = torch.nn.functional
        self._xyz = t["means3D"].clone().requires_grad_(True)
        self._scaling = torch.log(torch.expm1(t["scales"])).requires_grad_(True)  # inverse softplus
        self._rotation = t["rotations"].clone().requires_grad_(True)
        self._opacity = torch.logit(t["opacities"].clamp(1e-4, 1 - 1e-4)).requires_grad_(True)
        self._features = t["shs"].clone().requires_grad_(True)
        self.active_sh_degree = self.max_sh_degree = t["sh_degree"]
        self.scaling_activation, self.rotation_activation = f.softplus, f.normalize

    get_xyz = property(lambda s: s._xyz)
    get_opacity = property(lambda s: __import__("torch").sigmoid(s._opacity))
    get_scaling = property(lambda s: s.scaling_activation(s._scaling))
    get_rotation = property(lambda s: s.rotation_activation(s._rotation))
    get_features = property(lambda s: s._features)

    def params(self):
        return dict(xyz=self._xyz, scaling=self._scaling, rotation=self._rotation, opacity=self._opacity,
                    features=self._features)


def _motion_net(P, seed):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s, k=1.0: (torch.rand(*s, device="cuda", generator=g) - 0.5) * k
    fixed = dict(d_xyz=r(P, 3, k=2e-3), d_scale=r(P, 3, k=0.1), d_rot=r(P, 4, k=0.1))
    amb = dict(ambient_aud=torch.rand(P, 1, device="cuda", generator=g).requires_grad_(True),
               ambient_eye=torch.rand(P, 1, device="cuda", generator=g).requires_grad_(True))

    def net(xyz, aud, exp):
        return dict({k: v.clone() for k, v in fixed.items()}, **amb)
    return net, amb


def _run_caller(caller, t, weights, seed):
    import torch
    model, cam, pipe = _Model(t), _Cam(t), _Pipe()
    pkg = caller.render(cam, model, pipe, t["bg"])
    net, amb = _motion_net(t["means3D"].shape[0], seed)
    pkg2 = caller.render_motion(cam, model, net, pipe, t["bg"], return_attn=True)
    loss = (pkg["render"] * weights[0]).sum() + (pkg["depth"] * weights[1]).sum() + (pkg["normal"] * weights[2]).sum() + \
        (pkg["alpha"] * weights[3]).sum() + (pkg2["render"] * weights[0]).sum() * 0.5 + (pkg2["attn"] * weights[0]).sum()
    loss.backward()
    torch.cuda.synchronize()
    grads = {k: v.grad.detach().cpu().numpy() for k, v in model.params().items()}
    grads.update({k: v.grad.detach().cpu().numpy() for k, v in amb.items()})
    grads["viewspace_points"] = pkg["viewspace_points"].grad.detach().cpu().numpy()
    grads["viewspace_points_motion"] = pkg2["viewspace_points"].grad.detach().cpu().numpy()
    imgs = {k: pkg[k].detach().cpu().numpy() for k in ("render", "depth", "alpha", "normal", "radii")}
    imgs.update(motion=pkg2["render"].detach().cpu().numpy(), attn=pkg2["attn"].detach().cpu().numpy(),
                visibility=pkg2["visibility_filter"].cpu().numpy())
    return imgs, grads


def test_unmodified_gaussian_renderer_runs_on_the_drop_in():
    import torch
    if not os.path.isfile(CALLER):
        pytest.skip("baseline/_ref/instag_caller not present (baseline/build_ref.sh copies it)")
    ref = _ref()
    import diff_gauss as ours
    sc = U.make_scene(P=25000, W=160, H=160, sh_degree=1, M=4, seed=47, scale_mult=1.4)
    t = U.to_torch(sc)
    g = U.S.upstream_grads(160, 160, seed=8)
    weights = [torch.from_numpy(g[k]).cuda() for k in ("color", "depth", "norm", "alpha")]
    a_img, a_g = _run_caller(_load_caller(ours, "ours"), t, weights, 9)
    b_img, b_g = _run_caller(_load_caller(ref, "ref"), t, weights, 9)
    assert np.array_equal(a_img["radii"], b_img["radii"]) and np.array_equal(a_img["visibility"], b_img["visibility"])
    for k in ("render", "depth", "alpha", "normal", "motion", "attn"):
        assert np.abs(a_img[k] - b_img[k]).max() <= U.IMG_ATOL, k
    for k in b_g:
        assert U.rel_err(a_g[k], b_g[k]) <= U.GRAD_RTOL, (k, U.rel_err(a_g[k], b_g[k]))
