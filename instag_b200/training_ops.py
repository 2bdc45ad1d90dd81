"""The steps either side of the rasterizer call in training, each as one
kernel of libigs_raster behind the reference's own expressions:

* `activate_gaussians` = gaussian_renderer/__init__.py:227-235 (means3D = xyz + d_xyz,
  scales = softplus(_scaling + d_scale), rotations = normalize(_rotation + d_rot),
  opacity = sigmoid(_opacity)) as ONE autograd.Function: one kernel forward, one backward,
  instead of ~8 element-wise launches and their autograd nodes;
* `add_densification_stats` = train_face.py:675-682 + GaussianModel.add_densification_stats
  (scene/gaussian_model.py:683-685), in place, without the boolean-mask gathers/scatters.

No CPU fallback: CUDA tensors only.
"""
import torch

from . import _native as N


def _c(t):
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def _ptr(t):
    return None if t is None else t.data_ptr()


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, d_xyz, scaling, d_scale, rotation, d_rot, opacity):
        if not xyz.is_cuda:
            raise RuntimeError("libigs_raster is CUDA-only: activate_gaussians needs CUDA tensors (no CPU fallback)")
        P, dev = xyz.shape[0], xyz.device
        ins = [_c(t) for t in (xyz, d_xyz, scaling, d_scale, rotation, d_rot, opacity)]
        means3D = torch.empty((P, 3), dtype=torch.float32, device=dev)
        scales = torch.empty((P, 3), dtype=torch.float32, device=dev)
        rotations = torch.empty((P, 4), dtype=torch.float32, device=dev)
        opac = torch.empty((P, 1), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            N.check(N.lib().igs_activate_forward(P, *[_ptr(t) for t in ins], means3D.data_ptr(), scales.data_ptr(),
                                                 rotations.data_ptr(), opac.data_ptr(), stream), "activate forward")
        ctx.save_for_backward(*[t for t in (ins[2], ins[3], ins[4], ins[5], ins[6]) if t is not None])
        ctx.have = [t is not None for t in (ins[3], ins[5])]
        ctx.has_dxyz = d_xyz is not None
        return means3D, scales, rotations, opac

    @staticmethod
    def backward(ctx, g_means3D, g_scales, g_rotations, g_opac):
        saved = list(ctx.saved_tensors)
        scaling = saved.pop(0)
        d_scale = saved.pop(0) if ctx.have[0] else None
        rotation = saved.pop(0)
        d_rot = saved.pop(0) if ctx.have[1] else None
        opacity = saved.pop(0)
        P, dev = scaling.shape[0], scaling.device
        ups = [_c(g) for g in (g_means3D, g_scales, g_rotations, g_opac)]
        g_xyz = torch.empty((P, 3), dtype=torch.float32, device=dev)
        g_scaling = torch.empty((P, 3), dtype=torch.float32, device=dev)
        g_rotation = torch.empty((P, 4), dtype=torch.float32, device=dev)
        g_opacity = torch.empty((P, 1), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            N.check(N.lib().igs_activate_backward(P, _ptr(scaling), _ptr(d_scale), _ptr(rotation), _ptr(d_rot),
                                                  _ptr(opacity), *[_ptr(u) for u in ups], g_xyz.data_ptr(),
                                                  g_scaling.data_ptr(), g_rotation.data_ptr(), g_opacity.data_ptr(),
                                                  stream), "activate backward")
        return (g_xyz, g_xyz if ctx.has_dxyz else None, g_scaling, g_scaling if ctx.have[0] else None,
                g_rotation, g_rotation if ctx.have[1] else None, g_opacity)


def activate_gaussians(xyz, scaling, rotation, opacity, d_xyz=None, d_scale=None, d_rot=None):
    """(means3D, scales, rotations, opacity) of gaussian_renderer/__init__.py:227-235; `opacity`
    is the raw `_opacity` [P,1].  Differentiable w.r.t. every input."""
    return _Activate.apply(xyz, d_xyz, scaling, d_scale, rotation, d_rot, opacity)


def add_densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """In place, for the Gaussians with radii > 0 (the reference's visibility_filter):
    max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |viewspace_grad[:, :2]|; denom += 1
    (train_face.py:675-682, scene/gaussian_model.py:683-685).  viewspace_grad = means2D.grad [P,3]."""
    if not radii.is_cuda:
        raise RuntimeError("libigs_raster is CUDA-only: add_densification_stats needs CUDA tensors (no CPU fallback)")
    for t in (max_radii2D, xyz_gradient_accum, denom):
        if t.dtype != torch.float32 or not t.is_contiguous():
            raise ValueError("max_radii2D, xyz_gradient_accum and denom must be contiguous float32 (updated in place)")
    P, dev = radii.shape[0], radii.device
    r = radii if radii.dtype == torch.int32 and radii.is_contiguous() else radii.to(torch.int32).contiguous()
    g = _c(viewspace_grad)
    stream = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        N.check(N.lib().igs_densify_stats(P, r.data_ptr(), g.data_ptr(), max_radii2D.data_ptr(),
                                          xyz_gradient_accum.data_ptr(), denom.data_ptr(), stream), "densify stats")
