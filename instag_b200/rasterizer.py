"""Host side of the drop-in: the reference's Python API over libigs_raster.

Mirrors submodules/diff-gaussian-rasterization/diff_gauss/__init__.py of the
reference ("DGR"): same names, argument order, defaults, return tuple, error
texts and debug-snapshot behaviour, so that gaussian_renderer/__init__.py:15
(`from diff_gauss import GaussianRasterizationSettings, GaussianRasterizer`)
runs unmodified.  PyTorch supplies device memory, streams and autograd; every
kernel is in libigs_raster.so (no fallback path exists).
"""
import ctypes as C
import os
import threading
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _native as N


def cpu_deep_copy_tuple(input_tuple):
    # DGR diff_gauss/__init__.py:17-19
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


_DEFER_SLOTS = 8  # frames a run-ahead caller may leave unchecked


class _DeviceScratch:
    """Reusable scratch of one (device, stream, host thread): the unsorted instance words, the
    control block (tile counters and cursors), the gradient accumulators, the pinned num_rendered
    slot, and the instance capacity to allocate for each (P, H, W) seen so far."""

    def __init__(self, device):
        self.device = device
        self.sort = None
        self.ctrl = None
        self.accum = None
        # word 0: the count of a call that waits; words 4..: a ring for calls that do not (defer=True)
        self.pinned = torch.zeros(4 + _DEFER_SLOTS, dtype=torch.int32).pin_memory()
        self.pinned_ptr = self.pinned.data_ptr()
        self.defer_serial = 0
        self.defer_pending = []   # (serial, slot, capacity, hint_key), oldest first
        self.defer_last = None    # newest count held against its capacity
        self.capacity_hint = {}
        self.sizes = {}
        self.zero_imgs = {}
        self.dirty = set()
        self.ctrl_key = None
        # argument blocks are reused call after call (the library reads them before it returns)
        self.fwd_args, self.bwd_args = N.IgsFwdArgs(), N.IgsBwdArgs()
        self.fwd_args_ref, self.bwd_args_ref = C.byref(self.fwd_args), C.byref(self.bwd_args)
        self.n_out = C.c_int64(0)
        self.n_out_ref = C.byref(self.n_out)

    def buf(self, name, nbytes):
        """Reusable scratch.  `ctrl` and `accum` are handed to the library zeroed and come back
        clean from every successful call (the kernels put counters / rows back to zero), which the
        C ABI's ctrl_clean / accum_clean flags tell it; `self.dirty` names the ones a failed call
        may have left in an unknown state."""
        t = getattr(self, name)
        if t is None or t.numel() < nbytes:
            t = torch.zeros(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=self.device)
            setattr(self, name, t)
            self.dirty.discard(name)
        elif name in self.dirty:
            t.zero_()
            self.dirty.discard(name)
        return t

    def zeros_like_img(self, C_, H, W):
        key = (C_, H, W)
        z = self.zero_imgs.get(key)
        if z is None:
            z = torch.zeros((C_, H, W), dtype=torch.float32, device=self.device)
            self.zero_imgs[key] = z
        return z


_scratch = {}


def _get_scratch(device, stream=0):
    """Scratch is private to a (device, stream, host thread): calls on different streams may run
    concurrently on the GPU and must not share the sort / control / accumulator buffers."""
    key = (device, stream, threading.get_ident())
    s = _scratch.get(key)
    if s is None:
        s = _scratch[key] = _DeviceScratch(device)
    return s


class _NullCtx:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NULL_CTX = _NullCtx()


def _on_device(dev):
    """torch.cuda.device(dev) only when dev is not already the current device."""
    if dev.index is None or dev.index == torch.cuda.current_device():
        return _NULL_CTX
    return torch.cuda.device(dev)


def _ptr(t):
    """Device pointer of a tensor, NULL for the reference's empty-tensor 'None' sentinels."""
    return t.data_ptr() if (t is not None and t.numel() != 0) else None


def _f32c(t):
    """fp32, contiguous and 16-byte aligned (the kernels use vector accesses; a contiguous view at an
    odd storage offset -- which the reference accepts -- is copied)."""
    if t.dtype != torch.float32:
        t = t.float()
    if not t.is_contiguous():
        return t.contiguous()
    if t.numel() and t.data_ptr() % 16:
        return t.clone()
    return t


class GaussianRasterizationSettings(NamedTuple):
    # DGR diff_gauss/__init__.py:180-192
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _forward_native(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier,
                    cov3Ds_precomp, norm3Ds_precomp, extra_attrs, ED, viewmatrix, projmatrix, tanfovx,
                    tanfovy, image_height, image_width, sh, sh_degree, campos, prefiltered, debug,
                    normalize_norm=False, write_keys=False, defer=False):
    """Equivalent of _C.rasterize_gaussians (DGR rasterize_points.cu:35-133).  normalize_norm: the
    kernel also applies the F.normalize(norm, dim=0) of DGR diff_gauss/__init__.py:48.  write_keys:
    also leave the reference's sorted 64-bit keys in the binning buffer (tests compare them).
    defer: do not wait for the instance count (the reference's one device->host sync,
    rasterizer_impl.cu:284): the frame is queued with the instance workspaces of the last frame of
    this shape that did wait, num_rendered comes back as -1 (pass it on to `_backward_native`), and
    `check_deferred` later holds the posted count against those workspaces.  The first call of a
    shape waits whatever `defer` says."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("libigs_raster is CUDA-only: means3D must be a CUDA tensor (no CPU fallback)")
    lib = N.lib()
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    M = sh.size(1) if sh.numel() != 0 else 0
    empty = torch.empty
    color = empty((3, H, W), dtype=torch.float32, device=dev)
    depth = empty((1, H, W), dtype=torch.float32, device=dev)
    norm = empty((3, H, W), dtype=torch.float32, device=dev)
    alpha = empty((1, H, W), dtype=torch.float32, device=dev)
    extra = empty((ED, H, W) if ED > 0 else (0,), dtype=torch.float32, device=dev)
    radii = empty((P,), dtype=torch.int32, device=dev)

    stream = torch.cuda.current_stream(dev).cuda_stream
    sc = _get_scratch(dev, stream)
    hint_key = (P, H, W)
    sizes = sc.sizes.get(hint_key)
    if sizes is None:
        sizes = sc.sizes[hint_key] = (lib.igs_geom_bytes(P), lib.igs_img_bytes(W, H), lib.igs_ctrl_bytes(P, W, H))
    geom_bytes, img_bytes, ctrl_bytes = sizes
    capacity = sc.capacity_hint.get(hint_key)
    if capacity is None:
        capacity = max(8 * P, 1 << 16)
    # forward state kept for backward: one allocation, [geom | img] (both sizes are multiples of 256)
    state = empty((geom_bytes + img_bytes,), dtype=torch.uint8, device=dev)
    geom, img = state[:geom_bytes], state[geom_bytes:]
    binb = empty((lib.igs_bin_bytes(capacity),), dtype=torch.uint8, device=dev)
    sort = sc.buf("sort", lib.igs_sort_bytes(capacity))
    if sc.ctrl_key != hint_key:  # another (P, H, W) lays [ctrl] out differently: what is clean for one is not for the other
        sc.dirty.add("ctrl")
        sc.ctrl_key = hint_key
    ctrl = sc.buf("ctrl", ctrl_bytes)

    keep = (_f32c(bg), _f32c(viewmatrix), _f32c(projmatrix), _f32c(campos), _f32c(means3D), _f32c(sh),
            _f32c(colors_precomp), _f32c(opacities), _f32c(scales), _f32c(rotations),
            _f32c(cov3Ds_precomp), _f32c(norm3Ds_precomp), _f32c(extra_attrs))
    a = sc.fwd_args
    a.P, a.D, a.M, a.ED, a.W, a.H = P, int(sh_degree), M, ED, W, H
    a.tan_fovx, a.tan_fovy, a.scale_modifier = tanfovx, tanfovy, scale_modifier
    a.prefiltered, a.debug, a.normalize_norm = int(bool(prefiltered)), int(bool(debug)), int(bool(normalize_norm))
    (a.bg, a.viewmatrix, a.projmatrix, a.campos, a.means3D, a.shs, a.colors_precomp, a.opacities, a.scales,
     a.rotations, a.cov3Ds_precomp, a.norm3Ds_precomp, a.extra_attrs) = [_ptr(t) for t in keep]
    a.out_color, a.out_depth, a.out_norm, a.out_alpha = color.data_ptr(), depth.data_ptr(), norm.data_ptr(), alpha.data_ptr()
    a.out_extra = _ptr(extra)
    a.radii = radii.data_ptr() if P > 0 else None
    geom_ptr = state.data_ptr()
    a.geom, a.img, a.bin, a.sort, a.ctrl = geom_ptr, geom_ptr + geom_bytes, binb.data_ptr(), sort.data_ptr(), ctrl.data_ptr()
    a.capacity = capacity
    defer = bool(defer) and sc.capacity_hint.get(hint_key) is not None
    if defer:
        if len(sc.defer_pending) >= _DEFER_SLOTS:
            check_deferred(dev, keep=_DEFER_SLOTS - 1, stream=stream)
        slot = sc.defer_serial % _DEFER_SLOTS
        a.num_rendered_pinned = sc.pinned_ptr + 4 * (4 + slot)
    else:
        a.num_rendered_pinned = sc.pinned_ptr
    a.ctrl_clean = 1
    a.defer_count, a.write_keys = int(defer), int(bool(write_keys))
    sc.dirty.add("ctrl")  # until the call has returned without an exception

    n = sc.n_out
    with _on_device(dev):
        st = lib.igs_forward(sc.fwd_args_ref, stream, sc.n_out_ref)
        if st == N.IGS_ERR_CAPACITY:
            capacity = int(n.value * 1.1) + 4096
            binb = empty((lib.igs_bin_bytes(capacity),), dtype=torch.uint8, device=dev)
            sort = sc.buf("sort", lib.igs_sort_bytes(capacity))
            a.bin, a.sort, a.capacity = binb.data_ptr(), sort.data_ptr(), capacity
            st = lib.igs_forward_bin_render(sc.fwd_args_ref, stream)
        if st != N.IGS_OK:
            N.check(st, "forward")
    sc.dirty.discard("ctrl")
    if defer:
        sc.defer_pending.append((sc.defer_serial, slot, capacity, hint_key))
        sc.defer_serial += 1
        return -1, color, depth, norm, alpha, radii, extra, geom, binb, img
    num_rendered = int(n.value)
    # next call with this shape: 25 % head-room over what this one needed
    sc.capacity_hint[hint_key] = max(int(num_rendered * 1.25) + 4096, 1 << 14)
    return num_rendered, color, depth, norm, alpha, radii, extra, geom, binb, img


def check_deferred(device, keep=0, stream=None):
    """Hold the counts of this (device, stream, thread)'s `defer=True` frames, all but the newest `keep`,
    against the instance workspaces they were queued with.  Waits for a count that has not been posted
    yet.  Returns the newest count seen (None if there is none); raises RuntimeError if a frame outgrew
    its workspaces -- it was not rendered and its backward stood down -- after raising the size the next
    frame of that shape gets."""
    device = torch.device(device)
    if stream is None:
        stream = torch.cuda.current_stream(device).cuda_stream
    sc = _get_scratch(device, stream)
    while len(sc.defer_pending) > keep:
        serial, slot, capacity, hint_key = sc.defer_pending.pop(0)
        word = sc.pinned[4 + slot]
        spins = 0
        while int(word) == -1:
            spins += 1
            if spins % 1024 == 0 and torch.cuda.current_stream(device).query():
                if int(word) == -1:
                    raise RuntimeError("libigs_raster: a deferred frame finished without posting its instance count")
        n = int(word) & 0xFFFFFFFF
        if n > capacity:
            sc.capacity_hint[hint_key] = max(int(n * 1.25) + 4096, 1 << 14)
            raise RuntimeError(f"libigs_raster: deferred frame {serial} has {n} instances, its workspaces hold "
                               f"{capacity}; it was not rendered (render it again)")
        sc.defer_last = n
    return sc.defer_last


def gradient_layout(P, M, ED):
    """Layout of the ONE allocation behind the ten gradient tensors `_backward_native` returns, in
    its return order (means2D, colours, opacity, means3D, cov3D, norm3D, sh, scales, rotations,
    extra): (starts, widths, total floats, floats of the parameter prefix).  Every block starts
    16-byte aligned, as the C ABI requires.  The gradients of trainable parameters come first and
    are contiguous (means3D, colours, opacity, SH, scales, rotations, extra), so that a view-sharded
    batch can all-reduce just that prefix (sharding.py); means2D (per-view densification statistics
    only) and the cov3D / norm3D intermediates follow."""
    widths = [3, 3, 1, 3, 6, 3, 3 * M, 3, 4, ED]
    starts, o, param_end = [0] * len(widths), 0, 0
    for i in (3, 1, 2, 6, 7, 8, 9, 0, 4, 5):
        if i == 0:
            param_end = o
        starts[i] = o
        o += (P * widths[i] + 3) // 4 * 4
    return starts, widths, o, param_end


# optional allocator of the packed gradient block: fn(n_floats, device) -> flat float32 tensor.
# sharding.NvlsGradientBlock installs one so that the block lives in NVLink multicast memory.
_grad_alloc = None


def set_gradient_allocator(fn):
    global _grad_alloc
    prev, _grad_alloc = _grad_alloc, fn
    if _host_binding is not None or host_path() == "cxx":
        binding().set_gradient_allocator(fn)
    return prev


# ---- the compiled twin of _forward_native / _backward_native / _RasterizeGaussians ----
_host_binding = None


def host_path():
    """Which host side runs the autograd node: "cxx" (csrc/igs_torch.cpp, the default) or "python"
    (this file; IGS_HOST=python, and always with raster_settings.debug, whose snapshot dumps live
    here).  Both end in the same igs_forward / igs_backward calls."""
    return os.environ.get("IGS_HOST", "cxx")


def binding():
    """The torch extension module built from csrc/igs_torch.cpp; raises when it has not been built."""
    global _host_binding
    if _host_binding is None:
        N.lib()  # libigs_raster.so first: the extension links against it
        try:
            from . import _igs_torch
        except ImportError as ex:
            raise RuntimeError("instag_b200/_igs_torch.so is missing or does not load: build it with "
                               "`python -m instag_b200.build` (or run with IGS_HOST=python)") from ex
        _host_binding = _igs_torch
        if _grad_alloc is not None:
            _igs_torch.set_gradient_allocator(_grad_alloc)
    return _host_binding


def _backward_native(bg, means3D, radii, colors_precomp, scales, rotations, extra_attrs, scale_modifier,
                     cov3Ds_precomp, norm3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy,
                     dL_dcolor, dL_ddepth, dL_dnorm, dL_dalpha, dL_dextra, sh, sh_degree, campos, geom,
                     num_rendered, binb, img, alpha, debug, norm_out=None):
    """Equivalent of _C.rasterize_gaussians_backward (DGR rasterize_points.cu:135-239).  norm_out: the
    unit normals a normalize_norm forward returned; dL_dnorm is then the gradient w.r.t. them."""
    lib = N.lib()
    dev = means3D.device
    P = means3D.size(0)
    H, W = alpha.size(1), alpha.size(2)
    ED = extra_attrs.size(1) if extra_attrs.numel() != 0 else 0
    M = sh.size(1) if sh.numel() != 0 else 0
    stream = torch.cuda.current_stream(dev).cuda_stream
    sc = _get_scratch(dev, stream)
    # one allocation for all per-Gaussian gradients, contiguous sub-blocks
    starts, widths, total, _ = gradient_layout(P, M, ED)
    flat = _grad_alloc(total, dev) if _grad_alloc is not None else torch.empty((total,), dtype=torch.float32, device=dev)
    outs = [flat[st:st + P * w].view(P, w) for st, w in zip(starts, widths)]
    (g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_norm3D, g_sh, g_scales, g_rot, g_extra) = outs
    g_sh = g_sh.view(P, M, 3)
    if ED == 0:
        g_extra = torch.empty(tuple(extra_attrs.shape), dtype=torch.float32, device=dev)  # no elements
    accum = sc.buf("accum", lib.igs_accum_bytes(P, ED))

    def grad(t, C_):
        return sc.zeros_like_img(C_, H, W) if t is None else _f32c(t)

    keep = [_f32c(bg), _f32c(viewmatrix), _f32c(projmatrix), _f32c(campos), _f32c(means3D), _f32c(sh),
            _f32c(colors_precomp), _f32c(scales), _f32c(rotations), _f32c(cov3Ds_precomp),
            _f32c(norm3Ds_precomp), _f32c(extra_attrs), radii, geom, img, binb, _f32c(alpha),
            grad(dL_dcolor, 3), grad(dL_ddepth, 1), grad(dL_dnorm, 3), grad(dL_dalpha, 1),
            grad(dL_dextra, ED) if ED > 0 else None]
    a = sc.bwd_args
    a.P, a.D, a.M, a.ED, a.W, a.H = P, int(sh_degree), M, ED, W, H
    a.tan_fovx, a.tan_fovy, a.scale_modifier, a.debug = tanfovx, tanfovy, scale_modifier, int(bool(debug))
    a.normalize_norm = int(norm_out is not None)
    a.out_norm = _ptr(_f32c(norm_out)) if norm_out is not None else None
    a.num_rendered = num_rendered
    (a.bg, a.viewmatrix, a.projmatrix, a.campos, a.means3D, a.shs, a.colors_precomp, a.scales, a.rotations,
     a.cov3Ds_precomp, a.norm3Ds_precomp, a.extra_attrs, a.radii, a.geom, a.img, a.bin, a.out_alpha,
     a.dL_dcolor, a.dL_ddepth, a.dL_dnorm, a.dL_dalpha, a.dL_dextra) = [_ptr(t) for t in keep]
    a.accum = accum.data_ptr()
    a.accum_clean = 1
    sc.dirty.add("accum")
    (a.dL_dmeans2D, a.dL_dcolors, a.dL_dopacity, a.dL_dmeans3D, a.dL_dcov3D, a.dL_dnorm3D, a.dL_dsh,
     a.dL_dscales, a.dL_drotations, a.dL_dextra_attrs) = [_ptr(t) for t in
                                                          (g_means2D, g_colors, g_opacity, g_means3D, g_cov3D,
                                                           g_norm3D, g_sh, g_scales, g_rot, g_extra)]
    with _on_device(dev):
        st = lib.igs_backward(sc.bwd_args_ref, stream)
        if st != N.IGS_OK:
            N.check(st, "backward")
    sc.dirty.discard("accum")
    return g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_norm3D, g_sh, g_scales, g_rot, g_extra


class _RasterizeGaussians(torch.autograd.Function):
    # DGR diff_gauss/__init__.py:52-178
    FUSE_NORMALIZE = False  # the reference Function returns the un-normalised normal image

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                norm3Ds_precomp, extra_attrs, raster_settings):
        assert extra_attrs.shape[0] == 0 or extra_attrs.shape[1] <= N.MAX_EXTRA_DIMS
        args = (
            raster_settings.bg, means3D, colors_precomp, opacities, scales, rotations,
            raster_settings.scale_modifier, cov3Ds_precomp, norm3Ds_precomp, extra_attrs,
            extra_attrs.shape[1] if extra_attrs.shape[0] != 0 else 0,
            raster_settings.viewmatrix, raster_settings.projmatrix, raster_settings.tanfovx,
            raster_settings.tanfovy, raster_settings.image_height, raster_settings.image_width, sh,
            raster_settings.sh_degree, raster_settings.campos, raster_settings.prefiltered,
            raster_settings.debug, ctx._forward_cls.FUSE_NORMALIZE)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # copy them before they can be corrupted
            try:
                res = _forward_native(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            res = _forward_native(*args)
        num_rendered, color, depth, norm, alpha, radii, extra, geomBuffer, binningBuffer, imgBuffer = res
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(radii)
        ctx.fused_norm = ctx._forward_cls.FUSE_NORMALIZE
        saved = (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, norm3Ds_precomp,
                 radii, extra_attrs, sh, geomBuffer, binningBuffer, imgBuffer, alpha)
        ctx.save_for_backward(*(saved + ((norm,) if ctx.fused_norm else ())))
        return color, depth, norm, alpha, radii, extra

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_depth, grad_out_norm, grad_out_alpha, _, grad_out_extra):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, norm3Ds_precomp, radii, extra_attrs, sh,
         geomBuffer, binningBuffer, imgBuffer, alpha) = ctx.saved_tensors[:13]
        norm_out = ctx.saved_tensors[13] if ctx.fused_norm else None
        args = (raster_settings.bg, means3D, radii, colors_precomp, scales, rotations, extra_attrs,
                raster_settings.scale_modifier, cov3Ds_precomp, norm3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy, grad_out_color,
                grad_out_depth, grad_out_norm, grad_out_alpha, grad_out_extra, sh, raster_settings.sh_degree,
                raster_settings.campos, geomBuffer, num_rendered, binningBuffer, imgBuffer, alpha,
                raster_settings.debug, norm_out)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                res = _backward_native(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            res = _backward_native(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
         grad_norm3Ds_precomp, grad_sh, grad_scales, grad_rotations, grad_extra_attrs) = res
        grads = (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                 grad_rotations, grad_cov3Ds_precomp, grad_norm3Ds_precomp, grad_extra_attrs, None)
        # gradients for inputs that do not require one (incl. the empty 'None'
        # sentinels) are dropped, as autograd would do with the reference's
        return tuple(g if (g is not None and need) else None for g, need in zip(grads, ctx.needs_input_grad))


class _RasterizeGaussiansUnitNormal(_RasterizeGaussians):
    """Same Function with the reference's trailing `norm = F.normalize(norm, p=2, dim=0)`
    (DGR diff_gauss/__init__.py:48) done inside the render kernels, forward and backward: three
    element-wise kernels, six backward kernels and four autograd nodes less per call."""
    FUSE_NORMALIZE = True


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        norm3Ds_precomp, extra_attrs, raster_settings):
    # DGR diff_gauss/__init__.py:21-50; the F.normalize of :48 is fused into the kernels
    rs = raster_settings
    if not rs.debug and host_path() == "cxx":
        return tuple(binding().rasterize(
            means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, norm3Ds_precomp,
            extra_attrs, rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos, float(rs.tanfovx), float(rs.tanfovy),
            float(rs.scale_modifier), int(rs.image_height), int(rs.image_width), int(rs.sh_degree),
            bool(rs.prefiltered), True))
    return _RasterizeGaussiansUnitNormal.apply(
        means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, norm3Ds_precomp,
        extra_attrs, raster_settings)


_NONE = None


def _none_tensor():
    """The reference's stand-in for an absent input, `torch.Tensor([])` (DGR diff_gauss/__init__.py:
    232-246); it is only ever read for its emptiness, so one instance serves every call."""
    global _NONE
    if _NONE is None:
        _NONE = torch.Tensor([])
    return _NONE


class GaussianRasterizer(nn.Module):
    # DGR diff_gauss/__init__.py:194-250
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            if not positions.is_cuda:
                raise RuntimeError("libigs_raster is CUDA-only: positions must be a CUDA tensor")
            pos, view, proj = _f32c(positions), _f32c(rs.viewmatrix), _f32c(rs.projmatrix)
            P = pos.size(0)
            visible = torch.empty((P,), dtype=torch.bool, device=pos.device)
            if P:
                with torch.cuda.device(pos.device):
                    N.check(N.lib().igs_mark_visible(P, pos.data_ptr(), view.data_ptr(), proj.data_ptr(),
                                                     visible.data_ptr(),
                                                     torch.cuda.current_stream(pos.device).cuda_stream),
                            "mark_visible")
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3Ds_precomp=None, norm3Ds_precomp=None, extra_attrs=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3Ds_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3Ds_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = _none_tensor()
        if colors_precomp is None:
            colors_precomp = _none_tensor()
        if scales is None:
            raise ValueError('To support norm and depth prediction, scales == None is not allowed')
        if rotations is None:
            raise ValueError('To support norm and depth prediction, rotations == None is not allowed')
        if cov3Ds_precomp is None:
            cov3Ds_precomp = _none_tensor()
        if norm3Ds_precomp is None:
            norm3Ds_precomp = _none_tensor()
        if extra_attrs is None:
            extra_attrs = _none_tensor()
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3Ds_precomp, norm3Ds_precomp, extra_attrs, raster_settings)
