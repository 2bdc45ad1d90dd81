"""TEST/BASELINE INFRASTRUCTURE -- torch front end of oracle/_ref/libdgr_ref.so,
the UNMODIFIED reference CUDA rasterizer compiled by oracle/Makefile.

Plays the role of the reference's rasterize_points.cu + diff_gauss/__init__.py
(DGR rasterize_points.cu:35-239, diff_gauss/__init__.py:21-250): allocates and
ZERO-FILLS the same tensors the reference's torch shim does (torch::full /
torch::zeros), hands raw pointers to CudaRasterizer::Rasterizer, and exposes the
intermediates (tiles_touched, point_offsets, keys, point_list, ranges,
n_contrib) by parsing the scratch buffers with the reference's own fromChunk.

Only tests/ and bench.py (--impl reference, GPU parity) import this module.
"""
import ctypes as C
import os
from typing import NamedTuple

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libdgr_ref.so")
_lib = None
_ALLOC = C.CFUNCTYPE(C.c_void_p, C.c_size_t, C.c_int)
_vp, _f, _i = C.c_void_p, C.c_float, C.c_int


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{LIB_PATH} missing: run `make -C oracle ref` where /root/reference exists")
        l = C.CDLL(LIB_PATH)
        l.dgr_ref_forward.restype = _i
        l.dgr_ref_forward.argtypes = [_ALLOC, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp,
                                      _vp, _vp, _vp, _vp, _vp, _f, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i]
        l.dgr_ref_backward.restype = _i
        l.dgr_ref_backward.argtypes = [_i, _i, _i, _i, _i, _vp, _i, _i] + [_vp] * 4 + [_f] + [_vp] * 7 + [_f, _f] + \
                                      [_vp] * 22 + [_i]
        l.dgr_ref_mark_visible.restype = None
        l.dgr_ref_mark_visible.argtypes = [_i, _vp, _vp, _vp, _vp]
        for n in ("dgr_ref_geom_layout", "dgr_ref_binning_layout", "dgr_ref_image_layout"):
            getattr(l, n).restype = None
            getattr(l, n).argtypes = [_vp, C.c_size_t, C.POINTER(C.c_int64)]
        _lib = l
    return _lib


def _p(t):
    return t.data_ptr() if (t is not None and t.numel() != 0) else None


def _c(t):
    return t.contiguous()


def forward_raw(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3Ds_precomp,
                norm3Ds_precomp, extra_attrs, ED, viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh, degree,
                campos, prefiltered, debug):
    """What _C.rasterize_gaussians returns (DGR rasterize_points.cu:35-133)."""
    dev = means3D.device
    P = means3D.size(0)
    f32 = dict(dtype=torch.float32, device=dev)
    out_color = torch.full((3, H, W), 0.0, **f32)
    out_depth = torch.full((1, H, W), 0.0, **f32)
    out_alpha = torch.full((1, H, W), 0.0, **f32)
    out_norm = torch.full((3, H, W), 0.0, **f32)
    radii = torch.full((P,), 0, dtype=torch.int32, device=dev)
    out_extra = torch.full((ED, H, W), 0.0, **f32) if ED > 0 else torch.empty((0,), **f32)
    bufs = [torch.empty((0,), dtype=torch.uint8, device=dev) for _ in range(3)]

    def alloc(nbytes, which):
        bufs[which] = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
        return bufs[which].data_ptr()

    cb = _ALLOC(alloc)
    rendered = 0
    if P != 0:
        M = sh.size(1) if sh.numel() != 0 else 0
        keep = [_c(t) for t in (bg, means3D, sh, colors, opacity, scales, rotations, cov3Ds_precomp,
                                norm3Ds_precomp, extra_attrs, viewmatrix, projmatrix, campos)]
        (bg_, m_, sh_, col_, op_, sc_, rot_, cov_, nrm_, ex_, v_, pr_, cam_) = keep
        rendered = lib().dgr_ref_forward(
            cb, P, degree, M, ED, _p(bg_), W, H, _p(m_), _p(sh_), _p(col_), _p(op_), _p(sc_), scale_modifier,
            _p(rot_), _p(cov_), _p(nrm_), _p(ex_), _p(v_), _p(pr_), _p(cam_), tanfovx, tanfovy,
            int(bool(prefiltered)), _p(out_color), _p(out_depth), _p(out_norm), _p(out_alpha), _p(out_extra),
            _p(radii), int(bool(debug)))
        if rendered < 0:
            raise RuntimeError("reference forward threw")
    return rendered, out_color, out_depth, out_norm, out_alpha, radii, out_extra, bufs[0], bufs[1], bufs[2]


def backward_raw(bg, means3D, radii, colors, scales, rotations, extra_attrs, scale_modifier, cov3Ds_precomp,
                 norm3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy, dL_dcolor, dL_ddepth, dL_dnorm,
                 dL_dalpha, dL_dextra, sh, degree, campos, geom, R, binning, img, out_alpha, debug):
    """What _C.rasterize_gaussians_backward returns (DGR rasterize_points.cu:135-239)."""
    P = means3D.size(0)
    H, W = dL_dcolor.size(1), dL_dcolor.size(2)
    ED = extra_attrs.size(1) if extra_attrs.numel() != 0 else 0
    M = sh.size(1) if sh.numel() != 0 else 0
    o = dict(dtype=means3D.dtype, device=means3D.device)
    z = lambda *s: torch.zeros(s, **o)
    dL_dmeans3D, dL_dmeans2D, dL_dcolors = z(P, 3), z(P, 3), z(P, 3)
    dL_ddepths, dL_dconic, dL_dopacity = z(P, 1), z(P, 2, 2), z(P, 1)
    dL_dcov3D, dL_dnorm3D, dL_dsh = z(P, 6), z(P, 3), z(P, M, 3)
    dL_dscales, dL_drot = z(P, 3), z(P, 4)
    dL_dextra_attrs = z(P, ED) if ED > 0 else torch.empty((0,), **o)
    if P != 0:
        keep = [_c(t) for t in (bg, means3D, sh, colors, scales, rotations, cov3Ds_precomp, norm3Ds_precomp,
                                extra_attrs, viewmatrix, projmatrix, campos, radii, out_alpha, dL_dcolor,
                                dL_ddepth, dL_dnorm, dL_dalpha, dL_dextra)]
        (bg_, m_, sh_, col_, sc_, rot_, cov_, nrm_, ex_, v_, pr_, cam_, rad_, al_, dc_, dd_, dn_, da_, de_) = keep
        r = lib().dgr_ref_backward(
            P, degree, M, R, ED, _p(bg_), W, H, _p(m_), _p(sh_), _p(col_), _p(sc_), scale_modifier, _p(rot_),
            _p(cov_), _p(nrm_), _p(ex_), _p(v_), _p(pr_), _p(cam_), tanfovx, tanfovy, _p(rad_), _p(geom),
            _p(binning), _p(img), _p(al_), _p(dc_), _p(dd_), _p(dn_), _p(da_), _p(de_), _p(dL_dmeans2D),
            _p(dL_dconic), _p(dL_dopacity), _p(dL_dcolors), _p(dL_ddepths), _p(dL_dmeans3D), _p(dL_dcov3D),
            _p(dL_dnorm3D), _p(dL_dsh), _p(dL_dscales), _p(dL_drot), _p(dL_dextra_attrs), int(bool(debug)))
        if r < 0:
            raise RuntimeError("reference backward threw")
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dnorm3D, dL_dsh, dL_dscales,
            dL_drot, dL_dextra_attrs)


def intermediates(P, H, W, num_rendered, geom, binning, img):
    """Views of the reference's scratch buffers (layout by its own fromChunk)."""
    def view(buf, off, dtype, count, shape=None):
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        t = buf[off:off + nbytes].view(dtype)
        return t.view(*shape) if shape else t

    off = (C.c_int64 * 9)()
    lib().dgr_ref_geom_layout(geom.data_ptr(), P, off)
    out = dict(
        depths=view(geom, off[0], torch.float32, P), clamped=view(geom, off[1], torch.uint8, 3 * P, (P, 3)),
        means2D=view(geom, off[2], torch.float32, 2 * P, (P, 2)), cov3D=view(geom, off[3], torch.float32, 6 * P, (P, 6)),
        conic_opacity=view(geom, off[4], torch.float32, 4 * P, (P, 4)), rgb=view(geom, off[5], torch.float32, 3 * P, (P, 3)),
        norm3D=view(geom, off[6], torch.float32, 3 * P, (P, 3)), tiles_touched=view(geom, off[7], torch.int32, P),
        point_offsets=view(geom, off[8], torch.int32, P))
    if num_rendered > 0:
        offb = (C.c_int64 * 4)()
        lib().dgr_ref_binning_layout(binning.data_ptr(), num_rendered, offb)
        out["point_list"] = view(binning, offb[0], torch.int32, num_rendered)
        out["keys"] = view(binning, offb[2], torch.int64, num_rendered)
    else:
        out["point_list"] = torch.empty(0, dtype=torch.int32, device=geom.device)
        out["keys"] = torch.empty(0, dtype=torch.int64, device=geom.device)
    offi = (C.c_int64 * 2)()
    lib().dgr_ref_image_layout(img.data_ptr(), H * W, offi)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out["n_contrib"] = view(img, offi[0], torch.int32, H * W)
    out["ranges"] = view(img, offi[1], torch.int32, 2 * T, (T, 2))
    return out


class _RefRasterize(torch.autograd.Function):
    """DGR diff_gauss/__init__.py:52-178 over the compiled reference."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                norm3Ds_precomp, extra_attrs, rs):
        ED = extra_attrs.shape[1] if extra_attrs.shape[0] != 0 else 0
        (n, color, depth, norm, alpha, radii, extra, geom, binning, img) = forward_raw(
            rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            norm3Ds_precomp, extra_attrs, ED, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
            rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        ctx.rs, ctx.num_rendered = rs, n
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, norm3Ds_precomp, radii,
                              extra_attrs, sh, geom, binning, img, alpha)
        return color, depth, norm, alpha, radii, extra

    @staticmethod
    def backward(ctx, g_color, g_depth, g_norm, g_alpha, _, g_extra):
        rs = ctx.rs
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, norm3Ds_precomp, radii, extra_attrs, sh, geom,
         binning, img, alpha) = ctx.saved_tensors
        (g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_norm3D, g_sh, g_scales, g_rot, g_ex) = backward_raw(
            rs.bg, means3D, radii, colors_precomp, scales, rotations, extra_attrs, rs.scale_modifier,
            cov3Ds_precomp, norm3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, g_color, g_depth,
            g_norm, g_alpha, g_extra, sh, rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, alpha,
            rs.debug)
        grads = (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rot, g_cov3D, g_norm3D, g_ex, None)
        return tuple(g if need else None for g, need in zip(grads, ctx.needs_input_grad))


class RefGaussianRasterizer(torch.nn.Module):
    """Same call surface as the reference's GaussianRasterizer (DGR diff_gauss/__init__.py:194-250)."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            P = positions.size(0)
            present = torch.full((P,), False, dtype=torch.bool, device=positions.device)
            if P:
                pos, v, pr = _c(positions), _c(rs.viewmatrix), _c(rs.projmatrix)
                lib().dgr_ref_mark_visible(P, pos.data_ptr(), v.data_ptr(), pr.data_ptr(), present.data_ptr())
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3Ds_precomp=None, norm3Ds_precomp=None, extra_attrs=None):
        e = lambda t: torch.Tensor([]) if t is None else t
        color, depth, norm, alpha, radii, extra = _RefRasterize.apply(
            means3D, means2D, e(shs), e(colors_precomp), opacities, scales, rotations, e(cov3Ds_precomp),
            e(norm3Ds_precomp), e(extra_attrs), self.raster_settings)
        norm = torch.nn.functional.normalize(norm, p=2, dim=0)
        return color, depth, norm, alpha, radii, extra
