"""ctypes binding of libigs_raster.so (include/igs_raster.h).

The library is the product; there is NO fallback: if it is missing or cannot be
loaded every entry point raises.  PyTorch is only used by the callers for device
memory and streams.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# IGS_LIB: another build of the library (kernel experiments, profiles/tools); the default is the in-tree one
LIB_PATH = os.environ.get("IGS_LIB") or os.path.join(_HERE, "libigs_raster.so")

IGS_OK, IGS_ERR_INVALID, IGS_ERR_CUDA, IGS_ERR_CAPACITY = 0, -1, -2, -3
MAX_EXTRA_DIMS = 34

_p = C.c_void_p


class IgsFwdArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("ED", C.c_int32),
        ("W", C.c_int32), ("H", C.c_int32),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32), ("normalize_norm", C.c_int32),
        ("bg", _p), ("viewmatrix", _p), ("projmatrix", _p), ("campos", _p),
        ("means3D", _p), ("shs", _p), ("colors_precomp", _p), ("opacities", _p), ("scales", _p),
        ("rotations", _p), ("cov3Ds_precomp", _p), ("norm3Ds_precomp", _p), ("extra_attrs", _p),
        ("out_color", _p), ("out_depth", _p), ("out_norm", _p), ("out_alpha", _p), ("out_extra", _p),
        ("radii", _p),
        ("geom", _p), ("img", _p), ("bin", _p), ("sort", _p), ("ctrl", _p),
        ("capacity", C.c_int64),
        ("num_rendered_pinned", _p),
        ("ctrl_clean", C.c_int32),
        ("defer_count", C.c_int32),
        ("write_keys", C.c_int32),
    ]


class IgsBwdArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("ED", C.c_int32),
        ("W", C.c_int32), ("H", C.c_int32),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("debug", C.c_int32), ("normalize_norm", C.c_int32),
        ("num_rendered", C.c_int64),
        ("bg", _p), ("viewmatrix", _p), ("projmatrix", _p), ("campos", _p),
        ("means3D", _p), ("shs", _p), ("colors_precomp", _p), ("scales", _p), ("rotations", _p),
        ("cov3Ds_precomp", _p), ("norm3Ds_precomp", _p), ("extra_attrs", _p), ("radii", _p),
        ("geom", _p), ("img", _p), ("bin", _p), ("out_alpha", _p), ("out_norm", _p),
        ("dL_dcolor", _p), ("dL_ddepth", _p), ("dL_dnorm", _p), ("dL_dalpha", _p), ("dL_dextra", _p),
        ("accum", _p),
        ("dL_dmeans2D", _p), ("dL_dcolors", _p), ("dL_dopacity", _p), ("dL_dmeans3D", _p),
        ("dL_dcov3D", _p), ("dL_dnorm3D", _p), ("dL_dsh", _p), ("dL_dscales", _p),
        ("dL_drotations", _p), ("dL_dextra_attrs", _p),
        ("accum_clean", C.c_int32),
    ]


# every symbol include/igs_raster.h declares, with its ctypes signature
SYMBOLS = {
    "igs_version": (C.c_char_p, []),
    "igs_last_error": (C.c_char_p, []),
    "igs_geom_bytes": (C.c_size_t, [C.c_int32]),
    "igs_img_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "igs_img_info_offset": (C.c_size_t, [C.c_int32, C.c_int32]),
    "igs_bin_bytes": (C.c_size_t, [C.c_int64]),
    "igs_sort_bytes": (C.c_size_t, [C.c_int64]),
    "igs_ctrl_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "igs_accum_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "igs_geom_layout": (None, [C.c_int32, C.POINTER(C.c_int64)]),
    "igs_bin_layout": (None, [C.c_int64, C.POINTER(C.c_int64)]),
    "igs_img_layout": (None, [C.c_int32, C.c_int32, C.POINTER(C.c_int64)]),
    "igs_forward": (C.c_int, [C.POINTER(IgsFwdArgs), _p, C.POINTER(C.c_int64)]),
    "igs_forward_bin_render": (C.c_int, [C.POINTER(IgsFwdArgs), _p]),
    "igs_forward_count": (C.c_int, [C.POINTER(IgsFwdArgs), _p, C.POINTER(C.c_int64)]),
    "igs_point_offsets": (C.c_int, [C.c_int32, _p, _p]),
    "igs_backward": (C.c_int, [C.POINTER(IgsBwdArgs), _p]),
    "igs_profile_enable": (C.c_int, [C.c_int]),
    "igs_profile_read": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "igs_launch_count": (C.c_uint64, []),
    "igs_debug_trace": (C.c_int, [_p]),
    "igs_mark_visible": (C.c_int, [C.c_int32, _p, _p, _p, _p, _p]),
    "igs_compose_frame": (C.c_int, [_p, _p, _p, _p, _p, C.c_int32, C.c_int32, C.c_int32, _p, _p, _p]),
    "igs_activate_forward": (C.c_int, [C.c_int32] + [_p] * 12),
    "igs_activate_backward": (C.c_int, [C.c_int32] + [_p] * 14),
    "igs_densify_stats": (C.c_int, [C.c_int32] + [_p] * 6),
    "igs_allreduce_nvls": (C.c_int, [_p, _p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _p]),
    "igs_l1_ssim_partials": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "igs_l1_ssim_forward": (C.c_int, [_p, _p, C.c_int32, C.c_int32, C.c_int32, _p, _p, _p, _p, _p]),
    "igs_l1_ssim_finish": (C.c_int, [_p, C.c_int32, C.c_int32, C.c_int32, C.c_float, _p, _p]),
    "igs_l1_ssim_backward": (C.c_int, [_p, _p, _p, _p, _p, C.c_int32, C.c_int32, C.c_int32, C.c_float, _p, _p, _p]),
    "igs_l1_ssim_forward_static": (C.c_int, [_p, _p, C.c_int32, C.c_int32, C.c_int32, C.c_float, _p, _p, _p, _p, _p]),
    "igs_geom_loss_workspace": (C.c_size_t, [C.c_int32]),
    "igs_geom_loss_forward": (C.c_int, [_p] * 8 + [C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_float, _p, _p, _p]),
    "igs_geom_loss_backward": (C.c_int, [_p] * 6 + [C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_float] + [_p] * 6),
}

_lib = None


def lib():
    """Load libigs_raster.so; raises (never falls back) when it is unavailable."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m instag_b200.build` "
                "(there is no CPU or PyTorch fallback for the rasterizer)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def last_error():
    return lib().igs_last_error().decode("utf-8", "replace")


def check(status, what):
    if status != IGS_OK:
        raise RuntimeError(f"libigs_raster {what} failed ({status}): {last_error()}")


NUM_STAGES = 7
STAGE_NAMES = ("preprocess", "duplicate", "sort", "tile_ranges", "render", "render_backward",
               "preprocess_backward")


def profile_enable(on=True):
    lib().igs_profile_enable(int(bool(on)))


def profile_read():
    """(mean stage durations in ms since profile_enable, frames averaged); waits for the events."""
    ms = (C.c_float * NUM_STAGES)()
    n = C.c_int32(0)
    check(lib().igs_profile_read(ms, C.byref(n)), "profile_read")
    return dict(zip(STAGE_NAMES, [float(x) for x in ms])), int(n.value)


def launch_count():
    return int(lib().igs_launch_count())


def geom_layout(P):
    off = (C.c_int64 * 4)()
    lib().igs_geom_layout(P, off)
    return dict(splat=off[0], tiles_touched=off[1], point_offsets=off[2], clamped=off[3])


def bin_layout(capacity):
    off = (C.c_int64 * 2)()
    lib().igs_bin_layout(capacity, off)
    return dict(point_list=off[0], keys=off[1])


def img_layout(W, H):
    off = (C.c_int64 * 2)()
    lib().igs_img_layout(W, H, off)
    return dict(ranges=off[0], n_contrib=off[1])
