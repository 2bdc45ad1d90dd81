"""TEST INFRASTRUCTURE -- numpy/ctypes front end of the C restatement.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
legs may import this module (see oracle/cpu_raster.h for the header that says
what the oracle restates and how its parity is pinned).  The product path
(instag_b200/, diff_gauss/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class _Scene(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("ED", C.c_int),
        ("W", C.c_int), ("H", C.c_int),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("campos", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p), ("cov3Ds_precomp", C.c_void_p),
        ("norm3Ds_precomp", C.c_void_p), ("extra_attrs", C.c_void_p),
    ]


class _Geom(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "radii", "means2D", "depths", "cov3D", "norm3D", "conic_opacity", "rgb", "clamped",
        "tiles_touched", "point_offsets")]


def lib():
    """Load (building if needed) oracle/liboracle_cpu.so."""
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle_cpu.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", _HERE, "cpu"])
        _LIB = C.CDLL(path)
        _LIB.orc_scan.restype = C.c_int64
        _LIB.orc_num_threads.restype = C.c_int
    return _LIB


def num_threads():
    return int(lib().orc_num_threads())


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Scene:
    """Inputs of one rasterization, as contiguous float32 numpy arrays."""

    def __init__(self, *, W, H, tanfovx, tanfovy, bg, viewmatrix, projmatrix, campos, means3D,
                 opacities, scales, rotations, sh_degree=0, shs=None, colors_precomp=None,
                 cov3Ds_precomp=None, norm3Ds_precomp=None, extra_attrs=None, scale_modifier=1.0):
        self.W, self.H = int(W), int(H)
        self.tanfovx, self.tanfovy = float(tanfovx), float(tanfovy)
        self.scale_modifier = float(scale_modifier)
        self.bg = _f32(bg)
        self.viewmatrix = _f32(viewmatrix).reshape(16)
        self.projmatrix = _f32(projmatrix).reshape(16)
        self.campos = _f32(campos)
        self.means3D = _f32(means3D)
        self.opacities = _f32(opacities).reshape(-1)
        self.scales = _f32(scales)
        self.rotations = _f32(rotations)
        self.shs = _f32(shs)
        self.colors_precomp = _f32(colors_precomp)
        self.cov3Ds_precomp = _f32(cov3Ds_precomp)
        self.norm3Ds_precomp = _f32(norm3Ds_precomp)
        self.extra_attrs = _f32(extra_attrs)
        self.P = int(self.means3D.shape[0])
        self.D = int(sh_degree)
        self.M = int(self.shs.shape[1]) if self.shs is not None else 0
        self.ED = int(self.extra_attrs.shape[1]) if self.extra_attrs is not None else 0
        assert (self.shs is None) != (self.colors_precomp is None)

    def c_struct(self):
        s = _Scene()
        s.P, s.D, s.M, s.ED, s.W, s.H = self.P, self.D, self.M, self.ED, self.W, self.H
        s.tan_fovx, s.tan_fovy, s.scale_modifier = self.tanfovx, self.tanfovy, self.scale_modifier
        for name in ("bg", "viewmatrix", "projmatrix", "campos", "means3D", "shs", "colors_precomp",
                     "opacities", "scales", "rotations", "cov3Ds_precomp", "norm3Ds_precomp",
                     "extra_attrs"):
            setattr(s, name, _ptr(getattr(self, name)))
        return s


def _geom_struct(g):
    s = _Geom()
    for name, _ in _Geom._fields_:
        setattr(s, name, _ptr(g[name]))
    return s


def preprocess(scene):
    """K1 + K2.  Returns the per-Gaussian state dict (+ num_rendered)."""
    P = scene.P
    g = {
        "radii": np.zeros(P, np.int32), "means2D": np.zeros((P, 2), np.float32),
        "depths": np.zeros(P, np.float32), "cov3D": np.zeros((P, 6), np.float32),
        "norm3D": np.zeros((P, 3), np.float32), "conic_opacity": np.zeros((P, 4), np.float32),
        "rgb": np.zeros((P, 3), np.float32), "clamped": np.zeros((P, 3), np.uint8),
        "tiles_touched": np.zeros(P, np.uint32), "point_offsets": np.zeros(P, np.uint32),
    }
    cs, gs = scene.c_struct(), _geom_struct(g)
    lib().orc_preprocess(C.byref(cs), C.byref(gs))
    g["num_rendered"] = int(lib().orc_scan(P, _ptr(g["tiles_touched"]), _ptr(g["point_offsets"])))
    return g


def forward(scene, want_border=True):
    """Full forward.  Returns a dict with every intermediate the reference has."""
    g = preprocess(scene)
    N = g["num_rendered"]
    gx, gy = (scene.W + 15) // 16, (scene.H + 15) // 16
    HW = scene.H * scene.W
    out = dict(g)
    out["keys_unsorted"] = np.zeros(max(N, 1), np.uint64)
    out["vals_unsorted"] = np.zeros(max(N, 1), np.uint32)
    out["keys"] = np.zeros(max(N, 1), np.uint64)
    out["point_list"] = np.zeros(max(N, 1), np.uint32)
    out["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    cs, gs = scene.c_struct(), _geom_struct(g)
    lib().orc_bin(C.byref(cs), C.byref(gs), C.c_int64(N), _ptr(out["keys_unsorted"]),
                  _ptr(out["vals_unsorted"]), _ptr(out["keys"]), _ptr(out["point_list"]),
                  _ptr(out["ranges"]))
    ED = scene.ED
    out["color"] = np.zeros((3, scene.H, scene.W), np.float32)
    out["depth"] = np.zeros((1, scene.H, scene.W), np.float32)
    out["norm"] = np.zeros((3, scene.H, scene.W), np.float32)
    out["alpha"] = np.zeros((1, scene.H, scene.W), np.float32)
    out["extra"] = np.zeros((max(ED, 1), scene.H, scene.W), np.float32)
    out["n_contrib"] = np.zeros(HW, np.uint32)
    out["borderline"] = np.zeros(HW, np.uint8) if want_border else None
    counters = np.zeros(2, np.int64)
    lib().orc_render(C.byref(cs), C.byref(gs), _ptr(out["ranges"]), _ptr(out["point_list"]),
                     _ptr(out["color"]), _ptr(out["depth"]), _ptr(out["norm"]), _ptr(out["alpha"]),
                     _ptr(out["extra"]), _ptr(out["n_contrib"]), _ptr(out["borderline"]),
                     _ptr(counters))
    if ED == 0:
        out["extra"] = np.zeros((0,), np.float32)
    for k in ("keys_unsorted", "vals_unsorted", "keys", "point_list"):
        out[k] = out[k][:N]
    out["pairs_evaluated"], out["pairs_blended"] = int(counters[0]), int(counters[1])
    return out


def backward(scene, fwd, dL_dcolor, dL_ddepth, dL_dnorm, dL_dalpha, dL_dextra=None):
    """K7 + K8 + K9 on the state returned by forward().  Gradients w.r.t. the
    (un-normalised) outputs of the rasterizer, as in _RasterizeGaussians.backward."""
    P, ED, M = scene.P, scene.ED, scene.M
    z = lambda *s: np.zeros(s, np.float32)
    r = {"dL_dmeans2D": z(P, 3), "dL_dconic": z(P, 4), "dL_dopacity": z(P, 1),
         "dL_dcolors": z(P, 3), "dL_ddepths": z(P, 1), "dL_dnorm3D": z(P, 3),
         "dL_dextra": z(P, max(ED, 1)), "dL_dmeans3D": z(P, 3), "dL_dcov3D": z(P, 6),
         "dL_dsh": z(P, max(M, 1), 3), "dL_dscales": z(P, 3), "dL_drotations": z(P, 4)}
    g = {k: fwd[k] for k, _ in _Geom._fields_}
    cs, gs = scene.c_struct(), _geom_struct(g)
    dpe = _f32(dL_dextra) if ED > 0 else None
    dc, dd, dn, da = _f32(dL_dcolor), _f32(dL_ddepth), _f32(dL_dnorm), _f32(dL_dalpha)
    lib().orc_render_backward(
        C.byref(cs), C.byref(gs), _ptr(fwd["ranges"]), _ptr(np.ascontiguousarray(fwd["point_list"])),
        _ptr(fwd["alpha"]), _ptr(fwd["n_contrib"]), _ptr(dc), _ptr(dd), _ptr(dn), _ptr(da), _ptr(dpe),
        _ptr(r["dL_dmeans2D"]), _ptr(r["dL_dconic"]), _ptr(r["dL_dopacity"]), _ptr(r["dL_dcolors"]),
        _ptr(r["dL_ddepths"]), _ptr(r["dL_dnorm3D"]), _ptr(r["dL_dextra"]))
    lib().orc_preprocess_backward(
        C.byref(cs), C.byref(gs), _ptr(r["dL_dmeans2D"]), _ptr(r["dL_dconic"]), _ptr(r["dL_dcolors"]),
        _ptr(r["dL_ddepths"]), _ptr(r["dL_dnorm3D"]), _ptr(r["dL_dmeans3D"]), _ptr(r["dL_dcov3D"]),
        _ptr(r["dL_dsh"]), _ptr(r["dL_dscales"]), _ptr(r["dL_drotations"]))
    if ED == 0:
        r["dL_dextra"] = z(0)
    if M == 0:
        r["dL_dsh"] = z(P, 0, 3)
    return r


def mark_visible(means3D, viewmatrix, projmatrix):
    m = _f32(means3D)
    out = np.zeros(m.shape[0], np.uint8)
    lib().orc_mark_visible(m.shape[0], _ptr(m), _ptr(_f32(viewmatrix).reshape(16)),
                           _ptr(_f32(projmatrix).reshape(16)), _ptr(out))
    return out.astype(bool)
