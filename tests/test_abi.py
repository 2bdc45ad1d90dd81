"""The C-ABI library loads without a GPU and exports every symbol that
include/igs_raster.h declares; struct layouts of the ctypes mirror match the
header as gcc sees it; argument validation fails loudly (no compute calls)."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import util as U  # noqa: F401
from instag_b200 import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "igs_raster.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(igs_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    names = _declared()
    assert len(names) >= 15
    lib = N.lib()
    for n in names:
        assert n in N.SYMBOLS, f"{n} declared in the header but not bound in _native.SYMBOLS"
        getattr(lib, n)
    assert set(N.SYMBOLS) == set(names)
    assert b"sm_100a" in lib.igs_version()


def test_struct_layout_matches_header():
    prog = r'''
#include <stdio.h>
#include <stddef.h>
#include "igs_raster.h"
int main(void){
  printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(IgsFwdArgs), offsetof(IgsFwdArgs,bg), offsetof(IgsFwdArgs,capacity),
         offsetof(IgsFwdArgs,num_rendered_pinned), sizeof(IgsBwdArgs), offsetof(IgsBwdArgs,num_rendered),
         offsetof(IgsBwdArgs,dL_dextra_attrs));
  return 0; }
'''
    with tempfile.TemporaryDirectory() as d:
        src, exe = os.path.join(d, "t.c"), os.path.join(d, "t")
        open(src, "w").write(prog)
        subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), src, "-o", exe])
        got = [int(x) for x in subprocess.check_output([exe]).split()]
    F, B = N.IgsFwdArgs, N.IgsBwdArgs
    want = [C.sizeof(F), F.bg.offset, F.capacity.offset, F.num_rendered_pinned.offset, C.sizeof(B),
            B.num_rendered.offset, B.dL_dextra_attrs.offset]
    assert got == want


def test_workspace_sizes_are_monotonic_and_aligned():
    lib = N.lib()
    prev = 0
    for P in (0, 1, 255, 256, 257, 100_000, 1_000_000):
        b = lib.igs_geom_bytes(P)
        assert b % 256 == 0 and b >= prev and b >= 72 * P
        prev = b
    for cap in (0, 1, 4096, 1 << 20):
        assert lib.igs_bin_bytes(cap) >= 12 * cap and lib.igs_sort_bytes(cap) >= 8 * cap
        lay = N.bin_layout(max(cap, 1))
        assert lay["point_list"] == 0 and lay["keys"] % 256 == 0
    assert lib.igs_img_bytes(512, 512) >= 4 * 512 * 512 + 8 * 1024
    assert lib.igs_accum_bytes(1000, 1) >= 64 * 1000
    assert lib.igs_accum_bytes(1000, 3) > lib.igs_accum_bytes(1000, 1)
    g = N.geom_layout(1000)
    assert g["splat"] == 0 and g["tiles_touched"] >= 64 * 1000 and g["point_offsets"] > g["tiles_touched"]


def test_invalid_arguments_fail_loudly():
    lib = N.lib()
    n = C.c_int64(0)
    assert lib.igs_forward(None, None, C.byref(n)) == N.IGS_ERR_INVALID
    assert "null" in N.last_error()
    a = N.IgsFwdArgs()
    a.P, a.W, a.H, a.ED = 10, 64, 64, 99
    assert lib.igs_forward(C.byref(a), None, C.byref(n)) == N.IGS_ERR_INVALID
    assert "ED" in N.last_error()
    a.ED = 0
    assert lib.igs_forward(C.byref(a), None, C.byref(n)) == N.IGS_ERR_INVALID  # neither shs nor colors
    assert "exactly one" in N.last_error()
    assert lib.igs_backward(None, None) == N.IGS_ERR_INVALID
    assert lib.igs_mark_visible(5, None, None, None, None, None) == N.IGS_ERR_INVALID


def test_missing_library_raises(monkeypatch):
    import pytest
    monkeypatch.setattr(N, "_lib", None)
    monkeypatch.setattr(N, "LIB_PATH", "/nonexistent/libigs_raster.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        N.lib()


def test_compiled_host_side_loads_and_rejects_cpu_tensors():
    """csrc/igs_torch.cpp (the autograd node in C++) imports without a GPU, links the same library
    and, like the Python host side, refuses anything that is not on a CUDA device."""
    import pytest
    import torch
    from instag_b200 import rasterizer as R
    b = R.binding()
    assert b.library_version() == N.lib().igs_version().decode()
    e = torch.Tensor([])
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="CUDA-only"):
        b.rasterize(m, m, e, m, torch.zeros(4, 1), m, torch.zeros(4, 4), e, e, e, torch.zeros(3), torch.eye(4),
                    torch.eye(4), torch.zeros(3), 0.5, 0.5, 1.0, 16, 16, 0, False, True)
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        b.rasterize(torch.zeros(4, 2), m, e, m, torch.zeros(4, 1), m, torch.zeros(4, 4), e, e, e, torch.zeros(3),
                    torch.eye(4), torch.eye(4), torch.zeros(3), 0.5, 0.5, 1.0, 16, 16, 0, False, True)


def test_host_path_switch(monkeypatch):
    from instag_b200 import rasterizer as R
    monkeypatch.delenv("IGS_HOST", raising=False)
    assert R.host_path() == "cxx"
    monkeypatch.setenv("IGS_HOST", "python")
    assert R.host_path() == "python"


def test_deferred_counts_context_without_a_gpu(monkeypatch):
    """synthesis.DeferredCounts is inert with the Python host side, and with the compiled one it
    only flips this thread's switch (restored on exit); no frame is outstanding, so nothing is cut short."""
    from instag_b200 import rasterizer as R
    from instag_b200 import synthesis as SY
    monkeypatch.setenv("IGS_HOST", "python")
    dc = SY.DeferredCounts()
    with dc:
        tok = dc.end(dc.begin())
        assert dc.b is None and not dc.cut_short(tok)
    monkeypatch.setenv("IGS_HOST", "cxx")
    b = R.binding()
    assert b.set_deferred_counts(False) is False
    with SY.DeferredCounts() as dc:
        assert b.set_deferred_counts(True) is True  # the context switched it on
        tok = dc.end(dc.begin())
        assert tok[0] == tok[1] == b.deferred_mark() and not dc.cut_short(tok)
        with dc.waiting():
            assert b.set_deferred_counts(False) is False
        assert b.set_deferred_counts(True) is True  # waiting() put it back
    assert b.set_deferred_counts(False) is False  # and the context restored the caller's setting
