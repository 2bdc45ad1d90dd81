"""GPU: igs_l1_ssim_forward_static -- the L1 + D-SSIM forward with its scratch in a file-scope device
array of the library (include/igs_raster.h; no counterpart in the reference) -- against the existing
igs_l1_ssim_forward + igs_l1_ssim_finish pair, through the C ABI, on calls in a row that share the
array.  Tolerances are those of tests/test_gpu_adjacent.py (loss terms 2e-6, maps 1e-4 of their max);
the two paths run the same kernels, so in practice they agree bit for bit.

This file is named to be collected last.
"""
import numpy as np
import pytest

import util as U  # noqa: F401

pytestmark = pytest.mark.gpu

LAMBDA = 0.2


def _pair(shape, seed):
    import torch
    rng = np.random.default_rng(seed)
    x = rng.random(shape, dtype=np.float32)
    y = np.clip(x + 0.1 * rng.standard_normal(shape).astype(np.float32), 0, 1).astype(np.float32)
    return torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()


def _existing(lib, N, x, y):
    import torch
    C_, H, W = (int(v) for v in x.shape)
    maps = torch.empty((3, C_, H, W), dtype=torch.float32, device=x.device)
    part = torch.empty((lib.igs_l1_ssim_partials(C_, H, W), 2), dtype=torch.float32, device=x.device)
    out3 = torch.empty((3,), dtype=torch.float32, device=x.device)
    stream = torch.cuda.current_stream().cuda_stream
    N.check(lib.igs_l1_ssim_forward(x.data_ptr(), y.data_ptr(), C_, H, W, maps[0].data_ptr(), maps[1].data_ptr(),
                                    maps[2].data_ptr(), part.data_ptr(), stream), "l1_ssim forward")
    N.check(lib.igs_l1_ssim_finish(part.data_ptr(), C_, H, W, LAMBDA, out3.data_ptr(), stream), "l1_ssim finish")
    torch.cuda.synchronize()
    return maps.cpu().numpy(), out3.cpu().numpy()


def test_static_scratch_forward_matches_the_workspace_forward_on_calls_in_a_row():
    import torch
    from instag_b200 import _native as N
    lib = N.lib()
    stream = torch.cuda.current_stream().cuda_stream
    # big, small, big again with other data, one tile: each call finds the last call's sums in the array
    cases = [((3, 512, 512), 11), ((3, 75, 50), 12), ((3, 512, 512), 13), ((1, 16, 16), 14), ((3, 75, 50), 12)]
    inputs = [_pair(shape, seed) for shape, seed in cases]
    want = [_existing(lib, N, x, y) for x, y in inputs]
    outs = [(torch.full((3,) + tuple(x.shape), float("nan"), device="cuda"), torch.full((3,), float("nan"), device="cuda"))
            for x, _ in inputs]
    torch.cuda.synchronize()
    free_before = None
    for i, ((x, y), (maps, out3)) in enumerate(zip(inputs, outs)):
        C_, H, W = (int(v) for v in x.shape)
        if i == 1:   # the first call has loaded what the driver loads lazily; from here nothing may be taken
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        N.check(lib.igs_l1_ssim_forward_static(x.data_ptr(), y.data_ptr(), C_, H, W, LAMBDA, maps[0].data_ptr(),
                                               maps[1].data_ptr(), maps[2].data_ptr(), out3.data_ptr(), stream),
                "l1_ssim forward_static")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "the static-scratch entry point took device memory"
    for (shape, _), (wm, w3), (maps, out3) in zip(cases, want, outs):
        gm, g3 = maps.cpu().numpy(), out3.cpu().numpy()
        assert np.isfinite(g3).all() and np.isfinite(gm).all(), shape
        assert np.abs(g3 - w3).max() <= 2e-6, (shape, g3, w3)
        for k in range(3):
            assert np.abs(gm[k] - wm[k]).max() <= 1e-4 * np.abs(wm[k]).max(), (shape, k)


def test_static_scratch_forward_refuses_frames_beyond_its_array():
    import torch
    from instag_b200 import _native as N
    lib = N.lib()
    assert lib.igs_l1_ssim_partials(3, 4096, 4096) <= 65536 < lib.igs_l1_ssim_partials(3, 4800, 4800)
    t = torch.zeros((16,), dtype=torch.float32, device="cuda")   # never dereferenced: refused before any launch
    p = t.data_ptr()
    rc = lib.igs_l1_ssim_forward_static(p, p, 3, 4800, 4800, LAMBDA, p, p, p, p, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.igs_last_error().decode()
    assert "static scratch" in msg and "65536" in msg, msg
    with pytest.raises(RuntimeError):
        N.check(rc, "l1_ssim forward_static")
    torch.cuda.synchronize()
    assert float(t.abs().max()) == 0.0
