"""GPU, >= 2 devices: the hand-written NVLS all-reduce (igs_allreduce_nvls) of the packed gradient
block against NCCL on the same data, and the gradient-allocator hook that puts the block into
multicast memory.  Skipped on a single-GPU box (the driver's -m gpu run); scratch/nvls_test.py is
the same check under torchrun (2 and 8 B200: bit-identical to NCCL)."""
import os
import socket

import pytest

import util as U  # noqa: F401

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from instag_b200 import sharding as SHD
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        blk = SHD.NvlsGradientBlock(5000, 4, 1, dev)
        g = torch.Generator(device=dev).manual_seed(rank)
        worst = 0.0
        for _ in range(4):
            src = torch.randn(blk.total, device=dev, generator=g)
            want = src.clone()
            dist.all_reduce(want[:blk.param_end])
            with blk:
                from instag_b200 import rasterizer as R
                buf = R._grad_alloc(blk.total, dev)  # what _backward_native does
            buf.copy_(src)
            out = blk.all_reduce()
            torch.cuda.synchronize()
            worst = max(worst, float((out[:blk.param_end] - want[:blk.param_end]).abs().max()))
            assert torch.equal(out[blk.param_end:], src[blk.param_end:])  # means2D / cov3D / norm3D gradients stay local
            assert blk.holds([out[:10]]) and not blk.holds([src[:10]])
        if rank == 0:
            q.put(worst)
    finally:
        dist.barrier()
        dist.destroy_process_group()


def test_nvls_allreduce_matches_nccl():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    worst = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert worst <= 1e-5
