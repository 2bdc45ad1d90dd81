"""igs_allreduce_nvls against ncclAllReduce: result and time, back to back, on the packed gradient block of
C2 (6 MB of parameter gradients) and C5 (252 MB).  Run under torchrun on 2..8 GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 profiles/tools/nvls_test.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from instag_b200 import sharding as SHD

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
for name, (P, M) in {"C2": (100_000, 0), "C5": (1_000_000, 16)}.items():
    blk = SHD.NvlsGradientBlock(P, M, 1, dev)
    n = blk.param_end
    g = torch.Generator(device=dev).manual_seed(rank)
    src = torch.randn(blk.total, device=dev, generator=g)
    want = src.clone()
    dist.all_reduce(want[:n])
    buf = blk._alloc(blk.total, dev)
    buf.copy_(src)
    torch.cuda.synchronize(); dist.barrier()
    out = blk.all_reduce()
    torch.cuda.synchronize()
    same = torch.equal(out[:n], want[:n])
    diff = float((out[:n] - want[:n]).abs().max())

    def timeit(fn, reps=50):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) * 1e3

    t_igs = timeit(lambda: blk.all_reduce())
    t_nccl = timeit(lambda: dist.all_reduce(want[:n]))
    if rank == 0:
        print(f"{name}: {n * 4 / 1e6:.1f} MB over {world} GPUs: igs_allreduce_nvls {t_igs:.1f} us "
              f"({n * 4 / t_igs / 1e3:.0f} GB/s algorithmic), ncclAllReduce {t_nccl:.1f} us; "
              f"bit identical {same}, max |diff| {diff:.2e}", flush=True)
    del blk
dist.barrier()
dist.destroy_process_group()
