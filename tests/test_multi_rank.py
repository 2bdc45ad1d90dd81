"""World-size-2 `gloo` tests (CPU) of the multi-GPU host logic (instag_b200/sharding.py):
view-sharded gradient reduction and frame-sharded reassembly.  SURVEY.md 8(e)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from instag_b200 import sharding as SH


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _packed_grads(P, M, ED, seed):
    """Same layout as rasterizer._backward_native: ten views of one flat block, the parameter
    gradients (SH.PARAMETER_GRADS) first and contiguous."""
    widths = [3, 3, 1, 3, 6, 3, 3 * M, 3, 4, ED]
    starts, o = [0] * len(widths), 0
    for i in (3, 1, 2, 6, 7, 8, 9, 0, 4, 5):
        starts[i] = o
        o += (P * widths[i] + 3) // 4 * 4
    g = torch.Generator().manual_seed(seed)
    flat = torch.randn(o, generator=g)
    return flat, [flat[s:s + P * w].view(P, w) for s, w in zip(starts, widths)]


def _worker(rank, world, port, n_views, n_frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        P, M, ED = 257, 4, 1
        # ---- view-sharded batch: local accumulation over this rank's views, then one all-reduce ----
        mine = SH.shard_indices(n_views, rank, world)
        flat, views = _packed_grads(P, M, ED, seed=1000)  # allocate the block, then zero it
        flat.zero_()
        stats_sum = torch.zeros(P, 1)
        stats_cnt = torch.zeros(P, 1)
        rmax = torch.zeros(P, dtype=torch.int32)
        for v in mine:
            f, vs = _packed_grads(P, M, ED, seed=v)
            radii = torch.randint(0, 5, (P,), generator=torch.Generator().manual_seed(100 + v), dtype=torch.int32)
            flat += f
            # per-view statistics accumulate locally first, one reduction per batch
            vis = radii > 0
            stats_sum[:, 0] += torch.where(vis, torch.linalg.vector_norm(vs[0][:, :2], dim=-1), torch.zeros(()))
            stats_cnt[:, 0] += vis.float()
            rmax = torch.maximum(rmax, radii)
        assert all(SH.packed_base(v).data_ptr() == flat.data_ptr() for v in views)
        SH.allreduce_view_gradients(views)
        # only the parameter gradients: the others keep this rank's local sums
        f2, v2 = _packed_grads(P, M, ED, seed=500 + rank)
        keep = f2.clone()
        SH.allreduce_view_gradients(v2, only_parameters=True)
        dist.all_reduce(stats_sum)
        dist.all_reduce(stats_cnt)
        dist.all_reduce(rmax, op=dist.ReduceOp.MAX)
        # ---- one call of the stats helper on a per-rank view (collective inside) ----
        f0, vs0 = _packed_grads(P, M, ED, seed=rank)
        rad0 = torch.randint(0, 5, (P,), generator=torch.Generator().manual_seed(100 + rank), dtype=torch.int32)
        hs, hc, hr = SH.reduce_densification_stats(vs0[0], rad0)
        # ---- frame-sharded synthesis: every rank "renders" its frames, rank 0 reassembles ----
        fr = SH.shard_indices(n_frames, rank, world)
        local = torch.stack([torch.full((2, 3), float(i)) for i in fr]) if fr else torch.zeros((0, 2, 3))
        full = SH.gather_frames(local, n_frames, rank, world)
        if rank == 0:
            q.put(dict(part=[t.numpy().copy() for t in v2], part_local=keep.numpy().copy(), flat=flat.numpy().copy(), ssum=stats_sum.numpy().copy(), scnt=stats_cnt.numpy().copy(),
                       rmax=rmax.numpy().copy(), full=full.numpy().copy(), hs=hs.numpy().copy(),
                       hc=hc.numpy().copy(), hr=hr.numpy().copy()))
    finally:
        dist.barrier()
        dist.destroy_process_group()


def test_shard_indices_partition():
    for n in (0, 1, 7, 1000):
        for world in (1, 2, 4, 8):
            parts = [SH.shard_indices(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_world2_gradient_reduction_and_frame_gather():
    world, n_views, n_frames = 2, 5, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_views, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process truth
    P, M, ED = 257, 4, 1
    flat = sum(_packed_grads(P, M, ED, seed=v)[0] for v in range(n_views))
    used = 3 * P  # the block's trailing alignment padding (after norm3D, the last view) is not reduced
    end = flat.numel() - ((3 * P + 3) // 4 * 4 - used)
    np.testing.assert_allclose(out["flat"][:end], flat.numpy()[:end], rtol=1e-6, atol=1e-6)
    # only_parameters: parameter gradients summed over the ranks, the rest untouched (rank 0's own)
    per_rank = [_packed_grads(P, M, ED, seed=500 + r)[1] for r in range(world)]
    for i in range(10):
        want = sum(v[i] for v in per_rank) if i in SH.PARAMETER_GRADS else per_rank[0][i]
        np.testing.assert_allclose(out["part"][i], want.numpy(), rtol=1e-6, atol=1e-6, err_msg=str(i))
    ssum = np.zeros((P, 1), np.float32)
    scnt = np.zeros((P, 1), np.float32)
    rmax = np.zeros(P, np.int32)
    for v in range(n_views):
        g2 = _packed_grads(P, M, ED, seed=v)[1][0]
        radii = torch.randint(0, 5, (P,), generator=torch.Generator().manual_seed(100 + v), dtype=torch.int32).numpy()
        vis = radii > 0
        ssum[:, 0] += np.where(vis, np.linalg.norm(g2[:, :2].numpy(), axis=-1), 0)
        scnt[:, 0] += vis
        rmax = np.maximum(rmax, radii)
    np.testing.assert_allclose(out["ssum"], ssum, rtol=1e-5, atol=1e-6)
    np.testing.assert_array_equal(out["scnt"], scnt)
    np.testing.assert_array_equal(out["rmax"], rmax)
    # the helper with its own collectives: two views (seed 0 on rank 0, seed 1 on rank 1)
    hs = np.zeros((P, 1), np.float32)
    hc = np.zeros((P, 1), np.float32)
    hr = np.zeros(P, np.int32)
    for r in range(world):
        g2 = _packed_grads(P, M, ED, seed=r)[1][0]
        radii = torch.randint(0, 5, (P,), generator=torch.Generator().manual_seed(100 + r), dtype=torch.int32).numpy()
        vis = radii > 0
        hs[:, 0] += np.where(vis, np.linalg.norm(g2[:, :2].numpy(), axis=-1), 0)
        hc[:, 0] += vis
        hr = np.maximum(hr, radii)
    np.testing.assert_allclose(out["hs"], hs, rtol=1e-5, atol=1e-6)
    np.testing.assert_array_equal(out["hc"], hc)
    np.testing.assert_array_equal(out["hr"], hr)
    # frames back in sequence order
    assert out["full"].shape == (n_frames, 2, 3)
    for i in range(n_frames):
        assert (out["full"][i] == float(i)).all()
