"""Multi-GPU host logic of the rasterizer path: one process per GPU.

Two partitionings exist, both over independent units of the path:

* frames of a synthesis sequence (reference synthesize_fuse.py:53-56: a frame depends only on
  read-only models and that frame's camera/audio) -- no collective at all;
* views of a multi-view training batch -- each view's forward/backward is independent, the
  per-Gaussian gradients add up over views: ONE all-reduce of the packed gradient block that
  `_backward_native` allocates (all ten gradient tensors are sub-blocks of one allocation), plus the
  densification statistics, which are NOT linear in the gradient (reference
  scene/gaussian_model.py:683-685 accumulates the per-view norm of means2D.grad[:, :2]).

Nothing here touches the kernels; `torch.distributed` is plumbing (NCCL on GPUs, gloo in the CPU
tests).
"""
import torch
import torch.distributed as dist


def shard_indices(n, rank, world):
    """Unit indices of `rank`: round-robin, so that consecutive frames (similar cost) spread evenly."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n, world))


def packed_base(grad):
    """The one allocation behind the gradient tensors `_backward_native` returns."""
    return grad._base if grad._base is not None else grad


def allreduce_view_gradients(grads, group=None):
    """Sum the per-Gaussian gradients of this rank's views with everybody else's, in place.

    grads: the tuple `_backward_native` returned (or any tensors that are views of one block).
    One collective for all of them when they share a base, else one per tensor."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return grads
    bases = {}
    for g in grads:
        if g is None or g.numel() == 0:
            continue
        b = packed_base(g)
        bases[b.data_ptr()] = b
    for b in bases.values():
        dist.all_reduce(b, op=dist.ReduceOp.SUM, group=group)
    return grads


def reduce_densification_stats(grad_means2D, radii, group=None):
    """What the trainer needs after a view-sharded batch (reference train_face.py:675-682,
    scene/gaussian_model.py:683-685), summed / maxed over ALL views:
    returns (sum over views of |means2D.grad[:, :2]| [P,1], number of views that saw the Gaussian
    [P,1], max radius over views [P]).  Call BEFORE allreduce_view_gradients (needs per-view grads)."""
    vis = radii > 0
    stats = torch.zeros((radii.shape[0], 2), dtype=torch.float32, device=radii.device)
    stats[:, 0] = torch.where(vis, torch.linalg.vector_norm(grad_means2D[:, :2], dim=-1),
                              torch.zeros((), device=radii.device))
    stats[:, 1] = vis.to(torch.float32)
    rmax = radii.clone()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(rmax, op=dist.ReduceOp.MAX, group=group)
    return stats[:, 0:1], stats[:, 1:2], rmax


def gather_frames(local_frames, n, rank, world, group=None):
    """Frame-sharded synthesis: put the ranks' frames back in sequence order on rank 0.
    local_frames: [len(shard_indices(n, rank, world)), ...] tensor.  Returns [n, ...] on rank 0,
    None elsewhere.  (Host-side reassembly; the render path itself has no collective.)"""
    if world == 1:
        return local_frames
    per = (n + world - 1) // world
    pad = torch.zeros((per,) + tuple(local_frames.shape[1:]), dtype=local_frames.dtype, device=local_frames.device)
    pad[:local_frames.shape[0]] = local_frames
    out = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, out, dst=0, group=group)
    if rank != 0:
        return None
    full = torch.empty((n,) + tuple(local_frames.shape[1:]), dtype=local_frames.dtype, device=local_frames.device)
    for r in range(world):
        idx = shard_indices(n, r, world)
        full[idx] = out[r][:len(idx)]
    return full
