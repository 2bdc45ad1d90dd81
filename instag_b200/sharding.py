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


# positions of the trainable-parameter gradients in the tuple `_backward_native` returns
# (means2D, colours, opacity, means3D, cov3D, norm3D, sh, scales, rotations, extra)
PARAMETER_GRADS = (1, 2, 3, 6, 7, 8, 9)


def allreduce_view_gradients(grads, group=None, only_parameters=False):
    """Sum the per-Gaussian gradients of this rank's views with everybody else's, in place.

    grads: the tuple `_backward_native` returned (or any tensors that are views of one block).
    One collective for all of them when they share a base, else one per tensor.
    only_parameters: reduce just the gradients an optimizer consumes (PARAMETER_GRADS: 15 + 3M of the
    27 + 3M packed floats per Gaussian).  The means2D gradient only feeds the per-view densification
    statistics (reduce_densification_stats, which needs it un-summed), and the cov3D / norm3D
    gradients belong to inputs InsTaG never passes (gaussian_renderer/__init__.py:237 cov3D_precomp =
    None); `_backward_native` lays the parameter gradients out as one contiguous prefix for this."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return grads
    sel = [g for i, g in enumerate(grads) if (not only_parameters or i in PARAMETER_GRADS)]
    sel = [g for g in sel if g is not None and g.numel() != 0]
    bases = {}
    for g in sel:
        b = packed_base(g)
        bases.setdefault(b.data_ptr(), (b, []))[1].append(g)
    for b, members in bases.values():
        if b.dim() == 1 and b.is_contiguous():
            # the smallest range of the block that covers the selected tensors (whole block if all are)
            lo = min(g.storage_offset() for g in members) - b.storage_offset()
            hi = max(g.storage_offset() + g.numel() for g in members) - b.storage_offset()
            dist.all_reduce(b[lo:hi], op=dist.ReduceOp.SUM, group=group)
        else:
            dist.all_reduce(b, op=dist.ReduceOp.SUM, group=group)
    return grads


class NvlsGradientBlock:
    """Home of the packed gradient block in NVLink multicast (symmetric) memory, and its all-reduce
    as ONE hand-written kernel (`igs_allreduce_nvls`: multimem.ld_reduce + multimem.st with flag
    barriers, csrc/igs_collective.cu) instead of ncclAllReduce.

        block = NvlsGradientBlock(P, M, ED, device)        # collective: every rank, once
        with block:                                         # backward writes its gradients into it
            loss.backward()   (or _backward_native(...))
        block.all_reduce()                                  # parameter-gradient prefix, in place

    Two buffers alternate, so the gradients of one step stay valid while the next step runs.
    Raises RuntimeError when symmetric memory / multicast is unavailable (callers fall back to
    `allreduce_view_gradients`, i.e. NCCL)."""

    def __init__(self, P, M, ED, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        from . import rasterizer as R
        from . import _native as N
        self.N = N
        self.R = R
        self.group = dist.group.WORLD if group is None else group
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        _, _, total, param_end = R.gradient_layout(P, M, ED)
        self.total, self.param_end = total, param_end
        self.device = torch.device(device)
        self.bufs, self.hdls = [], []
        for _ in range(2):
            t = symm_mem.empty(total, dtype=torch.float32, device=self.device)
            h = symm_mem.rendezvous(t, self.group.group_name)
            if not getattr(h, "multicast_ptr", 0):
                raise RuntimeError("NVLink multicast (NVLS) is not available for this group")
            if h.signal_pad_size < 1024 + 4 * self.world:
                raise RuntimeError("symmetric-memory signal pad too small")
            self.bufs.append(t)
            self.hdls.append(h)
        self.cur = 0
        self._prev = None

    def _alloc(self, n_floats, device):
        if n_floats != self.total or torch.device(device) != self.device:
            raise RuntimeError("NvlsGradientBlock was sized for another (P, M, ED) or device")
        self.cur ^= 1
        return self.bufs[self.cur]

    def __enter__(self):
        self._prev = self.R.set_gradient_allocator(self._alloc)
        return self

    def __exit__(self, *exc):
        self.R.set_gradient_allocator(self._prev)
        return False

    def holds(self, tensors):
        """True if every given tensor lives inside the buffer the last backward wrote (e.g. the .grad
        of the leaves after loss.backward()), i.e. one all_reduce() of the block covers them all."""
        buf = self.bufs[self.cur]
        lo, hi = buf.data_ptr(), buf.data_ptr() + 4 * (self.param_end)
        return all(t is not None and lo <= t.data_ptr() and t.data_ptr() + 4 * t.numel() <= hi for t in tensors)

    def all_reduce(self, only_parameters=True):
        """Sum the block written by the last backward over the ranks, in place (current stream)."""
        h = self.hdls[self.cur]
        n = self.param_end if only_parameters else self.total
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            self.N.check(self.N.lib().igs_allreduce_nvls(h.multicast_ptr + getattr(h, "offset", 0), h.signal_pad_ptrs_dev,
                                                         int(h.signal_pad_size), self.rank, self.world, n, stream),
                         "allreduce_nvls")
        return self.bufs[self.cur]


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
