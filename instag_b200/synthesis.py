"""Frame-sharded synthesis driver around the rasterizer.

Mirrors the per-frame body of the reference's `render_set` (synthesize_fuse.py:53-90): the
caller still produces the face and mouth render packages with its own `render_motion*`
functions (which call `diff_gauss`, i.e. this library); what changes is everything after them:

* `compose_frame` -- ONE kernel (`igs_compose_frame`) for the optional dilation of the mouth
  alpha, mouth-over-background, face-over-mouth, clamp, *255 and the uint8 HWC conversion
  (reference: ~12 element-wise launches and three full-frame temporaries per frame);
* `FrameRing` -- the uint8 frame is written into a device ring slot and copied to pinned host
  memory on a side stream; the reference blocks on `.cpu().numpy()` every frame
  (synthesize_fuse.py:76), which serialises rendering and the PCIe copy;
* `RecordedFrames` -- the device work of a frame recorded once into CUDA graphs and replayed, for
  callers whose per-frame inputs live in buffers of fixed address;
* `synthesize_sequence` -- the frames of a sequence are independent (read-only models, one
  camera/audio window per frame), so rank r of world renders frames r, r+world, ...; there is
  no collective on the render path.

No CPU fallback: CUDA tensors only.
"""
import contextlib

import numpy as np
import torch

from . import _native as N
from .sharding import shard_indices


def _f32(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def compose_frame(face_render, face_alpha, mouth_render, mouth_alpha, background=None, dilate=False,
                  ksize=13, out=None, out_image=None):
    """pred = uint8 HWC frame of synthesize_fuse.py:64-76.

    face_render / mouth_render [3,H,W], face_alpha / mouth_alpha [1,H,W] (the `render` and `alpha`
    entries of the two render packages), background = `view.background` ([3,H,W], 0..255) or None.
    out: optional uint8 [H,W,3] CUDA tensor (e.g. a FrameRing slot); out_image: optional float32
    [3,H,W] receiving the un-clamped `image`."""
    if not face_render.is_cuda:
        raise RuntimeError("libigs_raster is CUDA-only: compose_frame needs CUDA tensors (no CPU fallback)")
    H, W = int(face_render.shape[-2]), int(face_render.shape[-1])
    dev = face_render.device
    keep = [_f32(face_render[0:3]), _f32(face_alpha), _f32(mouth_render[0:3]), _f32(mouth_alpha),
            None if background is None else _f32(background.to(dev))]
    if out is None:
        out = torch.empty((H, W, 3), dtype=torch.uint8, device=dev)
    if tuple(out.shape) != (H, W, 3) or out.dtype != torch.uint8 or not out.is_contiguous():
        raise ValueError("out must be a contiguous uint8 [H,W,3] tensor")
    ptr = lambda t: None if t is None else t.data_ptr()
    stream = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        N.check(N.lib().igs_compose_frame(ptr(keep[0]), ptr(keep[1]), ptr(keep[2]), ptr(keep[3]), ptr(keep[4]), H, W,
                                          int(ksize) if dilate else 0, out.data_ptr(), ptr(out_image), stream),
                "compose_frame")
    return out


class FrameRing:
    """`depth` uint8 [H,W,3] slots on the device, mirrored in pinned host memory.  `slot()` hands
    out the next device slot (waiting only if its previous copy has not finished); `commit()`
    queues the device->host copy on a side stream; `collect()` returns finished frames in order."""

    def __init__(self, H, W, depth=8, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.dev = torch.empty((depth, H, W, 3), dtype=torch.uint8, device=self.device)
        self.host = torch.empty((depth, H, W, 3), dtype=torch.uint8).pin_memory()
        self.copy_stream = torch.cuda.Stream(self.device)
        self.rendered = [torch.cuda.Event() for _ in range(depth)]
        self.copied = [torch.cuda.Event() for _ in range(depth)]
        self.depth, self.head, self.tail = depth, 0, 0  # frames committed / collected
        self.tags = [None] * depth

    def slot(self):
        if self.head - self.tail >= self.depth:
            raise RuntimeError("FrameRing is full: collect() before asking for another slot")
        return self.dev[self.head % self.depth]

    def commit(self, tag=None):
        i = self.head % self.depth
        cur = torch.cuda.current_stream(self.device)
        self.rendered[i].record(cur)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.rendered[i])
            self.host[i].copy_(self.dev[i], non_blocking=True)
            self.copied[i].record(self.copy_stream)
        self.tags[i] = tag
        self.head += 1

    def collect(self, block=False):
        """[(tag, uint8 numpy [H,W,3])] of the frames whose copy has finished, oldest first (all
        outstanding frames with block=True).  The arrays are copies: the slot is free afterwards."""
        out = []
        while self.tail < self.head:
            i = self.tail % self.depth
            if block:
                self.copied[i].synchronize()
            elif not self.copied[i].query():
                break
            out.append((self.tags[i], self.host[i].numpy().copy()))
            self.tail += 1
        return out

    def full(self):
        return self.head - self.tail >= self.depth


class PairOverlap:
    """Runs the mouth pass of a frame on a side stream while the face pass runs on the current one.

    The two rasterizations of a frame (synthesize_fuse.py:57-58) do not read each other's results,
    and the front of each (per-Gaussian pass, binning, per-tile sort) leaves most of the GPU idle,
    so the two overlap well.  Only for callbacks that are independent on the device: the side
    stream sees everything queued before the frame, nothing queued by the face pass of this frame.
    The library keeps its scratch per (device, stream), so the two passes share no workspace."""

    def __init__(self, device):
        self.side = torch.cuda.Stream(device)
        self.fork, self.join = torch.cuda.Event(), torch.cuda.Event()

    def __call__(self, render_face, render_mouth, view):
        main = torch.cuda.current_stream()
        self.fork.record(main)
        self.side.wait_event(self.fork)
        # mouth first (measured: 265 us per frame; face first 286; serial 279): the short pass is
        # in its render kernel while the host queues the face pass behind its own instance count
        with torch.cuda.stream(self.side):
            mouth = render_mouth(view)
            self.join.record(self.side)
        face = render_face(view)
        main.wait_event(self.join)
        for t in (mouth["render"], mouth["alpha"]):
            t.record_stream(main)  # allocated on the side stream, read by the compose kernel on this one
        return face, mouth


class DeferredCounts:
    """Takes the wait for `num_rendered` off the host's critical path of a forward-only pipeline.

    A rasterizer call normally returns only after the per-Gaussian kernel has posted the number of
    (tile, Gaussian) instances, because only then is it known that the instance workspaces (sized
    from the previous frame's count + 25 %) were large enough; the host therefore meets the GPU
    once per call and cannot queue ahead.  Inside `with DeferredCounts() as dc:` the no-grad calls
    of this thread (compiled host side only) queue their frame and return at once.  The pipeline
    brackets the calls of a frame with `tok = dc.begin()` ... `dc.end(tok)` and asks
    `dc.cut_short(tok)` when it consumes the frame (after its device->host copy has finished, so
    the counts are there): True means an instance list did not fit and the frame must be rendered
    again -- `with dc.waiting():` restores the ordinary behaviour (which also grows the
    workspaces) for that.  Frames must be asked about in the order they were rendered."""

    def __init__(self):
        from . import rasterizer as R
        self.b = R.binding() if R.host_path() == "cxx" else None
        self.was = False

    def __enter__(self):
        if self.b is not None:
            self.was = self.b.set_deferred_counts(True)
        return self

    def __exit__(self, *exc):
        if self.b is not None:
            self.b.deferred_overflowed(0, self.b.deferred_mark())  # read what is outstanding: the slots come back
            self.b.set_deferred_counts(self.was)
        return False

    def begin(self):
        return [self.b.deferred_mark() if self.b is not None else 0, None]

    def end(self, tok):
        tok[1] = self.b.deferred_mark() if self.b is not None else 0
        return tok

    def cut_short(self, tok):
        return self.b is not None and tok[1] > tok[0] and bool(self.b.deferred_overflowed(tok[0], tok[1]))

    @contextlib.contextmanager
    def waiting(self):
        was = self.b.set_deferred_counts(False) if self.b is not None else False
        try:
            yield
        finally:
            if self.b is not None:
                self.b.set_deferred_counts(was)


class RecordedFrames:
    """The device work of one synthesis frame recorded ONCE per ring slot into CUDA graphs and replayed.

    Nothing in a frame of this library depends on the host (the instance counts are posted, not
    fetched), so a frame -- both rasterizations, the compose kernel into the ring slot -- can be recorded
    through the module API and replayed with one call per frame.  `frame_fn(out)` queues the frame's
    kernels on the current stream and composes into the uint8 [H,W,3] tensor `out`; everything that
    changes from frame to frame (cameras, deformed Gaussians) must reach it through buffers whose
    ADDRESSES stay the same: the caller copies the frame's inputs into them, then calls `render(tag)`.
    The reference cannot be driven this way (its forward sizes buffers on the host from a count it reads
    back in the middle, rasterizer_impl.cu:284).

    The recorded workspaces are those of the warm-up frames with the library's head-room (2x the largest
    instance count seen); `overflowed()` tells after the fact whether a replayed frame outgrew them -- its
    kernels then stood down and the frame is not valid."""

    def __init__(self, frame_fn, ring, warmup=3):
        from . import rasterizer as R
        if R.host_path() != "cxx":
            raise RuntimeError("RecordedFrames needs the compiled host side (IGS_HOST=python does not record)")
        self.ring, self.binding = ring, R.binding()
        dev = ring.device
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        self.graphs, pool = [], None
        with torch.no_grad(), torch.cuda.stream(side):
            for _ in range(max(1, warmup)):  # workspaces, capacity hints, streams of the frame function
                frame_fn(ring.dev[0])
            side.synchronize()
            for i in range(ring.depth):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side, pool=pool):
                    frame_fn(ring.dev[i])
                pool = g.pool() if pool is None else pool
                self.graphs.append(g)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)

    def render(self, tag=None):
        """Replay the frame into the ring's next slot (current stream) and queue its device->host copy."""
        if self.ring.full():
            raise RuntimeError("FrameRing is full: collect() before rendering another frame")
        self.graphs[self.ring.head % self.ring.depth].replay()
        self.ring.commit(tag)

    def overflowed(self):
        return bool(self.binding.captured_overflowed())


def synthesize_sequence(views, render_face, render_mouth, dilate=False, rank=0, world=1, ring_depth=8,
                        sink=None, overlap=False, deferred_counts=True):
    """The loop of `render_set` (synthesize_fuse.py:53-78) for this rank's frames.

    views: the sequence's cameras; render_face(view) / render_mouth(view) -> dicts with "render"
    [3,H,W] and "alpha" [1,H,W] (what render_motion / render_motion_mouth_con return);
    `view.background` is used when the view has one.  Frames idx = rank, rank+world, ... are
    rendered; each is handed to sink(idx, uint8 HWC numpy) in order, or collected and returned as
    (indices, uint8 array [n_local,H,W,3]) when sink is None.  overlap=True runs the two passes of a
    frame on two streams (PairOverlap: only if render_mouth does not read device results of the
    same frame's render_face).  deferred_counts=True: the rasterizer calls do not wait for their
    instance counts (DeferredCounts); a frame whose count turns out to have exceeded the workspaces
    is rendered again before it is handed on."""
    mine = shard_indices(len(views), rank, world)
    ring, frames, got, pair = None, [], [], None
    toks = {}
    dc = DeferredCounts()

    def render_pair(view):
        nonlocal pair
        if overlap:
            if pair is None:
                pair = PairOverlap(torch.device("cuda", torch.cuda.current_device()))
            return pair(render_face, render_mouth, view)
        return render_face(view), render_mouth(view)

    def again(idx):
        # the ordinary, waiting calls (they grow the workspaces), composed and fetched synchronously
        with dc.waiting():
            face, mouth = render_pair(views[idx])
            return compose_frame(face["render"], face["alpha"], mouth["render"], mouth["alpha"],
                                 getattr(views[idx], "background", None), dilate=dilate).cpu().numpy()

    def drain(block):
        for tag, fr in ring.collect(block=block):
            if dc.cut_short(toks.pop(tag)):
                fr = again(tag)
            if sink is not None:
                sink(tag, fr)
            else:
                got.append(tag)
                frames.append(fr)

    with torch.no_grad(), (dc if deferred_counts else contextlib.nullcontext()):
        for idx in mine:
            view = views[idx]
            tok = dc.begin()
            face, mouth = render_pair(view)
            toks[idx] = dc.end(tok)
            H, W = face["render"].shape[-2:]
            if ring is None:
                ring = FrameRing(int(H), int(W), depth=ring_depth, device=face["render"].device)
            if ring.full():
                drain(block=True)
            compose_frame(face["render"], face["alpha"], mouth["render"], mouth["alpha"],
                          getattr(view, "background", None), dilate=dilate, out=ring.slot())
            ring.commit(idx)
            drain(block=False)
        if ring is not None:
            drain(block=True)
    if sink is not None:
        return mine, None
    return got, (np.stack(frames) if frames else np.zeros((0, 0, 0, 3), np.uint8))
