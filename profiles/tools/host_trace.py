"""Where the host time of an end-to-end C2 step goes (perf_counter marks, no syncs inside)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench as B
from instag_b200 import synthetic as S
impl = sys.argv[1] if len(sys.argv) > 1 else "igs"
cfg = dict(S.CONFIGS["C2"])
cloud = S.make_cloud(cfg["P"], sh_degree=cfg["sh_degree"], M=cfg["M"] or None, seed=0)
sc = dict(cloud); sc.update(S.make_camera(cfg["W"], cfg["H"])); sc["bg"] = np.asarray(cfg["bg"], np.float32)
grads = S.upstream_grads(cfg["H"], cfg["W"], extra_dims=1, seed=1)
dev = torch.device("cuda", 0)
arm = B.Arm(impl, sc, grads, dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
pc = time.perf_counter
marks = []
inner = []
if impl == "igs":
    from instag_b200 import rasterizer as R0
    _b = R0.binding()
    class _Timed:  # the binding with rasterize() timed: what is left of Module() is Python
        def __getattr__(self, k):
            return getattr(_b, k)
        def rasterize(self, *a):
            t = pc(); r = _b.rasterize(*a); inner.append(pc() - t); return r
    R0._host_binding = _Timed()
def step():
    t = [pc()]
    k = arm.cur
    main = torch.cuda.current_stream(dev)
    main.wait_event(arm.ready[k]); arm._issue_copy(k ^ 1)
    leaves = arm.leaves[k]
    for v in leaves.values(): v.grad = None
    m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
    t.append(pc())
    color, depth, norm, alpha, radii, extra = arm.Module(arm.rs)(
        means3D=leaves["means3D"], means2D=m2, shs=leaves.get("shs"), colors_precomp=leaves.get("colors_precomp"),
        opacities=leaves["opacities"], scales=leaves["scales"], rotations=leaves["rotations"],
        extra_attrs=leaves.get("extra_attrs"))
    t.append(pc())
    outs = [color, depth, norm, alpha, extra]
    loss = torch.dot(torch.cat([o.reshape(-1) for o in outs]), arm.wflat)
    t.append(pc())
    loss.backward()
    t.append(pc())
    arm.free[k].record(main)
    arm.loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
    arm.cur = k ^ 1
    t.append(pc())
    marks.append(t)
for _ in range(30):
    flush.zero_(); step()
torch.cuda.synchronize(); marks.clear(); inner.clear()
if impl == "igs":
    from instag_b200 import rasterizer as R
    R.binding().host_trace()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(300)]
t0 = pc()
for s, e in ev:
    flush.zero_(); s.record(); step(); e.record()
t1 = pc()
torch.cuda.synchronize()
m = np.array(marks) * 1e6
d = np.diff(m, axis=1)
names = ["prologue", "Module() incl. poll", "cat+dot", "backward()", "epilogue"]
print(impl, "host us per phase (median):", {n: round(float(np.median(d[:, i])), 1) for i, n in enumerate(names)})
print("host loop per step us:", round((t1 - t0) / 300 * 1e6, 1), " between steps (flush, events):",
      round(float(np.median(m[1:, 0] - m[:-1, -1])), 1))
if impl == "igs":
    print("binding.rasterize() us (median):", round(float(np.median(inner)) * 1e6, 1))
    tr = R.binding().host_trace()
    print("inside the node, host us: forward {alloc, args, igs_forward, autograd}:", [round(x, 1) for x in tr[:4]],
          " backward {saved+count, alloc+args, igs_backward, views}:", [round(x, 1) for x in tr[4:8]],
          " forwards {waited, deferred, count read in backward, recorded}:", [int(x) for x in tr[8:]])
print("gpu us per step:", round(float(np.mean([s.elapsed_time(e) for s, e in ev])) * 1e3, 1))
# no flush: can the host keep the GPU fed?
torch.cuda.synchronize(); t0 = pc()
for _ in range(300): step()
torch.cuda.synchronize(); t1 = pc()
print("no flush, wall us per step:", round((t1 - t0) / 300 * 1e6, 1))
