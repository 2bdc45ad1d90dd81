"""Per-unit trace of the forward render kernel (igs_debug_trace): how the (tile, 8x4 block) units of a frame
spread over time, which ones make the tail, and what they did (list length, chunks walked, survivors).

    python profiles/tools/trace_units.py [workload]        # C2, shell, concentrated, ... (instag_b200/synthetic.py)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import util as U
from instag_b200 import synthetic as S, rasterizer as R, _native as N

wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
cfg = dict(S.CONFIGS[wl])
sc = dict(S.cloud_for(cfg)); sc.update(S.make_camera(cfg["W"], cfg["H"])); sc["bg"] = np.asarray(cfg["bg"], np.float32)
t = U.to_torch(sc)
args = U.raw_args_fwd(t)
T = ((cfg["W"] + 15) // 16) * ((cfg["H"] + 15) // 16)
for _ in range(3):
    R._forward_native(*args)
torch.cuda.synchronize()
buf = torch.zeros(4 * 8 * T, dtype=torch.int64, device="cuda")
N.check(N.lib().igs_debug_trace(buf.data_ptr()), "trace on")
out = R._forward_native(*args)
torch.cuda.synchronize()
N.check(N.lib().igs_debug_trace(None), "trace off")
tr = buf.cpu().numpy().reshape(-1, 4)
ran = tr[:, 1] > 0
t0 = tr[ran, 0].min()
start, end = (tr[:, 0] - t0) / 1e3, (tr[:, 1] - t0) / 1e3
total, sm = tr[:, 2] >> 32, tr[:, 2] & 0xffffffff
chunks, surv = tr[:, 3] & 0xffffffff, tr[:, 3] >> 32
dur = end - start
busy = ran & (total > 0)
print(f"{wl}: num_rendered {out[0]}, units {len(tr)}, non-empty {int(busy.sum())}, kernel span {end[ran].max():.1f} us")
print(f"unit us: mean {dur[busy].mean():.1f} p50 {np.percentile(dur[busy], 50):.1f} p90 {np.percentile(dur[busy], 90):.1f} "
      f"p99 {np.percentile(dur[busy], 99):.1f} max {dur[busy].max():.1f}; sum {dur[busy].sum() / 1e3:.2f} ms")
span = end[ran].max()
for frac in (0.25, 0.5, 0.6, 0.7, 0.8, 0.9, 0.95):
    tt = frac * span
    print(f"  running at {tt:6.1f} us: {int(((start <= tt) & (end > tt) & ran).sum())} warps")
print("longest units: dur us | start | list | chunks walked | survivors | us per chunk")
for i in np.argsort(-dur * busy)[:12]:
    print(f"  {dur[i]:7.1f} | {start[i]:6.1f} | {total[i]:6d} | {chunks[i]:5d} | {surv[i]:6d} | {dur[i] / max(chunks[i], 1):.3f}")
