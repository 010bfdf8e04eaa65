"""Scratch probe: batched-kernel parity summary and raw throughput on the box."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker, K_BATCH
from event_based_odomety_b200.records import SOLUTION

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ps = synth.make_problems(n, 346, 260, seed=4242, n_frames=2)
tr = Tracker(346, 260)
for f in ps.frames: tr.add_frame(f)
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 16
prob = np.tile(ps.problems, reps)
dp = torch.from_numpy(prob.view(np.uint8).copy()).cuda()
de = torch.from_numpy(ps.events.view(np.uint8).copy()).cuda()
ds = torch.zeros(len(prob) * SOLUTION.itemsize, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
tr.timing(True)
for it in range(4):
    tr.optimize_batch_device(dp.data_ptr(), len(prob), de.data_ptr(), ds.data_ptr())
    tr.sync()
    ms, nl = tr.timing_get(K_BATCH)
    print(f"iter {it}: cumulative {ms:.3f} ms over {nl} launches")
ms, nl = tr.timing_get(K_BATCH)
sol = ds.cpu().numpy().view(SOLUTION)
ev = sol["n_passes"]
njac, ncost = ev // 100, ev % 100
ns, nu = sol["n_successful"].astype(float), sol["n_unsuccessful"].astype(float)
flops = 625.0 * (400 * (1 + ns) + 230 * (ns + nu))  # SURVEY 8d (reference's evaluation count)
print(f"problems/launch {len(prob)}  avg ms/launch {ms/nl:.3f}  -> {len(prob)/(ms/nl)*1e3:.0f} opt/s")
print(f"mean full evals {njac.mean():.2f} cost-only evals {ncost.mean():.2f}; algorithmic TFLOP/s {flops.sum()/(ms/nl*1e-3)/1e12:.3f}")
pc = tr.phase_clocks()
if pc[5]:
    names = ["request", "pixel pass", "wait partials", "totals+final", "LM rest"]
    print("leader phase clocks per evaluation: " + ", ".join(f"{n} {pc[i]/pc[5]:.0f}" for i, n in enumerate(names)))
