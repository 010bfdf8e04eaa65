"""Measurement of the FP32-sampling variant (build with -DEBT_FP32_SAMPLING): run
with EBT_LIB pointing at that build (or at none, for the FP64 library as the control).  Against the CPU
oracle: how often the accept/reject path of the LM solve differs, how far position and flow angle move, and
whether the stream still optimises at the same events."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
from tests._oracle import CpuBatch, CpuTracker, load_oracle


def ang(a, b):
    d = np.abs(a - b) % (2 * np.pi)
    return np.rad2deg(np.minimum(d, 2 * np.pi - d))


out = {"lib": os.environ.get("EBT_LIB", "default (FP64)")}
ps = synth.make_problems(2000, 640, 480, seed=5000, n_frames=3)
ob, tr = CpuBatch(load_oracle(), 640, 480), Tracker(640, 480)
for f in ps.frames:
    ob.add_frame(f)
    tr.add_frame(f)
o, onab, _, _ = ob.optimize(ps.problems, ps.events)
g, gnab = tr.optimize_batch(ps.problems, ps.events)
tr.close()
same = (o["n_successful"] == g["n_successful"]) & (o["n_unsuccessful"] == g["n_unsuccessful"]) & (o["termination"] == g["termination"])
dpos = np.hypot(o["x"] - g["x"], o["y"] - g["y"])
dang = ang(o["flow_dir"], g["flow_dir"])
out["batch_2000"] = {"lm_path_flip_rate": float(1 - same.mean()), "max_dpos_px": float(dpos.max()),
                     "p99_dpos_px": float(np.percentile(dpos, 99)), "frac_within_0.05px": float((dpos <= 0.05).mean()),
                     "max_dangle_deg": float(dang.max()), "frac_within_0.5deg": float((dang <= 0.5).mean()),
                     "integer_patches_equal": bool(np.array_equal(onab.astype(np.int32), gnab)),
                     "max_dcost": float(np.abs(o["final_cost"] - g["final_cost"]).max())}
S = synth.make_stream(346, 260, 0.10, 2e6, seed=1002, max_corners=500)
cpu, gpu = CpuTracker(load_oracle(), 346, 260, 500), Tracker(346, 260, 500)
gpu.set_stream_kernel(2)
nc, ng = synth.run_stream(cpu, S), synth.run_stream(gpu, S)
rc, rg = cpu.records(), gpu.records()
n = min(len(rc), len(rg))
same_ev = (rc["event_index"][:n] == rg["event_index"][:n]) & (rc["track_id"][:n] == rg["track_id"][:n])
first_div = int(np.argmin(same_ev)) if not same_ev.all() else n
k = first_div
res = {"optimisations_cpu": int(nc), "optimisations_gpu": int(ng),
       "records_before_first_divergence": k, "same_optimisation_sequence": bool(nc == ng and same_ev.all())}
if k:
    d = np.hypot(rc["x"][:k] - rg["x"][:k], rc["y"][:k] - rg["y"][:k])
    res.update(max_dpos_px_before_divergence=float(d.max()),
               max_dangle_deg_before_divergence=float(ang(rc["flow_dir"][:k], rg["flow_dir"][:k]).max()),
               window_sizes_equal_before_divergence=bool(np.array_equal(rc["num_of_events"][:k], rg["num_of_events"][:k])))
# per track: final positions of the tracks both sides still hold
tc, tg = cpu.tracks(), gpu.tracks()
ids = np.intersect1d(tc["track_id"], tg["track_id"])
a = tc[np.isin(tc["track_id"], ids)]
b = tg[np.isin(tg["track_id"], ids)]
res["final_track_max_dpos_px"] = float(np.hypot(a["rect_x"] - b["rect_x"], a["rect_y"] - b["rect_y"]).max())
res["final_tracks_common"] = int(len(ids))
res["final_tracks_cpu"] = int(len(tc))
out["stream_c2_0.1s"] = res
cpu.close()
gpu.close()
print(json.dumps(out))
