"""Scratch probe: where does the host time of a c2 step go (per C-ABI call type)?"""
import sys, time, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker, K_TRACK
import bench

WL = sys.argv[1] if len(sys.argv) > 1 else "c3"
w, h, feat, rate, seed = bench.WORKLOADS[WL]
S = synth.make_stream(w, h, 0.30, rate, seed, max_corners=feat)
nF = len(S.frames)
bounds = [(int(np.searchsorted(S.events["t_us"], S.frame_ts[k], "left")),
           int(np.searchsorted(S.events["t_us"], S.frame_ts[k + 1], "left")) if k + 1 < nF else len(S.events))
          for k in range(nF)]
tr = Tracker(w, h, feat, record_capacity=1 << 20)
flows = bench.FlowCache(S)
d_events = torch.from_numpy(S.events.view(np.uint8).reshape(-1).copy()).cuda()
d_frames = [torch.from_numpy(f.copy()).cuda() for f in S.frames]
T = collections.defaultdict(float)
def timed(name, fn, *a):
    t = time.perf_counter(); r = fn(*a); T[name] += time.perf_counter() - t; return r
def step():
    timed("reset", tr.reset)
    for k in range(nF):
        timed("new_frame_begin", tr.new_frame_begin_device, d_frames[k].data_ptr(), S.frame_ts[k], S.corners[k])
        ids, xy = timed("get_uninit", tr.get_uninit)
        if len(ids):
            found, nxt = timed("flow_cache", flows.get, k, ids, xy)
            timed("set_flows", tr.set_flows, ids, found, nxt)
        timed("new_frame_end", tr.new_frame_end)
        lo, hi = bounds[k]
        timed("push_events", tr.push_events_device, d_events.data_ptr() + lo * 16, hi - lo)
for _ in range(3): step()
T.clear(); tr.timing(True)
t0 = time.perf_counter()
N = 5
for _ in range(N): step()
wall = time.perf_counter() - t0
ms, nl = tr.timing_get(K_TRACK)
print(f"wall per step {wall/N*1e3:.2f} ms; stream kernel {ms/N:.2f} ms per step")
for k, v in sorted(T.items(), key=lambda kv: -kv[1]): print(f"  {k:18s} {v/N*1e3:7.3f} ms per step")
