"""Scratch probe: where a feature's leader spends its cycles in the stream kernels.  Needs a library built with
-DEBT_WALK_CLOCKS (EBT_LIB=build/libebt_walk.so); prints, per workload and stream-kernel mode, cycles per
optimisation inside Optimizer::optimize / lm_solve / window-size update / mask-row rebuilds and what is left
(candidate walk + state traffic), plus rebuilds and mask segments per optimisation."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
import bench

out = []
for WL in sys.argv[1:] or ["c2", "c3"]:
    w, h, feat, rate, seed = bench.WORKLOADS[WL]
    S = synth.make_stream(w, h, 0.30, rate, seed, max_corners=feat)
    nF = len(S.frames)
    t = S.events["t_us"]
    bounds = [(int(np.searchsorted(t, S.frame_ts[k], "left")),
               int(np.searchsorted(t, S.frame_ts[k + 1], "left")) if k + 1 < nF else len(S.events)) for k in range(nF)]
    d_events = torch.from_numpy(S.events.view(np.uint8).reshape(-1).copy()).cuda()
    d_frames = [torch.from_numpy(f.copy()).cuda() for f in S.frames]
    flows = bench.FlowCache(S)
    for mode in (0, 2):
        tr = Tracker(w, h, feat, record_capacity=1 << 20)
        tr.set_stream_kernel(mode)

        def step():
            tr.reset()
            n = 0
            for k in range(nF):
                tr.new_frame_begin_device(d_frames[k].data_ptr(), S.frame_ts[k], S.corners[k])
                ids, xy = tr.get_uninit()
                if len(ids):
                    found, nxt = flows.get(k, ids, xy)
                    tr.set_flows(ids, found, nxt)
                tr.new_frame_end()
                lo, hi = bounds[k]
                n += tr.push_events_device(d_events.data_ptr() + lo * 16, hi - lo)
            return n

        for _ in range(2):
            step()
        tr.phase_clocks()  # reset
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = step()
        e1.record()
        torch.cuda.synchronize()
        c = [int(v) for v in tr.phase_clocks()]
        nopt = max(c[5], 1)
        r = dict(workload=WL, stream_kernel=mode, optimisations=n, counted=c[5], step_ms=e0.elapsed_time(e1),
                 cycles_per_opt=dict(total=c[0] / nopt, optimize=c[1] / nopt, lm_solve=c[2] / nopt,
                                     num_events=c[6] / nopt, rebuild=c[3] / nopt,
                                     walk_and_rest=(c[0] - c[1] - c[3] - c[6]) / nopt),
                 rebuilds_per_opt=c[4] / nopt, segments_per_opt=c[7] / nopt,
                 cycles_per_rebuild=c[3] / max(c[4], 1))
        if mode == 2:  # the multi-stream kernel reuses slots 3 / 4: mask wait + expansion, chunk phase
            r["cycles_per_opt"]["expand"] = c[3] / nopt
            r["cycles_per_opt"]["chunks_without_optimise"] = (c[4] - c[1] - c[6]) / nopt
            r["cycles_per_opt"]["feature_entry_exit"] = (c[0] - c[3] - c[4]) / nopt
            r["cycles_per_opt"]["chunk_record_wait_and_first_test"] = c[7] / nopt
            del r["segments_per_opt"]
            del r["cycles_per_opt"]["rebuild"], r["rebuilds_per_opt"], r["cycles_per_rebuild"]
        print(json.dumps(r))
        out.append(r)
        tr.close()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/walk_profile.json", "w"), indent=1)
