"""Scratch probe: the same stream step (device-resident inputs, as bench.py's `value`) on several builds of the
library.  python tools/variant_bench.py <workload> <lib.so> [<lib.so> ...]; the stream is generated once."""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ONE = r'''
import sys, os, json, time
sys.path.insert(0, %r)
import numpy as np, torch
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
import bench
S = synth.load_stream(sys.argv[1]); feat = int(sys.argv[2]); mode = int(sys.argv[3])
nF = len(S.frames); t = S.events["t_us"]
bounds = [(int(np.searchsorted(t, S.frame_ts[k], "left")), int(np.searchsorted(t, S.frame_ts[k + 1], "left")) if k + 1 < nF else len(S.events)) for k in range(nF)]
d_events = torch.from_numpy(S.events.view(np.uint8).reshape(-1).copy()).cuda()
d_frames = [torch.from_numpy(f.copy()).cuda() for f in S.frames]
flows = bench.FlowCache(S)
tr = Tracker(S.width, S.height, feat, record_capacity=1 << 20); tr.set_stream_kernel(mode)
def step():
    tr.reset(); n = 0
    for k in range(nF):
        bench.frame_phase(tr, S, flows, lambda kk: tr.new_frame_begin_device(d_frames[kk].data_ptr(), S.frame_ts[kk], S.corners[kk]), k)
        lo, hi = bounds[k]; n += tr.push_events_device(d_events.data_ptr() + lo * 16, hi - lo)
    return n
for _ in range(2): step()
tr.timing(True); torch.cuda.synchronize(); t0 = time.perf_counter(); n = 0
for _ in range(3): n += step()
torch.cuda.synchronize(); dt = time.perf_counter() - t0
ms, _ = tr.timing_get(0)
print(json.dumps({"lib": os.environ.get("EBT_LIB", "default"), "mode": mode, "opt_per_s": n / dt, "ms_per_step": dt / 3 * 1e3, "kernel_ms_per_step": ms / 3}))
''' % ROOT

if __name__ == "__main__":
    import bench
    from event_based_odomety_b200 import synth
    wl = sys.argv[1]
    w, h, feat, rate, seed = bench.WORKLOADS[wl]
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "s.npz")
        synth.save_stream(synth.make_stream(w, h, 0.30, rate, seed, max_corners=feat), path)
        for lib in sys.argv[2:]:
            mode = 2
            if ":" in lib:
                lib, mode = lib.split(":")
            env = dict(os.environ)
            if lib != "default":
                env["EBT_LIB"] = os.path.join(ROOT, lib)
            r = subprocess.run([sys.executable, "-c", ONE, path, str(feat), str(mode)], env=env, capture_output=True, text=True)
            print(r.stdout.strip() or r.stderr.strip()[-400:])
