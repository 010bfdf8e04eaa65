"""Scratch probe: do two builds of the library produce bit-identical optimisation records on the same stream?
python tools/records_equal.py <libA.so|default> <libB.so|default> [workload] [seconds]"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ONE = r'''
import sys
sys.path.insert(0, %r)
import numpy as np
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
import bench
w, h, feat, rate, seed = bench.WORKLOADS[sys.argv[2]]
S = synth.make_stream(w, h, float(sys.argv[3]), rate, seed, max_corners=feat)
for mode in (2, 0):
    g = Tracker(w, h, feat, record_capacity=1 << 20)
    g.set_stream_kernel(mode)
    synth.run_stream(g, S)
    np.save(sys.argv[1] + f".{mode}.npy", g.records())
    g.close()
''' % ROOT

if __name__ == "__main__":
    import numpy as np
    a, b = sys.argv[1:3]
    wl = sys.argv[3] if len(sys.argv) > 3 else "c2"
    secs = sys.argv[4] if len(sys.argv) > 4 else "0.1"
    with tempfile.TemporaryDirectory() as tmp:
        outs = []
        for i, lib in enumerate((a, b)):
            env = dict(os.environ)
            if lib != "default":
                env["EBT_LIB"] = os.path.join(ROOT, lib)
            out = os.path.join(tmp, f"r{i}")
            r = subprocess.run([sys.executable, "-c", ONE, out, wl, secs], env=env, capture_output=True, text=True)
            if r.returncode:
                print(r.stderr[-600:])
                sys.exit(1)
            outs.append(out)
        for mode in (2, 0):
            ra, rb = np.load(outs[0] + f".{mode}.npy"), np.load(outs[1] + f".{mode}.npy")
            print(f"stream kernel {mode}: {len(ra)} / {len(rb)} records, bit-identical: {ra.tobytes() == rb.tobytes()}")
