"""Golden result of BASELINE.json configs[4] (C5) on the CPU side: the C5 synthetic stream (tools/c5_case.py)
tracked by the CPU oracle — FeatureDetector::newImage's OpenCV calls answered by the numpy restatement
(oracle/frontend_oracle.py), the tracking by oracle/libebt_oracle.so — and its tracks fed to the reference's
UNMODIFIED visual-odometry back end (oracle/_ref/vo_driver, mode "tracks").  Authoring container only
(needs oracle/_ref/vo_driver, built by `make -C oracle vo` from /root/reference).

    python tests/golden/make_c5_golden.py   ->  tests/golden/c5_vo_oracle.json

tests/test_c5_vo.py runs the same stream through the GPU tracker behind the drop-in headers + the same
back end (vo_driver mode "gpu") and compares keyframe count, ATE and camera centres with this file.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from tests._oracle import CpuTracker, load_oracle  # noqa: E402
from tests.test_gpu_frontend import CpuFrontEnd  # noqa: E402
from tools import c5_case  # noqa: E402

DURATION = 3.0


def main():
    t0 = time.time()
    S = c5_case.make_stream(DURATION)
    cpu = CpuFrontEnd(CpuTracker(load_oracle(), c5_case.W, c5_case.H, c5_case.MAX_PATCHES))
    per = c5_case.tracks_of(cpu, S)
    n_opt = len(cpu.records())
    os.makedirs("/tmp/c5", exist_ok=True)
    c5_case.write_tracks_input("/tmp/c5/oracle.bin", S, per)
    r = c5_case.run_vo("tracks", "/tmp/c5/oracle.bin", "/tmp/c5/oracle.txt")
    out = {k: v for k, v in r.items() if k != "pose"}
    out["pose"] = r["pose"].tolist()
    out["duration_s"] = DURATION
    out["frames"] = len(S.frames)
    out["events"] = int(len(S.events))
    out["tracker_optimisations"] = int(n_opt)
    out["tracks_per_image"] = [int(len(i)) for i, _ in per]
    json.dump(out, open(os.path.join(HERE, "c5_vo_oracle.json"), "w"), indent=1)
    print({k: v for k, v in out.items() if k not in ("pose", "tracks_per_image")}, f"{time.time() - t0:.0f} s")


if __name__ == "__main__":
    main()
