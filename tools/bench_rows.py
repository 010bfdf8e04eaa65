"""Measurement of the rows either side of the hot path:
per-frame re-arm kernels, the event-text parser and the VO hand-off, each with
CUDA-event kernel time (ebt_timing_*), an end-to-end time through the C-ABI with
host buffers, and its CPU counterpart timed on this box's host.

Prints one JSON line per row.  `python tools/bench_rows.py [--no-cpu]`."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from event_based_odomety_b200 import synth  # noqa: E402
from event_based_odomety_b200.cabi import (Camera, K_CORNERS, K_LK, K_PARSE, K_PYRAMID,  # noqa: E402
                                           Tracker)

NO_CPU = "--no-cpu" in sys.argv
HBM_GBS = 6453.7
try:
    HBM_GBS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def wall(fn, reps):
    fn()
    t = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t) / reps


def frontend_row(name, w, h, seed, n_pts):
    S = synth.make_stream(w, h, 0.12, 500e3, seed=seed, max_corners=10)
    f0, f1 = S.frames[0], S.frames[1]
    tr = Tracker(w, h, 2000)
    corners = tr.detect_features(f0)
    rng = np.random.default_rng(1)
    pts = np.column_stack([rng.uniform(15, w - 15, n_pts), rng.uniform(15, h - 15, n_pts)]).astype(np.float32)
    tr.timing(True)
    reps = 50
    e2e_det = wall(lambda: tr.detect_features(f0), reps)
    e2e_lk = wall(lambda: tr.lk_flow(f0, f1, pts), reps)
    det_ms, det_n = tr.timing_get(K_CORNERS)
    pyr_ms, pyr_n = tr.timing_get(K_PYRAMID)
    lk_ms, lk_n = tr.timing_get(K_LK)
    row = {"row": "n1 per-frame re-arm", "workload": f"{name}: {w}x{h}, {len(corners)} corners, LK on {n_pts} points",
           "corner_detector_kernels_us": 1e3 * det_ms / det_n, "corner_detector_e2e_us": 1e6 * e2e_det,
           "pyramid_scharr_kernels_us": 1e3 * pyr_ms / (lk_n), "lk_kernel_us": 1e3 * lk_ms / lk_n,
           "lk_e2e_us": 1e6 * e2e_lk,
           # algorithmic bytes of the detector: 1 B/px read + 4 B/px response written (its 3 product planes
           # and the response re-read are L2 traffic at these sizes)
           "corner_detector_algorithmic_GBps": 5.0 * w * h / (1e-3 * det_ms / det_n) / 1e9}
    if not NO_CPU:
        try:
            import cv2
            mask = np.zeros((h, w), np.uint8)
            cv2.rectangle(mask, (12, 12, w - 24, h - 24), 1, cv2.FILLED)
            row["cpu_cv2_goodFeaturesToTrack_us"] = 1e6 * wall(
                lambda: cv2.goodFeaturesToTrack(f0, w * h // 625, 0.01, 10, mask=mask, blockSize=3,
                                                useHarrisDetector=True, k=0.04), 20)
            # the reference calls calcOpticalFlowPyrLK once per point (flow_estimator.cpp:84-103)
            def per_point():
                for p in pts[:50]:
                    cv2.calcOpticalFlowPyrLK(f0, f1, p.reshape(1, 1, 2), None, winSize=(21, 21), maxLevel=3)
            row["cpu_cv2_lk_per_point_call_us"] = 1e6 * wall(per_point, 3) / 50
            row["cpu_threads"] = cv2.getNumThreads()
        except ImportError:
            row["cpu"] = "cv2 not importable on this box"
    tr.close()
    return row


def parser_row(n_lines):
    from oracle import ingest_oracle as io
    rng = np.random.default_rng(0)
    ev = np.zeros(n_lines, synth.EVENT) if hasattr(synth, "EVENT") else None
    from event_based_odomety_b200.records import EVENT
    ev = np.zeros(n_lines, EVENT)
    ev["t_us"] = np.sort(rng.integers(0, 60_000_000, n_lines))
    ev["x"] = rng.integers(0, 640, n_lines)
    ev["y"] = rng.integers(0, 480, n_lines)
    ev["polarity"] = rng.choice([-1, 1], n_lines)
    text = io.format_events_text(ev)
    tr = Tracker(640, 480)
    tr.timing(True)
    reps = 10
    e2e = wall(lambda: tr.parse_events_text(text, cap=n_lines), reps)
    ms, n = tr.timing_get(K_PARSE)
    k_s = 1e-3 * ms / n
    # algorithmic bytes: the text read once by the count pass and once by the parse pass + 16 B/event written
    alg = 2 * len(text) + 16 * n_lines
    row = {"row": "n2 event-text parser", "workload": f"{n_lines} lines, {len(text)} bytes of 't_sec x y p' text",
           "kernels_ms": 1e3 * k_s, "kernels_Mlines_per_s": n_lines / k_s / 1e6,
           "roofline": {"bound": "hbm", "achieved": alg / k_s / 1e9, "peak": HBM_GBS, "unit": "GB/s",
                        "frac": alg / k_s / 1e9 / HBM_GBS},
           "e2e_ms_host_text_to_host_events": 1e3 * e2e, "e2e_Mlines_per_s": n_lines / e2e / 1e6,
           "h2d_bytes": len(text), "d2h_bytes": 16 * n_lines}
    if not NO_CPU:
        from tests._oracle import cpu_parse_text, load_ingest
        dll = load_ingest()
        sample = text[:text.index(b"\n", len(text) // 10) + 1]
        t = time.perf_counter()
        got = cpu_parse_text(dll, sample)
        dt = time.perf_counter() - t
        row["cpu_baseline"] = {"value": len(got) / dt / 1e6, "unit": "Mlines/s", "cores": 1, "kind": "port",
                               "sample": f"first {len(got)} lines, std::getline/stod/stoi as the reference reads them"}
    tr.close()
    return row


def handoff_row(n):
    cam = Camera(fx=199.092366542, fy=198.82882047, cx=132.192071378, cy=110.712660011,
                 k1=-0.368436311798, k2=0.150947243557, k3=0.0, p1=-0.000296130534385, p2=-0.000759431726241)
    rng = np.random.default_rng(2)
    pts = rng.uniform([0, 0], [240, 180], (n, 2))
    tr = Tracker(240, 180)
    e2e = wall(lambda: tr.unproject(cam, pts), 20)
    row = {"row": "n4 VO hand-off", "workload": f"CameraModel::unproject of {n} landmarks",
           "e2e_us": 1e6 * e2e, "e2e_Mpoints_per_s": n / e2e / 1e6}
    if not NO_CPU:
        from tests._oracle import cpu_unproject, load_oracle
        ora = load_oracle()
        cpu = wall(lambda: cpu_unproject(ora, cam, pts), 5)
        row["cpu_port_Mpoints_per_s"] = n / cpu / 1e6
    tr.close()
    return row


if __name__ == "__main__":
    for args in (("c1", 240, 180, 1001, 100), ("c2", 346, 260, 1002, 500), ("c3", 640, 480, 1003, 2000)):
        print(json.dumps(frontend_row(*args)), flush=True)
    print(json.dumps(parser_row(2_000_000)), flush=True)
    print(json.dumps(handoff_row(2000)), flush=True)
    print(json.dumps(handoff_row(1_000_000)), flush=True)
