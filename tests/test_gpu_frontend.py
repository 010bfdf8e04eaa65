"""GPU parity of the per-frame re-arm: the corner detector and
the pyramidal LK tracker of csrc/ebt_frontend.cu through the C-ABI
(ebt_detect_features, ebt_lk_flow, ebt_new_image) against

* the cv2 golden vectors (tests/golden/frontend_cv2.npz): same corner lists in
  the same order; LK points within 0.01 px with the same status, and
* the numpy restatement (oracle/frontend_oracle.py) on the same inputs, and
* a whole stream: ebt_new_image (detector + LK on the device) against the CPU
  oracle tracker fed by the numpy detector/LK — same tracks, same optimisations.
"""
import os

import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
from oracle import frontend_oracle as fo
from tests._oracle import CpuTracker, load_oracle
from tests.golden.tie_image import tie_image

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "frontend_cv2.npz"))
CASES = [(240, 180, 1001), (346, 260, 1002), (640, 480, 1003)]
LK_TOL_CV2_PX = 0.01     # against OpenCV (its SIMD float sums are ordered differently)
LK_TOL_ORACLE_PX = 1e-4  # against the restatement (same scalar order)


def frames_of(w, h, seed):
    return synth.make_stream(w, h, 0.12, 500e3, seed=seed, max_corners=10).frames[:3]


@pytest.mark.parametrize("w,h,seed", CASES)
def test_corner_lists_equal_cv2(w, h, seed):
    tr = Tracker(w, h)
    for k, img in enumerate(frames_of(w, h, seed)):
        c = tr.detect_features(img)
        ref = GOLD[f"corners_{w}_{k}"]
        assert c.shape == ref.shape and np.array_equal(c, ref.astype(np.float64)), f"frame {k}"
    tr.close()


def test_equal_responses_are_ordered_as_cv2_orders_them():
    img = tie_image()
    tr = Tracker(img.shape[1], img.shape[0])
    assert np.array_equal(tr.detect_features(img), GOLD["corners_ties"].astype(np.float64))
    tr.close()


def test_corner_edge_cases():
    tr = Tracker(240, 180)
    assert len(tr.detect_features(np.full((180, 240), 77, np.uint8))) == 0  # flat image: no response
    rng = np.random.default_rng(5)
    noise = rng.integers(0, 256, (180, 240), dtype=np.uint8)  # every pixel a candidate
    assert np.array_equal(tr.detect_features(noise), fo.detect_corners(noise).astype(np.float64))
    tr.close()


@pytest.mark.parametrize("w,h,seed", CASES)
def test_lk_matches_cv2_and_the_restatement(w, h, seed):
    fr = frames_of(w, h, seed)
    tr = Tracker(w, h)
    for k in range(2):
        pts = GOLD[f"lk_pts_{w}_{k}"]
        nxt, st = tr.lk_flow(fr[k], fr[k + 1], pts)
        assert np.array_equal(st, GOLD[f"lk_status_{w}_{k}"])
        ok = st == 1
        d = np.hypot(*(nxt[ok] - GOLD[f"lk_next_{w}_{k}"][ok]).T)
        assert d.max() <= LK_TOL_CV2_PX, d.max()
        if w <= 346:
            onxt, ost = fo.lk_flow(fr[k], fr[k + 1], pts)
            assert np.array_equal(st, ost)
            d2 = np.hypot(*(nxt[ok] - onxt[ok]).T)
            assert d2.max() <= LK_TOL_ORACLE_PX, d2.max()
    tr.close()


def test_lk_edge_cases():
    w, h = 240, 180
    fr = frames_of(w, h, 1001)
    tr = Tracker(w, h)
    # points outside the image, on the border, on a flat region (min-eigenvalue reject)
    flat = np.full((h, w), 90, np.uint8)
    pts = np.array([[-30.0, 10.0], [0.0, 0.0], [w - 1.0, h - 1.0], [w + 25.0, h + 25.0], [120.0, 90.0]], np.float32)
    for a, b in ((fr[0], fr[1]), (flat, flat)):
        nxt, st = tr.lk_flow(a, b, pts)
        onxt, ost = fo.lk_flow(a, b, pts)
        assert np.array_equal(st, ost)
        assert np.abs(nxt[st == 1] - onxt[st == 1]).max(initial=0) <= LK_TOL_ORACLE_PX
    nxt, st = tr.lk_flow(fr[0], fr[1], np.zeros((0, 2), np.float32))
    assert len(nxt) == 0 and len(st) == 0
    tr.close()


class CpuFrontEnd:
    """The CPU oracle tracker with FeatureDetector::newImage's OpenCV calls
    answered by the numpy restatement."""

    def __init__(self, cpu):
        self.cpu = cpu
        self.prev = None

    def __getattr__(self, k):
        return getattr(self.cpu, k)

    def new_image(self, gray, ts):
        self.cpu.new_frame_begin(gray, ts, fo.detect_corners(gray).astype(np.float64))
        ids, xy = self.cpu.get_uninit()
        if len(ids):
            nxt, st = fo.lk_flow(self.prev, gray, xy.astype(np.float32))
            for i, tid in enumerate(ids):
                self.cpu.set_flow(tid, int(st[i]), nxt[i, 0], nxt[i, 1])
        self.cpu.new_frame_end()
        self.prev = gray


def run_images(tracker, stream, n_frames):
    total = 0
    for k in range(n_frames):
        tracker.new_image(stream.frames[k], stream.frame_ts[k])
        total += tracker.push_events(stream.events_between(k))
    return total


def test_new_image_stream_parity():
    S = synth.make_stream(240, 180, 0.2, 500e3, seed=1001, max_corners=10)
    nf = min(4, len(S.frames))
    cpu = CpuFrontEnd(CpuTracker(load_oracle(), 240, 180, 100))
    gpu = Tracker(240, 180, 100)
    n_cpu = run_images(cpu, S, nf)
    n_gpu = run_images(gpu, S, nf)
    assert n_cpu == n_gpu and n_cpu > 100
    rc, rg = cpu.records(), gpu.records()
    assert np.array_equal(rc["track_id"], rg["track_id"])
    assert np.array_equal(rc["event_index"], rg["event_index"])
    assert np.hypot(rc["x"] - rg["x"], rc["y"] - rg["y"]).max() <= 0.05
    for arch in (False, True):
        tc, tg = cpu.tracks(arch), gpu.tracks(arch)
        assert np.array_equal(tc["track_id"], tg["track_id"])
        for k in ("lost", "init", "num_of_events", "n_optimizations", "time_without_update"):
            assert np.array_equal(tc[k], tg[k]), k
        assert np.abs(tc["flow_dir"] - tg["flow_dir"]).max(initial=0) <= np.deg2rad(0.5)
    cpu.close()
    gpu.close()
