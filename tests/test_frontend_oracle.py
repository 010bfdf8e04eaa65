"""The numpy restatement of the per-frame re-arm (oracle/frontend_oracle.py:
goodFeaturesToTrack/Harris and pyramidal LK with the reference's arguments)
against golden vectors recorded from cv2 (tests/golden/frontend_cv2.npz)."""
import os

import numpy as np
import pytest

from event_based_odomety_b200 import synth
from oracle import frontend_oracle as fo

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "frontend_cv2.npz"))
CASES = [(240, 180, 1001), (346, 260, 1002), (640, 480, 1003)]
LK_TOL_PX = 0.01


def frames_of(w, h, seed):
    return synth.make_stream(w, h, 0.12, 500e3, seed=seed, max_corners=10).frames[:3]


@pytest.mark.parametrize("w,h,seed", CASES)
def test_corners_match_cv2(w, h, seed):
    for k, img in enumerate(frames_of(w, h, seed)):
        r = fo.harris_response(img)
        ref = GOLD[f"harris_{w}_{k}"]
        assert np.abs(r[::7, ::5] - ref).max() <= 1e-5 * np.abs(ref).max()
        c = fo.detect_corners(img)
        assert np.array_equal(c, GOLD[f"corners_{w}_{k}"]), f"corner list of frame {k}"


def test_equal_responses_are_ordered_as_cv2_orders_them():
    from tests.golden.tie_image import tie_image
    assert np.array_equal(fo.detect_corners(tie_image()), GOLD["corners_ties"])


@pytest.mark.parametrize("w,h,seed", CASES[:2])
def test_lk_matches_cv2(w, h, seed):
    fr = frames_of(w, h, seed)
    for k in range(2):
        pts = GOLD[f"lk_pts_{w}_{k}"]
        nxt, st = fo.lk_flow(fr[k], fr[k + 1], pts)
        assert np.array_equal(st, GOLD[f"lk_status_{w}_{k}"])
        ok = st == 1
        assert ok.sum() >= len(pts) - 3
        d = np.hypot(*(nxt[ok] - GOLD[f"lk_next_{w}_{k}"][ok]).T)
        assert d.max() <= LK_TOL_PX, d.max()
