"""The reference's two known-answer tests of the per-event driver, replayed on both CPU
restatements (oracle/ebt_oracle.c and oracle/_ref's harness around the compiled Patch/Optimizer)
and, marked gpu, on the CUDA library through the C-ABI:

* FeatureDetector.updatePatchTest (ft/test/feature_detector_test.cpp:43-97): three patches, two of
  them overlapping; random events; every patch's window must hold exactly the events its rect
  contains, in stream order (Patch::isInPatch / addEvent, patch.cpp:37-47, 174-177).
* FeatureDetector.associatedPatchesTest (feature_detector_test.cpp:99-126): patches at (0,0), (5,5),
  (20,20), new corners at (3,0), (0,1), (18,18), associationDistance 5 -> 4 patches: the first and
  the third take a corner, (0,1) finds the first patch taken and becomes a new track.

The reference test uses patchExtent 5; the hot path fixes the reference's default 12 (25x25,
feature_detector.h:15), so the same layouts are replayed shifted into the image (patches must stay
5 px clear of the border, feature_detector.cpp:670-687) with the overlap pattern kept: centres 5 px
apart overlap, a third patch stands alone."""
import numpy as np
import pytest

from event_based_odomety_b200.records import EVENT
from tests._oracle import CpuTracker, load_oracle, load_ref

W, H = 120, 90
OFF = 30.0
CORNERS = np.array([[0, 0], [5, 5], [20, 20]], np.float64) * 1.0 + OFF
GRAY = (np.add.outer(np.arange(H), np.arange(W)) % 251).astype(np.uint8)


def _trackers(kind):
    if kind == "gpu":
        from event_based_odomety_b200.cabi import Tracker
        return Tracker(W, H, 10)
    lib = load_oracle() if kind == "oracle" else load_ref()
    if lib is None:
        pytest.skip("oracle/_ref/libebt_ref.so is not built here")
    return CpuTracker(lib, W, H, 10)


def _in_rect(c, x, y):
    # cv::Rect2d(corner - 12, 25 x 25).contains: x0 <= x < x0 + 25
    return (c[0] - 12 <= x < c[0] + 13) and (c[1] - 12 <= y < c[1] + 13)


def _update_patch(kind):
    t = _trackers(kind)
    t.new_frame_begin(GRAY, 0, CORNERS)
    t.new_frame_end()
    rng = np.random.Generator(np.random.PCG64(5))
    ev = np.zeros(60, EVENT)
    ev["t_us"] = np.arange(60)
    ev["x"] = rng.integers(int(OFF) - 15, int(OFF) + 36, 60)
    ev["y"] = rng.integers(int(OFF) - 15, int(OFF) + 36, 60)
    ev["polarity"] = rng.choice(np.array([-1, 1], np.int8), 60)
    # one event per call, as the reference test does, then the rest in one batch
    n = 0
    for i in range(5):
        n += t.push_events(ev[i:i + 1])
    n += t.push_events(ev[5:])
    assert n == 0  # no flow yet (isInit() false): the patches only collect events
    tr = t.tracks()
    assert list(tr["track_id"]) == [0, 1, 2]
    want = [sum(_in_rect(c, int(x), int(y)) for x, y in zip(ev["x"], ev["y"])) for c in CORNERS]
    assert min(want) > 0 and want[0] != want[2]
    assert list(tr["window_size"]) == want               # Patch::getEvents().size()
    if kind != "ref":                                    # Patch::counter_ has no getter in the reference
        assert list(tr["counter"]) == want
    # events inside both overlapping rects went to both patches
    both = sum(_in_rect(CORNERS[0], int(x), int(y)) and _in_rect(CORNERS[1], int(x), int(y))
               for x, y in zip(ev["x"], ev["y"]))
    assert both > 0
    t.close()


def _associated_patches(kind):
    t = _trackers(kind)
    t.new_frame_begin(GRAY, 0, CORNERS)
    t.new_frame_end()
    assert list(t.tracks()["track_id"]) == [0, 1, 2]
    new = np.array([[3, 0], [0, 1], [18, 18]], np.float64) + OFF
    t.new_frame_begin(GRAY, 44065, new)
    t.new_frame_end()
    tr = t.tracks()
    assert list(tr["track_id"]) == [0, 1, 2, 3]          # EXPECT_EQ(updatedPatches.size(), 4)
    # the three old patches keep their rects; the new track sits on the corner nobody took
    assert np.allclose(tr["rect_x"][:3] + 12, CORNERS[:, 0]) and np.allclose(tr["rect_y"][:3] + 12, CORNERS[:, 1])
    assert (tr["rect_x"][3] + 12, tr["rect_y"][3] + 12) == (OFF + 0, OFF + 1)
    t.close()


@pytest.mark.parametrize("kind", ["oracle", "ref"])
def test_update_patch_known_answer_cpu(kind):
    _update_patch(kind)


@pytest.mark.parametrize("kind", ["oracle", "ref"])
def test_associated_patches_known_answer_cpu(kind):
    _associated_patches(kind)


@pytest.mark.gpu
def test_update_patch_known_answer_gpu():
    _update_patch("gpu")


@pytest.mark.gpu
def test_associated_patches_known_answer_gpu():
    _associated_patches("gpu")
