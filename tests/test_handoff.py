"""Tracker -> VO hand-off: Keyframe landmarks (vo/src/keyframe.cpp:5-14)
and CameraModel::unproject (common/include/common/camera_model.h:79-114).

CPU part: the C restatement against the reference's own known-answer test
(common/test/camera_model_test.cpp:5-35: project -> unproject returns the
normalised point) and, where oracle/_ref exists, bit for bit against the
reference header compiled unmodified.  GPU part: the kernel through the C-ABI
against the restatement (bit-exact: same operation order, no FMA contraction)."""
import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Camera, Tracker
from tests._oracle import cpu_project, cpu_unproject, load_oracle, load_ref

# common/test/camera_model_test.cpp:8-16
CAM = Camera(fx=199.092366542, fy=198.82882047, cx=132.192071378, cy=110.712660011,
             k1=-0.368436311798, k2=0.150947243557, k3=0.0, p1=-0.000296130534385, p2=-0.000759431726241)


def _grid_points():
    return np.array([[x, y, 10.0] for x in range(-9, 10) for y in range(-9, 10)])


def test_reference_known_answer_round_trip():
    ora = load_oracle()
    p = _grid_points()
    back = cpu_unproject(ora, CAM, cpu_project(ora, CAM, p))
    want = p / np.linalg.norm(p, axis=1, keepdims=True)
    # Eigen isApprox(a, b, 0.01): |a - b| <= 0.01 * min(|a|, |b|), both unit vectors here
    assert np.linalg.norm(back - want, axis=1).max() <= 0.01


@pytest.mark.skipif(load_ref() is None, reason="oracle/_ref not built here")
def test_restatement_equals_reference_header():
    ora, ref = load_oracle(), load_ref()
    rng = np.random.default_rng(4)
    pts = rng.uniform([-20, -20], [260, 200], (4000, 2))
    assert np.array_equal(cpu_unproject(ora, CAM, pts), cpu_unproject(ref, CAM, pts))
    xyz = rng.uniform([-5, -5, 2], [5, 5, 20], (4000, 3))
    assert np.array_equal(cpu_project(ora, CAM, xyz), cpu_project(ref, CAM, xyz))


@pytest.mark.gpu
def test_gpu_unproject_and_landmarks():
    ora = load_oracle()
    tr = Tracker(240, 180, 100)
    rng = np.random.default_rng(9)
    pts = rng.uniform([-20, -20], [260, 200], (5000, 2))
    assert np.array_equal(tr.unproject(CAM, pts), cpu_unproject(ora, CAM, pts))
    assert len(tr.unproject(CAM, np.zeros((0, 2)))) == 0
    ids, xy, b = tr.landmarks(CAM)
    assert len(ids) == 0
    S = synth.make_stream(240, 180, 0.1, 300e3, seed=3, max_corners=40)
    synth.run_stream(tr, S)
    t = tr.tracks()
    ids, xy, b = tr.landmarks(CAM)
    assert np.array_equal(ids, t["track_id"]) and len(ids) > 10
    assert np.array_equal(xy[:, 0], t["rect_x"] + 12) and np.array_equal(xy[:, 1], t["rect_y"] + 12)
    assert np.array_equal(b, cpu_unproject(ora, CAM, xy))
    ids2, xy2, none = tr.landmarks(None)
    assert none is None and np.array_equal(ids2, ids) and np.array_equal(xy2, xy)
    tr.close()
