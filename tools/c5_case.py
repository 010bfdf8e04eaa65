"""configs[4] (C5) helper: the synthetic stream, its ground-truth camera poses and the binary input files of
oracle/_ref/vo_driver (the reference's visual-odometry back end behind the drop-in tracker headers).

The stream generator moves a textured plane under the camera: image point p(t) of a plane point follows
p = R(-theta)(Q - c - o(t)) + c, i.e. a camera at depth Z over a fronto-parallel plane, translating by
T(t) = o(t) * Z / f parallel to it and rolling by theta(t).  The renderer has no lens distortion, so the
calibration is the pinhole part of common/test/camera_model_test.cpp:8-16 (fx 199.09 fy 198.83 cx 132.19
cy 110.71) with the distortion coefficients set to zero.
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from event_based_odomety_b200 import synth  # noqa: E402

W, H = 240, 180
SEED = 1005
RATE = 0.5e6
MAX_PATCHES = 150
CALIB = np.array([199.09, 198.83, 132.19, 110.71, 0.0, 0.0, 0.0, 0.0, 0.0])  # fx fy cx cy k1 k2 k3 p1 p2
DEPTH = 1.0
VO_DRIVER = os.path.join(ROOT, "oracle", "_ref", "vo_driver")


def make_stream(duration):
    return synth.make_stream(W, H, duration, RATE, SEED, max_corners=69)


def ground_truth(S, step_us=5000):
    """rows: t_us tx ty tz qx qy qz qw (camera-to-world), every step_us and at every image."""
    ts = sorted(set(list(range(0, int(S.frame_ts[-1]) + step_us, step_us)) + [int(t) for t in S.frame_ts]))
    rows = []
    for t in ts:
        th, ox, oy = S.pose(t * 1e-6)
        rows.append([t, ox * DEPTH / CALIB[0], oy * DEPTH / CALIB[1], 0.0, 0.0, 0.0, np.sin(th / 2), np.cos(th / 2)])
    return np.array(rows, np.float64)


def _w(f, arr):
    arr = np.ascontiguousarray(arr)
    f.write(np.int64(len(arr) if arr.dtype.fields is not None else arr.size).tobytes())
    f.write(arr.tobytes())


def write_common(f, S, n_frames):
    _w(f, np.array([W, H, n_frames, MAX_PATCHES], np.int64))
    _w(f, CALIB)
    _w(f, np.array(S.frame_ts[:n_frames], np.int64))
    _w(f, ground_truth(S).reshape(-1))


def write_gpu_input(path, S, n_frames=None):
    n_frames = n_frames or len(S.frames)
    with open(path, "wb") as f:
        write_common(f, S, n_frames)
        _w(f, np.stack(S.frames[:n_frames]).reshape(-1))
        _w(f, S.events)


def write_tracks_input(path, S, per_frame, n_frames=None):
    """per_frame: list of (ids int32[n], xy float64[n, 2]) — the live patches after every image."""
    n_frames = n_frames or len(S.frames)
    with open(path, "wb") as f:
        write_common(f, S, n_frames)
        _w(f, np.array([len(i) for i, _ in per_frame[:n_frames]], np.int64))
        _w(f, np.concatenate([i for i, _ in per_frame[:n_frames]]).astype(np.int32))
        _w(f, np.concatenate([x.reshape(-1, 2) for _, x in per_frame[:n_frames]]).astype(np.float64).reshape(-1))


def run_vo(mode, inp, out, dump=None):
    subprocess.check_call([VO_DRIVER, mode, inp, out] + ([dump] if dump else []))
    res = {"pose": []}
    for line in open(out):
        k, *v = line.split()
        if k == "pose":
            res["pose"].append([float(x) for x in v])
        elif k == "mode":
            res[k] = v[0]
        else:
            res[k] = float(v[0])
    res["pose"] = np.array(res["pose"])
    return res


def tracks_of(tracker, S, n_frames=None):
    """Drive a tracker exposing the C-ABI call shapes (cabi.Tracker with its device detector and LK, or a
    CPU checker fed by the same corners / flows) through the stream; returns the live patches after every
    image as (ids, corners)."""
    n_frames = n_frames or len(S.frames)
    out = []
    for k in range(n_frames):
        tracker.new_image(S.frames[k], S.frame_ts[k])
        t = tracker.tracks()
        out.append((t["track_id"].astype(np.int32), np.stack([t["rect_x"] + 12.0, t["rect_y"] + 12.0], axis=1)))
        if k + 1 < n_frames:  # Replayer::next ends with the last image (replayer.cpp:40-43)
            tracker.push_events(S.events_between(k))
    return out


if __name__ == "__main__":
    # smoke: ideal tracks (ground-truth image motion of fixed plane points) through the VO back end
    dur = float(sys.argv[1]) if len(sys.argv) > 1 else 1.5
    S = make_stream(dur)
    rng = np.random.Generator(np.random.PCG64(3))
    pts0 = np.stack([rng.uniform(15, W - 15, 120), rng.uniform(15, H - 15, 120)], axis=1)
    q = S.to_texture(0.0, pts0)
    per = []
    for k in range(len(S.frames)):
        p = S.from_texture(S.frame_ts[k] * 1e-6, q)
        ok = (p[:, 0] > 8) & (p[:, 0] < W - 8) & (p[:, 1] > 8) & (p[:, 1] < H - 8)
        per.append((np.nonzero(ok)[0].astype(np.int32), p[ok]))
    os.makedirs("/tmp/c5", exist_ok=True)
    write_tracks_input("/tmp/c5/ideal.bin", S, per)
    r = run_vo("tracks", "/tmp/c5/ideal.bin", "/tmp/c5/ideal.txt")
    print({k: v for k, v in r.items() if k != "pose"})
    print(r["pose"][:5])
