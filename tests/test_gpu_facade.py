"""GPU test of the C++ facade (include/ebt_tracker.hpp): a C++ driver written
like the reference's Evaluator (one updatePatches(event) call per event) must
produce the same tracks as the Python binding pushing whole batches."""
import os
import subprocess

import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import LIB_PATH, Tracker

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _w(f, arr):
    arr = np.ascontiguousarray(arr)
    f.write(np.int64(arr.size if arr.dtype.fields is None else len(arr)).tobytes())
    f.write(arr.tobytes())


@pytest.mark.gpu
def test_cpp_facade_matches_binding(tmp_path):
    S = synth.make_stream(240, 180, 0.10, 300e3, seed=21, max_corners=30)
    flows = []

    class Rec(Tracker):
        k = -1

        def new_frame_begin(self, gray, ts, corners):
            Rec.k += 1
            super().new_frame_begin(gray, ts, corners)

        def get_uninit(self, cap=8192):
            ids, xy = super().get_uninit(cap)
            self._xy = {int(i): p for i, p in zip(ids, xy)}
            return ids, xy

        def set_flow(self, tid, found, nx, ny):
            x, y = self._xy[int(tid)]
            flows.append((Rec.k, x, y, float(found), float(nx), float(ny)))
            super().set_flow(tid, found, nx, ny)

    t = Rec(240, 180, 100)
    n = synth.run_stream(t, S)
    want = t.tracks()
    n_arch = len(t.tracks(True))
    dump = tmp_path / "stream.bin"
    with open(dump, "wb") as f:
        _w(f, np.array([240, 180, len(S.frames), 100], np.int64))
        _w(f, np.array(S.frame_ts, np.int64))
        _w(f, np.stack(S.frames))
        _w(f, S.events)
        _w(f, np.array([len(c) for c in S.corners], np.int64))
        _w(f, np.concatenate(S.corners).astype(np.float64))
        _w(f, np.array(flows, np.float64))
    exe = tmp_path / "facade_stream"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "facade_stream.cpp"), "-o", str(exe),
                           LIB_PATH, f"-Wl,-rpath,{os.path.dirname(LIB_PATH)}"])
    out = subprocess.run([str(exe), str(dump)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    rows = [l.split() for l in out.stdout.splitlines()]
    tr = [r for r in rows if r[0] == "T"]
    tail = [r for r in rows if r[0] == "N"][0]
    assert int(tail[1]) == len(want) and int(tail[2]) == n_arch and int(tail[3]) == n
    assert [int(r[1]) for r in tr] == list(want["track_id"])
    got = np.array([[float(r[3]), float(r[4])] for r in tr])
    assert np.abs(got[:, 0] - (want["rect_x"] + 12)).max() < 1e-9
    assert np.abs(got[:, 1] - (want["rect_y"] + 12)).max() < 1e-9
    assert [int(r[6]) for r in tr] == list(want["n_trajectory"])
    t.close()
