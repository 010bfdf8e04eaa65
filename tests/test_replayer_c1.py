"""BASELINE.json configs[0]: "feature_tracker via replayer on a synthetic shapes_6dof-format stream (DAVIS
240x180, ~100 Harris features, 25x25 patches)".

The C1 stream is written to disk in the reference's dataset layout — events.txt ("t_sec x y p"), images.txt
("t_sec file") and the image files — and replayed by the reference's OWN `tools::Davis240cReader` +
`tools::Replayer` (tools/dataset_reader/src/davis240c_reader.cpp, tools/replayer/src/replayer.cpp, compiled
unmodified into oracle/_ref/replay_driver by `make -C oracle replay`), whose callbacks feed the drop-in
tracker::FeatureDetector exactly as tools::Evaluator registers them.  The same recording goes through the
library's own path: the device text parser (ebt_parse_events_text) and ebt_replay.  Both must deliver the same
events and images in the same order, so the optimisation records (track, timestamp, final cost) and the final
tracks must be equal to the last bit — the same kernels saw the same inputs.

Image files are binary PGM behind the .png names of images.txt: OpenCV C++ is not in this image and the cv::
shim's imread decodes PGM only (the real cv::imread would read either)."""
import os
import subprocess

import numpy as np
import pytest

from event_based_odomety_b200 import synth
from oracle import ingest_oracle as io

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "replay_driver")


def write_dataset(path, S):
    """The reference's recording layout; returns (event text, image timestamps as the reader will read them)."""
    text = io.format_events_text(S.events)
    with open(os.path.join(path, "events.txt"), "wb") as f:
        f.write(text)
    os.makedirs(os.path.join(path, "images"))
    lines, ts_read = [], []
    for k, (fr, ts) in enumerate(zip(S.frames, S.frame_ts)):
        name = f"images/frame_{k:08d}.png"
        with open(os.path.join(path, name), "wb") as f:
            f.write(b"P5\n%d %d\n255\n" % (fr.shape[1], fr.shape[0]) + np.ascontiguousarray(fr, np.uint8).tobytes())
        sec = f"{ts // 1000000}.{ts % 1000000:06d}000"
        lines.append(f"{sec} {name}\n")
        ts_read.append(int(float(sec) * 1e6))  # stod, then duration_cast<microseconds> (davis240c_reader.cpp:96-101)
    with open(os.path.join(path, "images.txt"), "w") as f:
        f.write("".join(lines))
    return text, np.array(ts_read, np.int64)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(DRIVER), reason="oracle/_ref/replay_driver is built in the authoring container only")
def test_c1_through_the_reference_reader_and_replayer(tmp_path):
    from event_based_odomety_b200.cabi import Tracker
    S = synth.make_stream(240, 180, 0.30, 500e3, seed=1001, max_corners=69)
    text, frame_ts = write_dataset(str(tmp_path), S)
    out = str(tmp_path / "replayed.txt")
    subprocess.check_call([DRIVER, str(tmp_path), "240", "180", "100", out])
    got = {"rec": [], "track": [], "archived": []}
    for line in open(out):
        k, *v = line.split()
        if k in got:
            got[k].append([float(x) for x in v])
        else:
            got[k] = int(v[0])
    # the library's own path for the same recording
    tr = Tracker(240, 180, 100)
    events, used = tr.parse_events_text(text)
    assert used == len(text) and len(events) == len(S.events)
    delivered, n_opt = tr.replay(events, S.frames, frame_ts)
    rec, tracks, arch = tr.records(), tr.tracks(), tr.tracks(True)
    assert got["images"] == len(S.frames) and got["events"] == delivered
    assert n_opt == len(got["rec"]) > 3000
    r = np.array(got["rec"])
    assert np.array_equal(r[:, 0].astype(np.int32), rec["track_id"]) and np.array_equal(r[:, 1].astype(np.int64), rec["ts"])
    assert np.array_equal(r[:, 2], rec["final_cost"])          # %.17g round-trips a double: bit-identical
    t = np.array(got["track"])
    assert np.array_equal(t[:, 0].astype(np.int32), tracks["track_id"])
    assert np.array_equal(t[:, 1] != 0, tracks["lost"] != 0)
    assert np.array_equal(t[:, 2], tracks["rect_x"] + 12.0) and np.array_equal(t[:, 3], tracks["rect_y"] + 12.0)
    assert np.array_equal(t[:, 5].astype(np.int64), tracks["n_trajectory"])
    a = np.array(got["archived"]).reshape(-1, 2)
    # the driver's preExit is not called, so its archive is the library's
    assert np.array_equal(a[:, 0].astype(np.int32), arch["track_id"])
    tr.close()
