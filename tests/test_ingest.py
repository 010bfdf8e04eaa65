"""Event ingestion: the device text parser (ebt_parse_events_text)
and the Replayer order (ebt_replay).

CPU part pins the Python restatement (oracle/ingest_oracle.py) on the
reference's known-answer test Davis240cReader.eventsTest
(tools/dataset_reader/test/davis240c_reader_test.cpp:19-49; the five lines of
its events.txt are restated below).  GPU part: the kernels through the C-ABI
against the restatement — bit-exact, integer/byte work."""
import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import EbtError, Tracker
from event_based_odomety_b200.records import EVENT
from oracle import ingest_oracle as io

# tools/dataset_reader/test/test_data/events.txt (no newline after the last line)
REF_TEXT = (b"0.000000000 33 39 1\n0.000011001 158 145 1\n0.000050000 88 143 0\n"
            b"0.000055000 174 154 0\n0.000080001 112 139 1")
# davis240c_reader_test.cpp:26-38
REF_POINTS = [(33, 39), (158, 145), (88, 143), (174, 154), (112, 139)]
REF_SIGNS = [1, 1, -1, -1, 1]
REF_TS = [0, 11, 50, 55, 80]


def _check_reference_answer(ev):
    assert len(ev) == 5
    assert list(ev["t_us"]) == REF_TS
    assert [(int(x), int(y)) for x, y in zip(ev["x"], ev["y"])] == REF_POINTS
    assert list(ev["polarity"]) == REF_SIGNS


def tricky_text(n, seed=0):
    """Timestamps whose decimal -> double -> microsecond chain truncates below the
    printed value (0.000249 -> 248), long integer parts, 3..9 decimals, CRLF."""
    rng = np.random.default_rng(seed)
    t = np.sort(rng.integers(0, 4_000_000_000, n))
    lines = []
    for i, a in enumerate(t.tolist()):
        nd = (9, 6, 3)[i % 3] if i % 7 else 9
        if nd == 9:
            ts = f"{a // 1000000}.{a % 1000000:06d}{int(rng.integers(0, 1000)):03d}"
        elif nd == 6:
            ts = f"{a // 1000000}.{a % 1000000:06d}"
        else:
            ts = f"{a // 1000000}.{(a % 1000000) // 1000:03d}"
        eol = "\r\n" if i % 11 == 0 else "\n"
        lines.append(f"{ts} {int(rng.integers(0, 640))} {int(rng.integers(0, 480))} {int(rng.integers(0, 2))}{eol}")
    return "".join(lines).encode()


def test_restatement_on_reference_known_answer():
    _check_reference_answer(io.parse_events_text(REF_TEXT))
    ev = io.parse_events_text(b"0.000249000 1 2 1\n0.000249 1 2 0\n")
    assert list(ev["t_us"]) == [int(0.000249 * 1e6)] * 2 == [248, 248]  # (int64)(stod * 1e6) truncates
    with pytest.raises(ValueError):
        io.parse_events_text(b"0.1 1 2 2\n")


def test_cpp_restatement_equals_python_restatement():
    from tests._oracle import cpu_parse_text, load_ingest
    dll = load_ingest()
    _check_reference_answer(cpu_parse_text(dll, REF_TEXT))
    text = tricky_text(20_000, seed=5)
    a, b = cpu_parse_text(dll, text), io.parse_events_text(text)
    for k in ("t_us", "x", "y", "polarity"):
        assert np.array_equal(a[k], b[k]), k
    with pytest.raises(ValueError):
        cpu_parse_text(dll, b"0.1 1 2 2\n")


def test_replay_order_restatement():
    # event before image while t_ev < t_img; image first on a tie; stop when either runs out
    seq = io.replay_order([1, 2, 5, 5, 9, 12], [5, 9, 20, 30])
    assert seq == [("events", 0, 2), ("image", 0), ("events", 2, 4), ("image", 1), ("events", 4, 6)]
    assert io.replay_order([7, 8], [5, 6]) == [("image", 0), ("image", 1)]
    assert io.replay_order([], [5]) == []


@pytest.mark.gpu
def test_gpu_parser_reference_known_answer_and_edges():
    tr = Tracker(240, 180)
    ev, used = tr.parse_events_text(REF_TEXT)
    _check_reference_answer(ev)
    assert used == len(REF_TEXT)
    # non-final chunk: the unterminated last line is left to the caller
    ev, used = tr.parse_events_text(REF_TEXT, final=False)
    assert len(ev) == 4 and used == REF_TEXT.rindex(b"\n") + 1
    ev, used = tr.parse_events_text(b"")
    assert len(ev) == 0
    ev, _ = tr.parse_events_text(b"12.5 7 9 0\n")
    assert (int(ev["t_us"][0]), int(ev["x"][0]), int(ev["y"][0]), int(ev["polarity"][0])) == (12500000, 7, 9, -1)
    for bad in (b"0.1 1 2 2\n", b"0.1 1 2\n", b"\n", b"0.5 1 2 1\n0.6 70000 2 1\n", b"abc 1 2 1\n",
                b"12345678901234567.123 1 2 1\n"):
        with pytest.raises(EbtError):
            tr.parse_events_text(bad)
    with pytest.raises(EbtError):
        tr.parse_events_text(REF_TEXT, cap=3)
    tr.close()


@pytest.mark.gpu
def test_gpu_parser_bit_exact_on_large_text():
    tr = Tracker(240, 180)
    text = tricky_text(200_000, seed=3)
    want = io.parse_events_text(text)
    got, used = tr.parse_events_text(text)
    assert used == len(text) and len(got) == len(want)
    for k in ("t_us", "x", "y", "polarity"):
        assert np.array_equal(got[k], want[k]), k
    # streamed in chunks that cut lines anywhere: same events
    parts, pos = [], 0
    chunk = 1_000_003
    while pos < len(text):
        piece = text[pos:pos + chunk]
        final = pos + chunk >= len(text)
        ev, used = tr.parse_events_text(piece, final=final)
        parts.append(ev)
        pos += used
        assert used > 0
    cat = np.concatenate(parts)
    assert np.array_equal(cat["t_us"], want["t_us"]) and np.array_equal(cat["x"], want["x"])
    # round trip of a generated stream through the dataset's text format
    S = synth.make_stream(240, 180, 0.1, 300e3, seed=3, max_corners=10)
    text = io.format_events_text(S.events)
    back, _ = tr.parse_events_text(text)
    want = io.parse_events_text(text)
    for k in ("t_us", "x", "y", "polarity"):
        assert np.array_equal(back[k], want[k]), k
    # the reference's seconds -> microseconds chain may land one below the printed value, never elsewhere
    d = S.events["t_us"] - back["t_us"]
    assert d.min() >= 0 and d.max() <= 1 and np.array_equal(back["x"], S.events["x"])
    tr.close()


@pytest.mark.gpu
def test_gpu_replay_follows_the_replayer_order():
    S = synth.make_stream(240, 180, 0.2, 400e3, seed=1001, max_corners=10)
    a, b = Tracker(240, 180, 100), Tracker(240, 180, 100)
    nd, no = a.replay(S.events, S.frames, S.frame_ts)
    total = 0
    for item in io.replay_order(S.events["t_us"], S.frame_ts):
        if item[0] == "events":
            total += b.push_events(S.events[item[1]:item[2]])
        else:
            b.new_image(S.frames[item[1]], S.frame_ts[item[1]])
    assert no == total > 100 and nd <= len(S.events)
    ra, rb = a.records(), b.records()
    assert np.array_equal(ra["event_index"], rb["event_index"]) and np.array_equal(ra["track_id"], rb["track_id"])
    assert np.array_equal(ra["x"], rb["x"])
    a.close()
    b.close()
