"""CPU restatement (pure Python/numpy) of the reference's event ingestion — TEST
INFRASTRUCTURE ONLY (only tests/ may import this).

* parse_events_text: tools::Davis240cReader::getEventSample
  (/root/reference/tools/dataset_reader/src/davis240c_reader.cpp:60-92): per line
  std::stod(t) seconds -> duration_cast<microseconds>(duration<double>) =
  (int64)(sec * 1e6) (truncation), std::stoi x, y, sign; sign 0 -> NEGATIVE,
  1 -> POSITIVE, anything else throws.  Python's float() is correctly rounded,
  like glibc's strtod.
* replay_order: tools::Replayer::next (tools/replayer/src/replayer.cpp:56-98):
  the next sample is the event while event.ts < image.ts, otherwise the image;
  replay ends when either sequence runs out (finished(), :40-43).

Pinned by the reference's own known-answer test Davis240cReader.eventsTest
(tools/dataset_reader/test/davis240c_reader_test.cpp:19-49), restated in
tests/test_ingest.py.
"""
import numpy as np

from event_based_odomety_b200.records import EVENT


def parse_events_text(text: bytes) -> np.ndarray:
    lines = text.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    out = np.zeros(len(lines), EVENT)
    for i, ln in enumerate(lines):
        tok = ln.rstrip(b"\r").split(b" ")
        if len(tok) != 4:
            raise ValueError(f"line {i}: malformed")
        sec = float(tok[0])
        sign = int(tok[3])
        if sign not in (0, 1):
            raise ValueError(f"line {i}: Sign is not equal to 0/1")
        out["t_us"][i] = int(sec * 1e6)
        out["x"][i] = int(tok[1])
        out["y"][i] = int(tok[2])
        out["polarity"][i] = 1 if sign else -1
    return out


def format_events_text(events: np.ndarray) -> bytes:
    """The dataset's text format: seconds with 9 decimals, x, y, polarity 0/1."""
    t = events["t_us"].astype(np.int64)
    return "".join(f"{a // 1000000}.{a % 1000000:06d}000 {x} {y} {1 if p > 0 else 0}\n"
                   for a, x, y, p in zip(t.tolist(), events["x"].tolist(), events["y"].tolist(),
                                         events["polarity"].tolist())).encode()


def replay_order(event_ts, frame_ts):
    """List of ("events", lo, hi) / ("image", k) in delivery order."""
    seq, pos, n = [], 0, len(event_ts)
    for k, ts in enumerate(frame_ts):
        if pos >= n:
            break
        lo = pos
        while pos < n and event_ts[pos] < ts:
            pos += 1
        if pos > lo:
            seq.append(("events", lo, pos))
        if pos >= n:
            break
        seq.append(("image", k))
    return seq
