"""ebt_parse_events_text_static — the event-text parser with its scratch in a
fixed __device__ array, for callers that must not allocate (include/ebt.h).
No counterpart in the reference.

Collected last on purpose.  Checks, through the
C-ABI, that the no-allocation entry point returns exactly what the existing
ebt_parse_events_text returns (bit-exact: integer/byte work) on two calls in a
row that share the static scratch — the second on a shorter text, so stale tile
counts of the first call are still in the array — plus the chunked mode, the
error path and the size limit."""
import numpy as np
import pytest

from event_based_odomety_b200 import cabi
from event_based_odomety_b200.cabi import EbtError, Tracker
from event_based_odomety_b200.records import EVENT
from tests.test_ingest import REF_TEXT, tricky_text

FIELDS = ("t_us", "x", "y", "polarity")


def _static_parse(torch, text, final=True, cap=None):
    """Device buffers come from torch (the caller's allocations); the entry point itself allocates nothing."""
    if cap is None:
        cap = text.count(b"\n") + 1
    d_text = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()
    d_out = torch.zeros(max(cap, 1) * EVENT.itemsize, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    n, used = cabi.parse_events_text_static(0, d_text.data_ptr(), len(text), d_out.data_ptr(), cap, final=final)
    out = np.frombuffer(d_out.cpu().numpy().tobytes(), EVENT)[:n].copy()
    return out, used


def _same(a, b):
    assert len(a) == len(b)
    for k in FIELDS:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.gpu
def test_static_parser_equals_the_allocating_entry_point_on_two_calls_in_a_row():
    import torch
    tr = Tracker(240, 180)
    first = tricky_text(120_000, seed=21)   # ~3.4 MB: several hundred tiles
    second = tricky_text(7_001, seed=22)[:-1]  # shorter, no newline after the last line
    for text in (first, second):
        want, want_used = tr.parse_events_text(text)
        got, used = _static_parse(torch, text)
        _same(got, want)
        assert used == want_used == len(text)
    # a third call on the first text again: the scratch now holds the second call's words
    want, _ = tr.parse_events_text(first)
    got, _ = _static_parse(torch, first)
    _same(got, want)
    # the reference's known-answer file, whole and as a non-final chunk
    want, want_used = tr.parse_events_text(REF_TEXT, final=False)
    got, used = _static_parse(torch, REF_TEXT, final=False)
    _same(got, want)
    assert used == want_used == REF_TEXT.rindex(b"\n") + 1
    got, used = _static_parse(torch, REF_TEXT)
    assert list(got["t_us"]) == [0, 11, 50, 55, 80] and used == len(REF_TEXT)
    # a text without any newline in a non-final chunk: nothing parsed, nothing consumed
    got, used = _static_parse(torch, b"0.5 1 2", final=False)
    assert len(got) == 0 and used == 0
    tr.close()


@pytest.mark.gpu
def test_static_parser_errors_and_limit():
    import torch
    for bad in (b"0.1 1 2 2\n", b"0.5 1 2 1\n0.6 70000 2 1\n", b"abc 1 2 1\n"):
        with pytest.raises(EbtError):
            _static_parse(torch, bad)
    with pytest.raises(EbtError, match="output holds"):
        _static_parse(torch, REF_TEXT, cap=3)
    # after failures the next call is clean again (the error word is re-armed every call)
    got, _ = _static_parse(torch, REF_TEXT)
    assert len(got) == 5
    # beyond the stated maximum: rejected before anything is touched
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(EbtError, match="at most"):
        cabi.parse_events_text_static(0, d.data_ptr(), cabi.STATIC_PARSE_MAX_BYTES + 1, d.data_ptr(), 1)
    # empty text is fine
    n, used = cabi.parse_events_text_static(0, d.data_ptr(), 0, d.data_ptr(), 1)
    assert (n, used) == (0, 0)
