"""CPU tests of the drop-in boundary: the C-ABI library loads without a GPU,
exports every symbol include/ebt.h declares, its records have the sizes the
bindings assume, and computing entry points fail loudly without a device."""
import ctypes as C
import os
import re

import numpy as np

from event_based_odomety_b200 import cabi, records

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ebt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ebt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = cabi.load()
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ebt.h but not exported"
    assert set(cabi.EXPORTS) == set(names), set(cabi.EXPORTS) ^ set(names)


def test_record_sizes_match_bindings():
    lib = cabi.load()
    assert lib.ebt_abi_version() == 1
    assert lib.ebt_sizeof_event() == records.EVENT.itemsize == 16
    assert lib.ebt_sizeof_track() == records.TRACK.itemsize
    assert lib.ebt_sizeof_opt_record() == records.OPT_RECORD.itemsize
    assert lib.ebt_sizeof_problem() == records.PROBLEM.itemsize
    assert lib.ebt_sizeof_solution() == records.SOLUTION.itemsize
    assert lib.ebt_sizeof_params() == C.sizeof(cabi.Params)


def test_default_params_are_the_reference_defaults():
    lib = cabi.load()
    p = cabi.Params()
    assert lib.ebt_default_params(C.byref(p), 240, 180) == 0
    # feature_detector.h:10-31, optimizer.h:16-27, patch.h:113-114, patch.cpp:22,298
    assert (p.width, p.height, p.max_patches, p.init_num_events) == (240, 180, 100, 75)
    assert (p.min_num_events, p.max_num_events, p.batch_stride, p.max_iterations) == (100, 300, 30, 10)
    assert (p.association_distance, p.huber_loss, p.optimizer_threshold) == (5.0, 0.3, 0.6)
    assert lib.ebt_default_params(None, 240, 180) < 0
    assert lib.ebt_default_params(C.byref(p), 0, 180) < 0


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return  # covered by the gpu tests
    try:
        cabi.Tracker(240, 180)
    except cabi.EbtError as e:
        assert "no CUDA device" in str(e) or "cuda" in str(e).lower()
    else:
        raise AssertionError("Tracker() must fail without a GPU: there is no CPU fallback")


def test_product_path_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "event_based_odomety_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.lower() or f == "synth.py", f
    assert "libebt_oracle" not in open(os.path.join(pkg, "synth.py")).read()
