"""GPU parity of Optimizer::drawCostMap (ft/src/optimizer.cpp:32-60; SURVEY 8f
n3): with OptimizerParams::drawCostMap on, every optimisation ends with 121
cost-only pixel passes of the solve's functor on the 11x11 grid of translation
offsets, read back through ebt_get_cost_map (Patch::getCostMap) and
ebt_get_batch_cost_maps.

The CPU oracle's version is pinned on the reference build's golden vectors
(tests/test_oracle.py::test_cost_map_against_reference_golden).  Gate: 1e-9 on
values of order 1 (FP64 both sides; the kernel forms the norm from the two
reduced sums of the pass instead of from the 625 residuals)."""
import os

import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import EbtError, Tracker
from tests._oracle import CpuBatch, CpuTracker, load_oracle

pytestmark = pytest.mark.gpu
TOL = 1e-9
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_batch_cost_maps_match_reference_golden_and_oracle():
    g = np.load(os.path.join(G, "problems_ref.npz"))
    cm_ref = np.load(os.path.join(G, "costmap_ref.npz"))["cost_maps"]
    tr = Tracker(240, 180)
    tr.add_frame(g["frame"])
    with pytest.raises(EbtError):
        tr.batch_cost_maps(1)  # nothing computed yet
    tr.set_cost_map(True)
    sol, nab = tr.optimize_batch(g["problems"], g["events"])
    cm = tr.batch_cost_maps(len(g["problems"]))
    assert np.abs(cm - cm_ref).max() <= TOL
    assert np.array_equal(nab.astype(np.int16), g["nabla"])
    assert np.hypot(sol["x"] - g["solutions"]["x"], sol["y"] - g["solutions"]["y"]).max() < 1e-6
    # a second, larger seeded batch against the oracle, incl. patches at the image border
    ps = synth.make_problems(200, 240, 180, seed=77)
    ob = CpuBatch(load_oracle(), 240, 180)
    ob.add_frame(g["frame"])
    _, cm_o = ob.optimize_cost_maps(ps.problems, ps.events)
    tr.optimize_batch(ps.problems, ps.events)
    cm_g = tr.batch_cost_maps(200)
    assert np.abs(cm_g - cm_o).max() <= TOL
    # switching it off stops the maps (and the 121 extra passes)
    tr.set_cost_map(False)
    tr.optimize_batch(ps.problems, ps.events)
    with pytest.raises(EbtError):
        tr.batch_cost_maps(200)
    tr.close()


def test_stream_cost_maps_match_oracle():
    ora = load_oracle()
    S = synth.make_stream(240, 180, 0.12, 400e3, seed=1001, max_corners=30)
    ora.set_draw_cost_map(1)
    try:
        cpu = CpuTracker(ora, 240, 180, 100)
        n_cpu = synth.run_stream(cpu, S)
    finally:
        ora.set_draw_cost_map(0)
    gpu = Tracker(240, 180, 100)
    gpu.set_cost_map(True)
    n_gpu = synth.run_stream(gpu, S)
    assert n_cpu == n_gpu > 300
    rc, rg = cpu.records(), gpu.records()
    assert np.array_equal(rc["event_index"], rg["event_index"])
    seen = 0
    for arch in (False, True):
        tc, tg = cpu.tracks(arch), gpu.tracks(arch)
        assert np.array_equal(tc["track_id"], tg["track_id"])
        for tid, nopt in zip(tc["track_id"], tc["n_optimizations"]):
            a, b = cpu.cost_map(tid), gpu.cost_map(tid)
            assert np.abs(a - b).max() <= TOL, f"cost map of track {tid}"
            if nopt > 0:
                assert a.min() > 0
                seen += 1
            else:
                assert not a.any() and not b.any()
    assert seen > 10
    cpu.close()
    gpu.close()


def test_cost_map_is_off_by_default_and_says_so():
    S = synth.make_stream(240, 180, 0.1, 300e3, seed=7, max_corners=20)
    gpu = Tracker(240, 180, 100)
    synth.run_stream(gpu, S, 2)
    with pytest.raises(EbtError):
        gpu.cost_map(int(gpu.tracks()["track_id"][0]))
    gpu.close()
