"""GPU parity of the two getter-only products of tracker::Patch (SURVEY 8a rows
a12 / a14): Patch::integrateMotionCompensatedEvents (patch.cpp:87-130) and
Patch::warpImage (patch.cpp:132-154), computed inside the stream kernel after
every optimisation (ebt_set_products) and read with ebt_get_patch(which = 1, 2).

The CPU oracle's versions are pinned against the reference build
(tests/test_oracle.py::test_products_match_reference_build).  Gates: the
motion-compensated patch is integer work -> bit-exact; predictedNabla within
1e-12 (it is a product of FP64 gradients and cos/sin of the flow angle)."""
import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import EbtError, Tracker
from tests._oracle import CpuTracker, load_oracle

pytestmark = pytest.mark.gpu
PRED_TOL = 1e-12


@pytest.mark.parametrize("mode", [1, 2])
def test_products_match_oracle(mode):
    S = synth.make_stream(240, 180, 0.2, 500e3, seed=1001, max_corners=40)
    cpu = CpuTracker(load_oracle(), 240, 180, 100, draw_images=(mode == 2))
    gpu = Tracker(240, 180, 100)
    gpu.set_products(mode)
    n_cpu = synth.run_stream(cpu, S)
    n_gpu = synth.run_stream(gpu, S)
    assert n_cpu == n_gpu > 1000
    seen = [0, 0]
    for arch in (False, True):
        tc, tg = cpu.tracks(arch), gpu.tracks(arch)
        assert np.array_equal(tc["track_id"], tg["track_id"])
        for tid in tc["track_id"]:
            assert np.array_equal(cpu.patch(tid, 0), gpu.patch(tid, 0))
            mc_c, mc_g = cpu.patch(tid, 1), gpu.patch(tid, 1)
            assert np.array_equal(mc_c, mc_g), f"motion-compensated patch of track {tid}"
            pr_c, pr_g = cpu.patch(tid, 2), gpu.patch(tid, 2)
            assert np.abs(pr_c - pr_g).max() <= PRED_TOL, f"predicted patch of track {tid}"
            seen[0] += int(np.abs(mc_c).sum() > 0)
            seen[1] += int(np.abs(pr_c).sum() > 0)
    assert seen[0] > 20 and seen[1] > 20  # the comparison saw real content
    cpu.close()
    gpu.close()


def test_products_are_off_by_default_and_say_so():
    S = synth.make_stream(240, 180, 0.1, 300e3, seed=7, max_corners=20)
    gpu = Tracker(240, 180, 100)
    synth.run_stream(gpu, S, 2)
    tid = int(gpu.tracks()["track_id"][0])
    gpu.patch(tid, 0)
    with pytest.raises(EbtError):
        gpu.patch(tid, 1)
    with pytest.raises(EbtError):
        gpu.patch(tid, 3)
    gpu.close()
