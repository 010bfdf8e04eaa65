"""GPU parity, batched entry: ebt_optimize_batch (CUDA, through the C-ABI) against
the CPU oracle on the same seeded problems.

Gates (BASELINE.json north_star): integer event patches bit-exact; feature
positions within 0.05 px; flow angle within 0.5 degrees; same iteration and
termination path (accepted/rejected step counts and termination type)."""
import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
from tests._oracle import CpuBatch, load_oracle, load_ref

POS_TOL_PX = 0.05
ANGLE_TOL_DEG = 0.5


def _angle_diff_deg(a, b):
    d = np.abs(a - b) % (2 * np.pi)
    return np.rad2deg(np.minimum(d, 2 * np.pi - d))


def run_case(n, w, h, seed, n_frames=1, **kw):
    ps = synth.make_problems(n, w, h, seed=seed, n_frames=n_frames, **kw)
    ob = CpuBatch(load_oracle(), w, h)
    tr = Tracker(w, h)
    for f in ps.frames:
        so = ob.add_frame(f)
        sg = tr.add_frame(f)
        assert so == sg
        go, gg = ob.gradients(so), tr.frame_gradients(sg)
        assert np.array_equal(go[0], gg[0]) and np.array_equal(go[1], gg[1]), "gradient grid not bit-exact"
    osol, onab, _, _ = ob.optimize(ps.problems, ps.events)
    gsol, gnab = tr.optimize_batch(ps.problems, ps.events)
    tr.close()
    return ps, osol, onab, gsol, gnab


@pytest.mark.gpu
@pytest.mark.parametrize("n,w,h,seed,nf", [(64, 240, 180, 3064, 1), (500, 346, 260, 3500, 2),
                                           (2000, 640, 480, 5000, 3)])
def test_batch_parity(n, w, h, seed, nf):
    ps, osol, onab, gsol, gnab = run_case(n, w, h, seed, nf)
    assert np.array_equal(onab.astype(np.int32), gnab), "integer event patches differ"
    assert np.array_equal(osol["ts"], gsol["ts"])
    dpos = np.hypot(osol["x"] - gsol["x"], osol["y"] - gsol["y"])
    dang = _angle_diff_deg(osol["flow_dir"], gsol["flow_dir"])
    same_path = ((osol["n_successful"] == gsol["n_successful"])
                 & (osol["n_unsuccessful"] == gsol["n_unsuccessful"])
                 & (osol["termination"] == gsol["termination"]))
    print(f"\n[n={n}] max dpos {dpos.max():.3e} px, max dangle {dang.max():.3e} deg, "
          f"max dcost {np.abs(osol['final_cost'] - gsol['final_cost']).max():.3e}, "
          f"same LM path {same_path.mean():.4f}")
    assert dpos.max() <= POS_TOL_PX
    assert dang.max() <= ANGLE_TOL_DEG
    assert np.abs(osol["final_cost"] - gsol["final_cost"]).max() < 1e-6
    assert same_path.mean() >= 0.99


@pytest.mark.gpu
def test_batch_edge_windows():
    """empty-in-rect window (NaN patch -> Ceres FAILURE path), minimum and maximum windows"""
    ps = synth.make_problems(8, 240, 180, seed=77)
    ev = ps.events.copy()
    p0 = ps.problems[0]
    sl = slice(p0["event_offset"], p0["event_offset"] + p0["n_events"])
    ev["x"][sl] = 0  # all events outside the rect -> ||nabla|| = 0 -> NaN
    ev["y"][sl] = 0
    ob = CpuBatch(load_oracle(), 240, 180)
    tr = Tracker(240, 180)
    ob.add_frame(ps.frames[0]); tr.add_frame(ps.frames[0])
    osol, onab, _, _ = ob.optimize(ps.problems, ev)
    gsol, gnab = tr.optimize_batch(ps.problems, ev)
    assert np.array_equal(onab.astype(np.int32), gnab)
    assert osol["termination"][0] == 2 and gsol["termination"][0] == 2
    assert gsol["final_cost"][0] == osol["final_cost"][0] == -1.0
    assert np.allclose(osol["x"], gsol["x"], atol=POS_TOL_PX) and np.allclose(osol["y"], gsol["y"], atol=POS_TOL_PX)


@pytest.mark.gpu
def test_batch_border_patches():
    """Patches hanging over the image border: pixels that warp outside the image
    contribute u = 0 without derivative (optimizer_cost.h:47-90) and the bicubic
    taps clamp to the edge (cubic_interpolation.h:403-421)."""
    ps, osol, onab, gsol, gnab = run_case(256, 240, 180, 3111, 1, border=2)
    assert np.array_equal(onab.astype(np.int32), gnab)
    dpos = np.hypot(osol["x"] - gsol["x"], osol["y"] - gsol["y"])
    same_path = ((osol["n_successful"] == gsol["n_successful"])
                 & (osol["n_unsuccessful"] == gsol["n_unsuccessful"])
                 & (osol["termination"] == gsol["termination"]))
    print(f"\n[border] max dpos {np.nanmax(dpos):.3e} px, same LM path {same_path.mean():.4f}")
    ok = np.isfinite(osol["x"])
    assert np.array_equal(ok, np.isfinite(gsol["x"]))
    assert dpos[ok].max() <= POS_TOL_PX
    assert _angle_diff_deg(osol["flow_dir"][ok], gsol["flow_dir"][ok]).max() <= ANGLE_TOL_DEG
    assert same_path.mean() >= 0.99


@pytest.mark.gpu
def test_batch_tile_fallbacks():
    """The solve samples the template grid from a 36x36 shared-memory tile staged around the initial
    estimate; two families of problems that leave it: (a) rotations of 25-60 degrees (the warped rect's
    bounding box does not fit: nothing is staged), (b) initial estimates displaced by ~8 px in the template
    (the iterates walk out of the tile: per-pixel fallback to global memory).  Same gates as everywhere."""
    w, h = 240, 180
    ps = synth.make_problems(192, w, h, seed=4711, border=45)
    pr = ps.problems.copy()
    rng = np.random.Generator(np.random.PCG64(5))
    for k in range(len(pr)):
        c, s, tx, ty = pr["warp"][k]
        init = np.array([pr["init_x"][k], pr["init_y"][k]])
        R = np.array([[c, -s], [s, c]])
        ctr = R.T @ (init - np.array([tx, ty]))  # Patch::updatePatchRect: centre = warp^-1 * initPoint
        if k % 2 == 0:
            th = np.arctan2(s, c) + np.deg2rad(rng.uniform(25, 60)) * rng.choice([-1, 1])
            R2 = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
            t2 = init - R2 @ ctr  # same centre, so the window's events stay in the rect
            pr["warp"][k] = [np.cos(th), np.sin(th), t2[0], t2[1]]
        else:
            d = rng.uniform(5, 8, size=2) * rng.choice([-1, 1], size=2)
            pr["init_x"][k] += d[0]
            pr["init_y"][k] += d[1]
            pr["warp"][k] = [c, s, tx + d[0], ty + d[1]]  # same centre, template sampled d away
    ob = CpuBatch(load_oracle(), w, h)
    tr = Tracker(w, h)
    ob.add_frame(ps.frames[0]); tr.add_frame(ps.frames[0])
    osol, onab, _, _ = ob.optimize(pr, ps.events)
    gsol, gnab = tr.optimize_batch(pr, ps.events)
    tr.close()
    assert np.array_equal(onab.astype(np.int32), gnab)
    assert np.abs(onab).sum(axis=1).min() > 0  # every window contributes: no trivial FAILURE cases
    same_path = ((osol["n_successful"] == gsol["n_successful"])
                 & (osol["n_unsuccessful"] == gsol["n_unsuccessful"])
                 & (osol["termination"] == gsol["termination"]))
    dpos = np.hypot(osol["x"] - gsol["x"], osol["y"] - gsol["y"])
    print(f"\n[tile fallbacks] max dpos {dpos.max():.3e} px, same LM path {same_path.mean():.4f}, "
          f"rejected steps in {np.count_nonzero(osol['n_unsuccessful'])} solves")
    assert dpos.max() <= POS_TOL_PX
    assert _angle_diff_deg(osol["flow_dir"], gsol["flow_dir"]).max() <= ANGLE_TOL_DEG
    assert same_path.mean() >= 0.99


@pytest.mark.gpu
def test_batch_shared_pixel_warp_kernel_edges():
    """More than 4 problems per SM select the shared-pixel-warp kernel (two leader warps share four pixel
    warps).  Border patches (clamped tile), an empty-in-rect window (NaN patch -> FAILURE) and the rotated /
    displaced problems that leave the tile, all inside one batch of 1200."""
    w, h = 240, 180
    ps = synth.make_problems(1200, w, h, seed=8123, border=2)
    pr, ev = ps.problems.copy(), ps.events.copy()
    p0 = pr[7]
    sl = slice(p0["event_offset"], p0["event_offset"] + p0["n_events"])
    ev["x"][sl] = 0
    ev["y"][sl] = 0
    rng = np.random.Generator(np.random.PCG64(9))
    for k in range(100, 400):  # rotate about the unchanged centre: nothing staged
        c, s_, tx, ty = pr["warp"][k]
        init = np.array([pr["init_x"][k], pr["init_y"][k]])
        R = np.array([[c, -s_], [s_, c]])
        ctr = R.T @ (init - np.array([tx, ty]))
        th = np.arctan2(s_, c) + np.deg2rad(rng.uniform(25, 60)) * rng.choice([-1, 1])
        R2 = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
        t2 = init - R2 @ ctr
        pr["warp"][k] = [np.cos(th), np.sin(th), t2[0], t2[1]]
    ob = CpuBatch(load_oracle(), w, h)
    tr = Tracker(w, h)
    ob.add_frame(ps.frames[0]); tr.add_frame(ps.frames[0])
    osol, onab, _, _ = ob.optimize(pr, ev)
    gsol, gnab = tr.optimize_batch(pr, ev)
    tr.close()
    assert np.array_equal(onab.astype(np.int32), gnab)
    assert osol["termination"][7] == 2 and gsol["termination"][7] == 2 and gsol["final_cost"][7] == -1.0
    ok = np.isfinite(osol["x"])
    assert np.array_equal(ok, np.isfinite(gsol["x"]))
    assert np.array_equal(osol["ts"], gsol["ts"])
    dpos = np.hypot(osol["x"] - gsol["x"], osol["y"] - gsol["y"])[ok]
    same_path = ((osol["n_successful"] == gsol["n_successful"])
                 & (osol["n_unsuccessful"] == gsol["n_unsuccessful"])
                 & (osol["termination"] == gsol["termination"]))
    print(f"\n[shared kernel edges] max dpos {dpos.max():.3e} px, same LM path {same_path.mean():.4f}")
    assert dpos.max() <= POS_TOL_PX
    assert _angle_diff_deg(osol["flow_dir"][ok], gsol["flow_dir"][ok]).max() <= ANGLE_TOL_DEG
    assert same_path.mean() >= 0.99


@pytest.mark.gpu
def test_batch_parity_against_the_reference_build():
    """500 problems against oracle/_ref (the reference's Patch + Optimizer + Ceres compiled unmodified),
    not its port: integer patches bit-exact, positions / angle within the north-star gates, costs 1e-6."""
    ref = load_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libebt_ref.so is not built here")
    ps = synth.make_problems(500, 346, 260, seed=3500, n_frames=2)
    ob = CpuBatch(ref, 346, 260)
    tr = Tracker(346, 260)
    for f in ps.frames:
        assert ob.add_frame(f) == tr.add_frame(f)
    osol, onab, _, _ = ob.optimize(ps.problems, ps.events)
    gsol, gnab = tr.optimize_batch(ps.problems, ps.events)
    tr.close()
    assert np.array_equal(onab.astype(np.int32), gnab), "integer event patches differ"
    dpos = np.hypot(osol["x"] - gsol["x"], osol["y"] - gsol["y"])
    dang = _angle_diff_deg(osol["flow_dir"], gsol["flow_dir"])
    print(f"\nvs reference build: max dpos {dpos.max():.3e} px, max dangle {dang.max():.3e} deg, "
          f"max dcost {np.abs(osol['final_cost'] - gsol['final_cost']).max():.3e}")
    assert dpos.max() <= POS_TOL_PX and dang.max() <= ANGLE_TOL_DEG
    assert np.abs(osol["final_cost"] - gsol["final_cost"]).max() < 1e-6


@pytest.mark.gpu
def test_batch_kernels_are_deterministic():
    """Stand-in for racecheck (closed on this pool): optimize_batch_shared_kernel hands jobs over through
    sequence numbers in shared memory.  2000 problems (shared-pixel-warp kernel) and 100 problems (one CTA
    per problem), 50 launches each: bit-identical solutions and integer patches."""
    for n in (2000, 100):
        ps = synth.make_problems(n, 346, 260, seed=4100 + n)
        tr = Tracker(346, 260)
        tr.add_frame(ps.frames[0])
        sol0, nab0 = tr.optimize_batch(ps.problems, ps.events)
        for _ in range(49):
            sol, nab = tr.optimize_batch(ps.problems, ps.events)
            assert sol.tobytes() == sol0.tobytes() and np.array_equal(nab, nab0)
        tr.close()
