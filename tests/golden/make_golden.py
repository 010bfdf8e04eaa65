"""Regenerates the golden fixtures in this directory.  Runs only in the
authoring container: it needs oracle/_ref/libebt_ref.so (the reference's own
Patch/Optimizer/Ceres build, `make -C oracle ref`) and Python cv2.

    python tests/golden/make_golden.py

Fixtures (all inputs + the reference's outputs, so tests need neither
/root/reference nor cv2):
  problems_ref.npz   24 independent feature-patch problems -> Optimizer::optimize
  costmap_ref.npz    the same 24 problems with OptimizerParams::drawCostMap on ->
                     Patch::getCostMap() (11x11) of Optimizer::drawCostMap
  functor_ref.npz    residuals + Ceres-autodiff Jacobians of OptimizerCostFunctor
  stream_ref.npz     a 240x180 stream driven per event through the reference
                     harness (records, tracks, patches)
  warp_cv2.npz       cv2.warpAffine(INTER_NEAREST | WARP_INVERSE_MAP) index maps
  patch_known.npz    the reference's known-answer test Patch.integrateEventsTest
                     (patch_test.cpp:35-60) replayed through the reference build
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from event_based_odomety_b200 import synth  # noqa: E402
from event_based_odomety_b200.records import EVENT  # noqa: E402
from tests._oracle import CpuBatch, CpuTracker, _ptr, load_ref  # noqa: E402


def main():
    ref = load_ref()
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"

    # ---- independent problems ------------------------------------------
    ps = synth.make_problems(24, 240, 180, seed=3040)
    b = CpuBatch(ref, 240, 180)
    b.add_frame(ps.frames[0])
    sol, nab, _, _ = b.optimize(ps.problems, ps.events)
    gx, gy = b.gradients(0)
    np.savez_compressed(os.path.join(HERE, "problems_ref.npz"), frame=ps.frames[0], problems=ps.problems,
                        events=ps.events, solutions=sol, nabla=nab.astype(np.int16),
                        grad_probe=np.stack([gx[::17, ::13], gy[::17, ::13]]))

    # ---- Optimizer::drawCostMap on the same problems ----------------------
    _, cm = b.optimize_cost_maps(ps.problems, ps.events)
    np.savez_compressed(os.path.join(HERE, "costmap_ref.npz"), cost_maps=cm)

    # ---- functor residuals / autodiff Jacobians --------------------------
    rng = np.random.Generator(np.random.PCG64(9))
    cases = []
    for k in range(4):
        m = rng.standard_normal(625)
        m /= np.linalg.norm(m)
        th = np.deg2rad(rng.uniform(-20, 20))
        warp = np.array([np.cos(th), np.sin(th), rng.uniform(-4, 4), rng.uniform(-4, 4)])
        rx, ry = (rng.uniform(20, 190), rng.uniform(20, 130)) if k < 3 else (225.3, 166.1)  # last: leaves the image
        flow = rng.uniform(-3, 3)
        r = np.zeros(625); jw = np.zeros((625, 4)); jf = np.zeros(625); jl = np.zeros((625, 3))
        rc = ref.eval_functor(b.f_, 0, _ptr(m), rx, ry, _ptr(warp), flow, _ptr(r), _ptr(jw), _ptr(jf), _ptr(jl))
        assert rc == 0
        cases.append(dict(m=m, warp=warp, rect=np.array([rx, ry]), flow=flow, r=r, jl=jl, jf=jf))
    np.savez_compressed(os.path.join(HERE, "functor_ref.npz"), frame=ps.frames[0],
                        **{f"{k}_{i}": c[k] for i, c in enumerate(cases) for k in c})

    # ---- stream -----------------------------------------------------------
    S = synth.make_stream(240, 180, 0.14, 300e3, seed=1001, max_corners=40)
    flows = []

    class Rec(CpuTracker):
        def set_flow(self, tid, found, nx, ny):
            flows.append((len(frames_seen) - 1, int(tid), int(found), float(nx), float(ny)))
            super().set_flow(tid, found, nx, ny)

        def new_frame_begin(self, gray, ts, corners):
            frames_seen.append(ts)
            super().new_frame_begin(gray, ts, corners)

    frames_seen = []
    t = Rec(ref, 240, 180, 100, warp_image=True)  # Patch::warpImage after every optimisation, as the reference does
    n = synth.run_stream(t, S)
    tracks = np.concatenate([t.tracks(False), t.tracks(True)])
    patches = np.stack([t.patch(i) for i in tracks["track_id"]]).astype(np.int16)
    # getter-only products of the same run: motionCompensatedIntegratedNabla_ (integer valued), predictedNabla_
    np.savez_compressed(os.path.join(HERE, "products_ref.npz"),
                        mc=np.stack([t.patch(i, 1) for i in tracks["track_id"]]).astype(np.int16),
                        pred=np.stack([t.patch(i, 2) for i in tracks["track_id"]]))
    cmax = max(len(c) for c in S.corners)
    corners = np.full((len(S.corners), cmax, 2), np.nan)
    for i, c in enumerate(S.corners):
        corners[i, :len(c)] = c
    np.savez_compressed(os.path.join(HERE, "stream_ref.npz"), frames=np.stack(S.frames),
                        frame_ts=np.array(S.frame_ts), events=S.events, corners=corners,
                        flows=np.array(flows, dtype=np.float64), records=t.records(), tracks=tracks,
                        patches=patches, n_active=len(t.tracks(False)), n_opt=n)

    # ---- cv2 warpAffine nearest index maps ---------------------------------
    import cv2
    w, h = 120, 90
    idx = np.arange(w * h, dtype=np.float32).reshape(h, w)
    Ms, maps = [], []
    for k in range(8):
        th = np.deg2rad(rng.uniform(-25, 25))
        M = np.array([[np.cos(th), -np.sin(th), rng.uniform(-9, 9)], [np.sin(th), np.cos(th), rng.uniform(-9, 9)]])
        out = cv2.warpAffine(idx, M, (w, h), flags=cv2.INTER_NEAREST | cv2.WARP_INVERSE_MAP,
                             borderMode=cv2.BORDER_CONSTANT, borderValue=-1)
        Ms.append(M.ravel())
        maps.append(out.astype(np.int32))
    np.savez_compressed(os.path.join(HERE, "warp_cv2.npz"), M=np.array(Ms), maps=np.stack(maps), w=w, h=h)

    # ---- reference known-answer test (extent 3 at (10,10), 30 events) ------
    ev = np.zeros(30, EVENT)
    for i in range(30):
        ev[i]["x"], ev[i]["y"] = 7 + i // 7, 7 + i % 7
        ev[i]["polarity"] = 1 if i % 2 else -1
        ev[i]["t_us"] = i
    out = np.zeros(49)
    fb = np.zeros(4, np.int64)
    ref.patch_integrate(10.0, 10.0, 3, _ptr(ev), 30, _ptr(out), _ptr(fb))
    np.savez_compressed(os.path.join(HERE, "patch_known.npz"), events=ev, nabla=out.reshape(7, 7), front_back=fb)
    print("golden fixtures written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
