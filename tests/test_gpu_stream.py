"""GPU parity, stream mode: the tracker driven frame by frame / event batch by
event batch through the C-ABI (ebt_new_frame_* + ebt_push_events) against the
CPU oracle's per-event FeatureDetector::updatePatches loop on the same seeded
synthetic stream.

Gates: same optimisations at the same stream events for the same tracks (the
trigger rule is integer logic), integer event patches bit-exact, window sizes
equal, positions within 0.05 px, flow angle within 0.5 degrees, same lost
flags and archive."""
import numpy as np
import pytest

from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
from tests._oracle import CpuTracker, load_oracle, load_ref

POS_TOL_PX = 0.05
ANGLE_TOL_DEG = 0.5


def _angle_diff_deg(a, b):
    d = np.abs(a - b) % (2 * np.pi)
    return np.rad2deg(np.minimum(d, 2 * np.pi - d))


def checker(kind):
    """'oracle' = the plain-C restatement; 'ref' = oracle/_ref/libebt_ref.so, the reference's own
    Patch + Optimizer + vendored Ceres compiled unmodified (travels to the GPU box prebuilt)."""
    if kind == "ref":
        lib = load_ref()
        if lib is None:
            pytest.skip("oracle/_ref/libebt_ref.so is not built here")
        return lib
    return load_oracle()


def compare(stream, max_patches, n_frames=None, split=None, margin=None, stream_kernel=0, against="oracle"):
    cpu = CpuTracker(checker(against), stream.width, stream.height, max_patches)
    gpu = Tracker(stream.width, stream.height, max_patches)
    if margin is not None:
        gpu.set_prefilter_margin(margin)
    gpu.set_stream_kernel(stream_kernel)
    n_cpu = synth.run_stream(cpu, stream, n_frames)
    if split:
        # same stream, but every inter-frame batch pushed in `split` pieces:
        # batching must not change results
        class Split:
            def __getattr__(self, k):
                return getattr(gpu, k)

            def push_events(self, ev):
                n = 0
                for part in np.array_split(ev, split):
                    n += gpu.push_events(part)
                return n
        n_gpu = synth.run_stream(Split(), stream, n_frames)
    else:
        n_gpu = synth.run_stream(gpu, stream, n_frames)
    assert n_cpu == n_gpu and n_cpu > 0
    check_equal(cpu, gpu, n_cpu)
    cpu.close()
    gpu.close()


def check_equal(cpu, gpu, n_cpu):
    rc, rg = cpu.records(), gpu.records()
    assert len(rc) == len(rg) == n_cpu
    assert np.array_equal(rc["track_id"], rg["track_id"])
    assert np.array_equal(rc["event_index"], rg["event_index"])
    assert np.array_equal(rc["ts"], rg["ts"])
    assert np.array_equal(rc["num_of_events"], rg["num_of_events"])
    assert np.array_equal(rc["lost"] != 0, rg["lost"] != 0)
    dpos = np.hypot(rc["x"] - rg["x"], rc["y"] - rg["y"])
    dang = _angle_diff_deg(rc["flow_dir"], rg["flow_dir"])
    if (rc["n_successful"] >= 0).all():
        same = ((rc["n_successful"] == rg["n_successful"]) & (rc["n_unsuccessful"] == rg["n_unsuccessful"])
                & (rc["termination"] == rg["termination"]))
    else:
        same = np.ones(len(rc), bool)  # the reference build reports final costs only, not Ceres' iteration counts
    assert np.abs(rc["final_cost"] - rg["final_cost"]).max() < 1e-6
    print(f"\n{n_cpu} optimisations: max dpos {dpos.max():.3e} px, max dangle {dang.max():.3e} deg, "
          f"same LM path {same.mean():.5f}, rejected steps seen {int(rc['n_unsuccessful'].sum())}")
    assert dpos.max() <= POS_TOL_PX and dang.max() <= ANGLE_TOL_DEG
    assert same.mean() >= 0.99
    for arch in (False, True):
        tc, tg = cpu.tracks(arch), gpu.tracks(arch)
        assert np.array_equal(tc["track_id"], tg["track_id"])
        for k in ("lost", "init", "num_of_events", "window_size", "n_optimizations", "n_trajectory",
                  "current_ts", "time_last_update", "init_time"):
            assert np.array_equal(tc[k], tg[k]), k
        if (tc["counter"] >= 0).all():  # Patch::counter_ has no getter in the reference build
            assert np.array_equal(tc["counter"], tg["counter"])
        if len(tc):
            assert np.abs(tc["rect_x"] - tg["rect_x"]).max() <= POS_TOL_PX
        for tid in tc["track_id"]:
            assert np.array_equal(cpu.patch(tid), gpu.patch(tid)), f"integer patch of track {tid}"
            a, b = cpu.trajectory(tid), gpu.trajectory(tid)
            assert a.shape == b.shape and np.array_equal(a[:, 2], b[:, 2])
            assert np.abs(a[:, :2] - b[:, :2]).max() <= POS_TOL_PX


@pytest.mark.gpu
def test_stream_c1_shape():
    """BASELINE.json configs[0]: DAVIS 240x180, <= 100 patches."""
    S = synth.make_stream(240, 180, 0.30, 500e3, seed=1001, max_corners=69)
    compare(S, 100)


@pytest.mark.gpu
def test_stream_c1_shape_against_the_reference_build():
    """The same stream against oracle/_ref: the CUDA path meets the reference's own code, not only its port."""
    S = synth.make_stream(240, 180, 0.30, 500e3, seed=1001, max_corners=69)
    compare(S, 100, against="ref")


@pytest.mark.gpu
@pytest.mark.parametrize("stream_kernel", [0, 2])
def test_stream_c2_shape_against_the_reference_build(stream_kernel):
    S = synth.make_stream(346, 260, 0.10, 2e6, seed=1002, max_corners=500)
    compare(S, 500, stream_kernel=stream_kernel, against="ref")


@pytest.mark.gpu
def test_stream_kernels_are_deterministic():
    """Stand-in for racecheck (compute-sanitizer is closed on this pool): the hand-offs of the persistent-CTA
    kernel are sequence numbers in shared memory, no barriers — a race would show up as run-to-run
    differences.  The same c2 stream 12 times: bit-identical records every time."""
    S = synth.make_stream(346, 260, 0.10, 2e6, seed=1002, max_corners=500)
    first = None
    for _ in range(12):
        g = Tracker(S.width, S.height, 500)
        g.set_stream_kernel(2)
        synth.run_stream(g, S)
        r = g.records().copy()
        g.close()
        if first is None:
            first = r
            assert len(first) > 1000
        else:
            assert first.tobytes() == r.tobytes()


@pytest.mark.gpu
def test_stream_split_batches():
    S = synth.make_stream(240, 180, 0.15, 500e3, seed=1011, max_corners=69)
    compare(S, 100, split=7)


@pytest.mark.gpu
@pytest.mark.parametrize("stream_kernel", [0, 1, 2])
def test_stream_c2_shape(stream_kernel):
    """BASELINE.json configs[1] shape: 346x260, 500 features, ~2 Mev/s (short); both stream kernels
    (0 one CTA per feature, 2 persistent CTAs of two leader warps sharing four pixel warps, 1 the
    automatic choice between them)."""
    S = synth.make_stream(346, 260, 0.10, 2e6, seed=1002, max_corners=500)
    compare(S, 500, stream_kernel=stream_kernel)


@pytest.mark.gpu
def test_stream_shared_pixel_warps_split_and_margin():
    """The persistent-CTA kernel with every batch pushed in 5 pieces and a 0.5 px prefilter margin
    (mask rows rebuilt after nearly every optimisation)."""
    S = synth.make_stream(346, 260, 0.08, 2e6, seed=1052, max_corners=400)
    compare(S, 400, split=5, margin=0.5, stream_kernel=2)


@pytest.mark.gpu
def test_stream_shared_pixel_warps_few_features():
    """The persistent-CTA kernel forced onto a c1-sized stream: fewer work items than CTAs could hold,
    second leaders without a first feature, an empty queue."""
    S = synth.make_stream(240, 180, 0.15, 500e3, seed=1012, max_corners=30)
    compare(S, 100, stream_kernel=2)
    S = synth.make_stream(240, 180, 0.10, 300e3, seed=1013, max_corners=1)
    compare(S, 100, stream_kernel=2)


@pytest.mark.gpu
@pytest.mark.parametrize("one_with_products", [False, True])
def test_multi_stream_push_equals_the_streams_alone(one_with_products):
    """ebt_push_events_multi (configs[3]: independent streams in one launch): three streams of different
    seeds, feature counts and batch sizes — one of them runs out of events early — each equal to the
    CPU oracle's per-event loop on that stream alone.  With the getter-only patch products switched on in
    one context the call takes the one-CTA-per-feature kernel for every stream (the shared-pixel-warp
    kernel does not produce them) and must give the same tracks."""
    from event_based_odomety_b200.cabi import new_frame_end_multi, push_events_multi
    streams = [synth.make_stream(240, 180, 0.15, 500e3, seed=2000, max_corners=69),
               synth.make_stream(240, 180, 0.15, 300e3, seed=2001, max_corners=40),
               synth.make_stream(240, 180, 0.10, 600e3, seed=2002, max_corners=60)]
    caps = [100, 100, 60]
    cpus = [CpuTracker(load_oracle(), S.width, S.height, c) for S, c in zip(streams, caps)]
    gpus = [Tracker(S.width, S.height, c) for S, c in zip(streams, caps)]
    if one_with_products:
        gpus[1].set_products(1)
    n_cpu = [synth.run_stream(c, S) for c, S in zip(cpus, streams)]
    n_gpu = [0, 0, 0]
    n_frames = max(len(S.frames) for S in streams)
    for k in range(n_frames):
        batches = []
        for i, (g, S) in enumerate(zip(gpus, streams)):
            if k >= len(S.frames):
                batches.append(S.events[:0])  # this stream has ended: an empty batch in the same call
                continue
            g.new_frame_begin(S.frames[k], S.frame_ts[k], S.corners[k])
            ids, xy = g.get_uninit()
            if len(ids):
                nxt = S.lk_next(k, xy)
                for j, tid in enumerate(ids):
                    inside = (0 <= nxt[j, 0] < S.width) and (0 <= nxt[j, 1] < S.height)
                    g.set_flow(tid, int(inside), nxt[j, 0], nxt[j, 1])
            batches.append(S.events_between(k))
        new_frame_end_multi([g for g, S in zip(gpus, streams) if k < len(S.frames)])
        got = push_events_multi(gpus, batches)
        n_gpu = [a + b for a, b in zip(n_gpu, got)]
    assert n_gpu == n_cpu and min(n_cpu) > 0
    for c, g, n in zip(cpus, gpus, n_cpu):
        check_equal(c, g, n)
        c.close()
        g.close()


@pytest.mark.gpu
def test_multi_stream_push_rejects_mismatched_contexts():
    from event_based_odomety_b200.cabi import EbtError, push_events_multi
    a, b = Tracker(240, 180, 10), Tracker(346, 260, 10)
    ev = np.zeros(4, synth.EVENT)
    with pytest.raises(EbtError, match="must be equal"):
        push_events_multi([a, b], [ev, ev])
    with pytest.raises(EbtError, match="listed twice"):
        push_events_multi([a, a], [ev, ev])
    assert push_events_multi([a], [ev]) == [0]  # no patches yet: events consumed, nothing happens
    a.close()
    b.close()


@pytest.mark.gpu
def test_stream_two_features_per_sm_build():
    """149-296 live patches on a 148-SM device select the 2-CTAs-per-SM build of the stream kernel
    (<= 148 the 1-CTA build: the c1 tests above; more the 3-CTA build: the c2 / c3 tests)."""
    S = synth.make_stream(346, 260, 0.10, 2e6, seed=1042, max_corners=240)
    compare(S, 240)


@pytest.mark.gpu
@pytest.mark.parametrize("margin", [0.0, 0.5, 64.0])
def test_stream_prefilter_margin(margin):
    """The prefilter is an index, not a filter: results must not depend on how
    far a rect may drift before its candidate row is rebuilt.  0 rebuilds after
    nearly every optimisation, 64 px never does on this stream."""
    S = synth.make_stream(240, 180, 0.15, 500e3, seed=1021, max_corners=69)
    compare(S, 100, margin=margin)


@pytest.mark.gpu
def test_stream_c3_shape():
    """BASELINE.json configs[2] shape: 640x480, 2000 features, ~10 Mev/s (two
    frame intervals; the CPU oracle needs about a minute for them)."""
    S = synth.make_stream(640, 480, 0.09, 10e6, seed=1003, max_corners=2000)
    compare(S, 2000, n_frames=2)


@pytest.mark.gpu
def test_stream_c3_full_size_properties():
    """BASELINE.json configs[2] at the size bench.py runs it (0.3 s, ~3 M events, 2000 features, ~3e5
    optimisations) — too long for the CPU checkers inside a test, so properties that do not depend on the
    size: (i) the persistent shared-pixel-warp kernel and the one-CTA-per-feature kernel make the same
    optimisations at the same events and agree to 1e-9 px; (ii) cutting every batch in three pieces changes
    nothing, bit for bit (the kernels are sequential in a feature's events); (iii) every track's
    optimisations come at strictly increasing stream events inside the pushed range, on live tracks only."""
    S = synth.make_stream(640, 480, 0.30, 10e6, seed=1003, max_corners=2000)

    class Split:
        def __init__(self, g):
            self.g = g

        def __getattr__(self, k):
            return getattr(self.g, k)

        def push_events(self, ev):
            return sum(self.g.push_events(part) for part in np.array_split(ev, 3))

    def run(mode, split=False):
        g = Tracker(S.width, S.height, 2000, record_capacity=1 << 20)
        g.set_stream_kernel(mode)
        n = synth.run_stream(Split(g) if split else g, S)
        r = g.records().copy()
        live, arch = g.tracks(False).copy(), g.tracks(True).copy()
        g.close()
        assert n == len(r)
        return r, live, arch

    a, live_a, arch_a = run(2)
    b, live_b, arch_b = run(0)
    c, live_c, arch_c = run(2, split=True)
    assert len(a) > 200000
    # (ii)
    assert a.tobytes() == c.tobytes() and live_a.tobytes() == live_c.tobytes() and arch_a.tobytes() == arch_c.tobytes()
    # (i)
    for k in ("track_id", "event_index", "ts", "num_of_events", "lost", "n_successful", "n_unsuccessful", "termination"):
        assert np.array_equal(a[k], b[k]), k
    assert np.hypot(a["x"] - b["x"], a["y"] - b["y"]).max() < 1e-9
    assert _angle_diff_deg(a["flow_dir"], b["flow_dir"]).max() < 1e-7
    assert np.array_equal(live_a["track_id"], live_b["track_id"]) and np.array_equal(arch_a["track_id"], arch_b["track_id"])
    # (iii)
    assert a["event_index"].min() >= 0 and a["event_index"].max() < len(S.events)
    order = np.lexsort((np.arange(len(a)), a["track_id"]))
    t, e, lost = a["track_id"][order], a["event_index"][order], a["lost"][order]
    same = t[1:] == t[:-1]
    assert (e[1:][same] > e[:-1][same]).all()
    assert not lost[:-1][same].any()  # nothing is optimised after the record that lost it


@pytest.mark.gpu
def test_stream_empty_and_tiny_pushes():
    """Zero-length pushes, single-event pushes and a push before any frame."""
    S = synth.make_stream(240, 180, 0.10, 500e3, seed=1031, max_corners=40)
    gpu = Tracker(S.width, S.height, 100)
    assert gpu.push_events(S.events[:1000]) == 0  # no patches yet: events are consumed, nothing happens
    assert len(gpu.tracks()) == 0
    gpu.reset()
    cpu = CpuTracker(load_oracle(), S.width, S.height, 100)

    class Ragged:
        def __getattr__(self, k):
            return getattr(gpu, k)

        def push_events(self, ev):
            n = gpu.push_events(ev[:0])
            cuts = [0, 1, 2, 3, 130, 131, len(ev) // 2, len(ev)]
            for a, b in zip(cuts, cuts[1:]):
                n += gpu.push_events(ev[a:b])
            return n + gpu.push_events(ev[:0])
    n_cpu = synth.run_stream(cpu, S)
    n_gpu = synth.run_stream(Ragged(), S)
    assert n_cpu == n_gpu > 0
    rc, rg = cpu.records(), gpu.records()
    assert np.array_equal(rc["track_id"], rg["track_id"]) and np.array_equal(rc["event_index"], rg["event_index"])
    assert np.hypot(rc["x"] - rg["x"], rc["y"] - rg["y"]).max() <= POS_TOL_PX
    cpu.close()
    gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("against", ["oracle", "ref"])
def test_stream_more_than_64_live_template_frames(against):
    """FeatureDetector::optimizers_ is an unbounded map (one gradient grid per frame that still has a
    patch created at it).  150 frames, each creating one more long-lived patch, keep 150 template
    grids alive at once (the first build capped them at 64); then events make the oldest, a middle
    and the newest patches optimise against their own template frames."""
    W, H, N = 240, 180, 150
    rng = np.random.Generator(np.random.PCG64(404))
    base = synth.make_stream(W, H, 0.05, 100e3, seed=1404, max_corners=4).frames[0]
    spots = [(20.0 + 14 * (i % 15), 20.0 + 14 * (i // 15)) for i in range(N)]
    cpu = CpuTracker(checker(against), W, H, 200)
    gpu = Tracker(W, H, 200)
    gpu.set_stream_kernel(2)
    ts = 0
    for k in range(N):
        frame = np.roll(base, k, axis=1).copy()
        corners = np.array(spots[:k + 1], np.float64)
        for t in (cpu, gpu):
            t.new_frame_begin(frame, ts, corners)
            ids, xy = t.get_uninit()
            for tid, (x, y) in zip(ids, xy):
                t.set_flow(tid, 1, x + 0.75, y - 0.5)
            t.new_frame_end()
        ts += 44065
    tc, tg = cpu.tracks(), gpu.tracks()
    assert len(tc) == len(tg) == N and np.array_equal(tc["track_id"], tg["track_id"])
    assert not tc["lost"].any()
    # events on the oldest, a middle and the newest initialised patches
    ev = []
    for j in (0, 70, N - 2):
        cx, cy = spots[j]
        n = 400
        e = np.zeros(n, synth.EVENT)
        e["x"] = np.clip(np.rint(cx + rng.uniform(-12, 12, n)), 0, W - 1)
        e["y"] = np.clip(np.rint(cy + rng.uniform(-12, 12, n)), 0, H - 1)
        e["polarity"] = rng.choice(np.array([-1, 1], np.int8), n)
        ev.append(e)
    ev = np.concatenate(ev)
    ev = ev[rng.permutation(len(ev))]
    ev["t_us"] = ts + np.arange(len(ev))
    n_cpu, n_gpu = cpu.push_events(ev), gpu.push_events(ev)
    assert n_cpu == n_gpu and n_cpu >= 9
    check_equal(cpu, gpu, n_cpu)
    cpu.close()
    gpu.close()
