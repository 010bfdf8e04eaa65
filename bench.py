#!/usr/bin/env python
"""bench.py — feature-patch optimisations/s and events/s of the tracker hot path.

A "step" is one pass of the hot path over one synthetic stream segment: the
tracker is reset, then every frame (gradient grid, association, window-size
kernel) and every inter-frame event batch (stream kernel: event scan, window
integration, LM solves) is processed.  Workload at N=1 is BASELINE.json
configs[1]: DAVIS346-shaped stream, 346x260, up to 500 features, ~2 Mev/s.
At N>1 every rank tracks its own independent stream of the same shape
(configs[3], sharding by stream; no collective on the data path).

  value  : optimisations/s with frames and events already resident in HBM
           (ebt_new_frame_begin_device / ebt_push_events_device)
  e2e    : same metric through the host-buffer C-ABI calls (pinned host
           events, H2D inside the timed region, records/tracks D2H inside)
  roofline / cpu_baseline : see DESIGN.md "Measurement"

`--impl reference` times the reference's own CPU code for the same path
(oracle/_ref/libebt_ref.so: unmodified Patch + Optimizer + Ceres) on the host
cores, features sharded over threads, on a bounded sample of the same stream.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from event_based_odomety_b200 import multi, synth  # noqa: E402
from event_based_odomety_b200.records import EVENT, OPT_RECORD, TRACK  # noqa: E402

WORKLOADS = {
    # name: (width, height, max_features, target rate ev/s, base seed)
    "c1": (240, 180, 100, 0.5e6, 1001),
    "c2": (346, 260, 500, 2.0e6, 1002),
    "c3": (640, 480, 2000, 10.0e6, 1003),
}
METRIC = "feature-patch optimisations/sec"


def workload_name(key):
    w, h, feat, rate, _ = WORKLOADS[key]
    return f"{key}: {w}x{h} synthetic stream, <= {feat} features (25x25 patches), ~{rate / 1e6:g} Mev/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--duration", type=float, default=0.30, help="stream seconds per step")
    ap.add_argument("--stream-kernel", type=int, default=0, choices=[0, 1],
                    help="ebt_set_stream_kernel: 0 one CTA per feature, 1 two features per CTA sharing its pixel warps")
    ap.add_argument("--streams", type=int, default=1,
                    help="independent streams tracked concurrently per GPU (configs[3] throughput mode)")
    ap.add_argument("--cpu-frames", type=int, default=2, help="frames of the stream in the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class FlowCache:
    """The LK answers the host injects (flow_estimator.cpp:84-103) are a pure
    function of the stream; cache them so the timed loop does not run numpy."""

    def __init__(self, stream):
        self.S = stream
        self.cache = {}

    def get(self, k, ids, xy):
        key = (k, ids.tobytes())
        if key not in self.cache:
            nxt = self.S.lk_next(k, xy)
            found = ((nxt[:, 0] >= 0) & (nxt[:, 0] < self.S.width) & (nxt[:, 1] >= 0)
                     & (nxt[:, 1] < self.S.height)).astype(np.int32)
            self.cache[key] = (found, nxt)
        return self.cache[key]


def run_step(tr, S, flows, frame_fn, push_fn, n_frames):
    """Replayer/Evaluator order (replayer.cpp:56-98): image k, then the events
    before image k+1."""
    total = 0
    for k in range(n_frames):
        frame_fn(k)
        ids, xy = tr.get_uninit()
        if len(ids):
            found, nxt = flows.get(k, ids, xy)
            if hasattr(tr, "set_flows"):
                tr.set_flows(ids, found, nxt)
            else:
                for i in range(len(ids)):
                    tr.set_flow(ids[i], found[i], nxt[i, 0], nxt[i, 1])
        tr.new_frame_end()
        total += push_fn(k)
    return total


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.stop_evt = threading.Event()
        self.sm, self.mx, self.reasons = [], 0, set()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip().split(",")
                self.sm.append(float(out[0]))
                self.mx = max(self.mx, float(out[1]))
                for n, v in zip(names, out[2:]):
                    if "Active" in v and "Not" not in v:
                        self.reasons.add(n)
            except Exception:
                pass
            self.stop_evt.wait(0.2)

    def result(self):
        self.stop_evt.set()
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons)}


def lm_flops(rec):
    """625 * [400 * jacobian evals + 230 * cost-only evals] per solve."""
    ns = rec["n_successful"].astype(np.float64)
    nu = rec["n_unsuccessful"].astype(np.float64)
    return float((625.0 * (400.0 * (1 + ns) + 230.0 * (ns + nu))).sum())


# ---------------------------------------------------------------------------
# CPU arms (the only place bench.py touches oracle/)
# ---------------------------------------------------------------------------
def cpu_lib():
    from tests._oracle import load_oracle, load_ref
    ref = load_ref()
    return (ref, "reference") if ref is not None else (load_oracle(), "port")


def cpu_run(S, max_feat, n_frames, threads):
    """Reference CPU tracker on the first n_frames frames (+ their event
    batches).  threads > 1: the path shards by independent stream,
    so every host thread tracks its own copy of the stream sample with its own
    tracker instance (ctypes releases the GIL); units = threads x optimisations."""
    from tests._oracle import CpuTracker
    lib, kind = cpu_lib()
    flows = FlowCache(S)
    counts = [0] * threads

    def work(t):
        tr = CpuTracker(lib, S.width, S.height, max_feat)

        def frame_fn(k):
            tr.new_frame_begin(S.frames[k], S.frame_ts[k], S.corners[k])
        counts[t] = run_step(tr, S, flows, frame_fn, lambda k: tr.push_events(S.events_between(k)), n_frames)
        tr.close()

    t0 = time.perf_counter()
    if threads == 1:
        work(0)
    else:
        th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
        [x.start() for x in th]
        [x.join() for x in th]
    dt = time.perf_counter() - t0
    n_ev = sum(len(S.events_between(k)) for k in range(n_frames))
    return sum(counts), n_ev * threads, dt, kind


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w, h, feat, rate, seed = WORKLOADS[args.workload]
    S = synth.make_stream(w, h, args.duration, rate, seed, max_corners=feat)  # same stream as our arm
    args.cpu_frames = min(args.cpu_frames, len(S.frames))
    cores = os.cpu_count() or 1
    for _ in range(args.warmup if args.warmup < 2 else 1):  # CPU needs no clock warm-up; one pass
        cpu_run(S, feat, min(2, args.cpu_frames), cores)
    times, opts, nev = [], 0, 0
    for _ in range(args.steps):
        n, ne, dt, kind = cpu_run(S, feat, args.cpu_frames, cores)
        times.append(dt)
        opts, nev = n, ne
    T = float(np.sum(times))
    value = opts * args.steps / T
    sample = (f"first {args.cpu_frames} frames + event batches of the {args.workload} stream "
              f"({nev} events, {opts} optimisations per step summed over {cores} host threads, "
              f"one independent copy of the stream sample per thread)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "optimisations/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": T / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args.workload), "sample": sample},
        "events_per_s": nev * args.steps / T,
        "cpu_baseline": {"value": value, "unit": "optimisations/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "optimisations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------
# Our arm
# ---------------------------------------------------------------------------
class Lane:
    """One event stream of this rank: its tracker context (own CUDA stream), its
    frames and events resident in HBM and in pinned host memory."""

    def __init__(self, torch, Tracker, wl, duration, seed, device, stream_kernel=0):
        w, h, feat, rate, _ = wl
        self.S = S = synth.make_stream(w, h, duration, rate, seed, max_corners=feat)
        self.nF = len(S.frames)
        t = S.events["t_us"]
        self.bounds = [(int(np.searchsorted(t, S.frame_ts[k], "left")),
                        int(np.searchsorted(t, S.frame_ts[k + 1], "left")) if k + 1 < self.nF else len(S.events))
                       for k in range(self.nF)]
        self.n_events = self.bounds[-1][1] - self.bounds[0][0]
        self.tr = Tracker(w, h, feat, device=device, record_capacity=1 << 20)
        self.tr.set_stream_kernel(stream_kernel)
        self.flows = FlowCache(S)
        ev_bytes = torch.from_numpy(S.events.view(np.uint8).reshape(-1).copy())
        self.d_events = ev_bytes.cuda()
        self.h_events = ev_bytes.pin_memory()
        self.d_frames = [torch.from_numpy(f.copy()).cuda() for f in S.frames]

    def frame_dev(self, k):
        self.tr.new_frame_begin_device(self.d_frames[k].data_ptr(), self.S.frame_ts[k], self.S.corners[k])

    def frame_host(self, k):
        self.tr.new_frame_begin(self.S.frames[k], self.S.frame_ts[k], self.S.corners[k])

    def push_dev(self, k):
        lo, hi = self.bounds[k]
        return self.tr.push_events_device(self.d_events.data_ptr() + lo * 16, hi - lo)

    def push_host(self, k):
        lo, hi = self.bounds[k]
        return self.tr.push_events_ptr(self.h_events.data_ptr() + lo * 16, hi - lo)

    def step(self, host):
        self.tr.reset()
        return run_step(self.tr, self.S, self.flows, self.frame_host if host else self.frame_dev,
                        self.push_host if host else self.push_dev, self.nF)


def main_ours(args):
    import torch
    import torch.distributed as dist
    from event_based_odomety_b200.cabi import K_GRAD, K_NUMEV, K_TRACK, Tracker

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    wl = WORKLOADS[args.workload]
    w, h, feat, rate, seed = wl
    # global stream g = rank * streams + lane; stream 0 is the named stream, every
    # further one an independent stream of the same shape (configs[3], seeds 2000 + g)
    nS = max(1, args.streams)
    lanes = [Lane(torch, Tracker, wl, args.duration, multi.stream_seed(seed, rank * nS + s), local, args.stream_kernel)
             for s in range(nS)]
    L0 = lanes[0]
    S, nF, tr = L0.S, L0.nF, L0.tr
    n_events = sum(L.n_events for L in lanes)
    fp64_peak, fp32_peak = tr.measure_fma_peak()
    # L2 flush between steps: 256 MB (> the 126 MB L2) rewritten at the start of
    # every step, inside the timed region (~0.05 ms)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def all_lanes(host):
        """One step of every stream of this rank; streams run concurrently, one
        host thread and one CUDA stream each (ctypes releases the GIL)."""
        if nS == 1:
            return L0.step(host)
        out = [0] * nS

        def work(i):
            out[i] = lanes[i].step(host)
        th = [threading.Thread(target=work, args=(i,)) for i in range(nS)]
        [x.start() for x in th]
        [x.join() for x in th]
        return sum(out)

    def timed(host, steps, with_timing):
        """K steps bracketed by barrier + synchronize; device time by CUDA events
        (the library syncs its own stream inside every call, so the events on
        the current stream bracket all of its work)."""
        barrier()
        if with_timing and nS == 1:
            # per-kernel CUDA events on the library stream; with several streams per
            # GPU the kernels overlap and the events would serialise them, so the
            # launch counts are then taken from one extra untimed step below
            tr.timing(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        opts = 0
        for _ in range(steps):
            flush.zero_()
            torch.cuda.current_stream().synchronize()
            opts += all_lanes(host)
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms = max(e0.elapsed_time(e1), 0.0)
        return opts, max(ms * 1e-3, wall if ms <= 0 else 0.0), wall

    for _ in range(max(args.warmup, 0)):
        flush.zero_()  # also loads torch's fill kernel outside the timed region
        all_lanes(False)
    sampler = ClockSampler(local)
    sampler.start()
    opts, secs, wall = timed(False, args.steps, True)
    kt_ms = kt_n = kg_ms = kg_n = kn_ms = kn_n = 0
    for L in lanes:
        a, b = L.tr.timing_get(K_TRACK); kt_ms += a; kt_n += b
        a, b = L.tr.timing_get(K_GRAD); kg_ms += a; kg_n += b
        a, b = L.tr.timing_get(K_NUMEV); kn_ms += a; kn_n += b
    flops_step = sum(lm_flops(L.tr.records()) for L in lanes)  # records of the last step
    if nS > 1:
        for L in lanes:
            L.tr.timing(True)
        all_lanes(False)
        kt_n, kg_n, kn_n = (args.steps * sum(L.tr.timing_get(k)[1] for L in lanes) for k in (K_TRACK, K_GRAD, K_NUMEV))
        kg_ms, kn_ms = (args.steps * sum(L.tr.timing_get(k)[0] for L in lanes) for k in (K_GRAD, K_NUMEV))
    for L in lanes:
        L.tr.timing(False)
    all_lanes(True)  # e2e warm-up
    e_opts, e_secs, _ = timed(True, args.steps, False)
    clocks = sampler.result()
    n_tracks = sum(len(L.tr.tracks()) + len(L.tr.tracks(True)) for L in lanes)

    # whole-job aggregate: sum of units over ranks / max time over ranks
    def agg(units, seconds):
        return multi.aggregate(units, seconds, device="cuda")

    g_opts, g_secs = agg(opts, secs)
    g_events, _ = agg(n_events * args.steps, secs)
    ge_opts, ge_secs = agg(e_opts, e_secs)

    if rank == 0:
        opts_step = opts / args.steps
        # stream-kernel time per step; with several streams per GPU their kernels
        # overlap, so the kernel's time is the step itself
        kt_s = kt_ms * 1e-3 / args.steps if nS == 1 else secs / args.steps
        achieved = flops_step / kt_s / 1e12 if kt_s > 0 else 0.0
        hbm_bytes = 16.0 * n_events + (2500.0 + 128.0) * opts_step
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        traffic = None  # DRAM bytes of one stream-kernel launch, from the committed ncu --set full capture
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r1_kernel_traffic.json")))[
                "track_batch_kernel"]["dram_bytes_per_launch"]
        except Exception:
            pass
        launches_step = kt_n / args.steps
        if world > 1 or nS > 1:
            sharding = (f"{nS} independent stream(s) per GPU (one context + CUDA stream + host thread each), "
                        f"{world} GPU(s), no collective")
        else:
            sharding = "single stream"
        line = {
            "metric": METRIC, "value": g_opts / g_secs, "unit": "optimisations/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": g_secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": workload_name(args.workload),
                "segment": f"{args.duration:g} s = {nF} frames, {n_events} events, {int(opts_step)} optimisations, "
                           f"{n_tracks} tracks per step" + (f" over {nS} streams" if nS > 1 else ""),
                "sharding": sharding,
                "l2": "flushed: a 256 MB buffer (> 126 MB L2) is rewritten at the start of every step, inside the "
                      "timed region; a step then reads its event stream (%.1f MB) and template grids from HBM once, "
                      "the kernel is FP64-pipe bound (see roofline)" % (n_events * 16 / 1e6),
            },
            "events_per_s": g_events / g_secs,
            "e2e": {"value": ge_opts / ge_secs, "unit": "optimisations/s",
                    "h2d_bytes_per_step": int(n_events * 16 + sum(
                        L.nF * w * h + 8 * sum(len(c) * 2 for c in L.S.corners) for L in lanes)),
                    "d2h_bytes_per_step": int(opts_step * OPT_RECORD.itemsize + 2 * nF * feat * 288 * nS)},
            # every push launches prefilter_kernel + track_batch_kernel
            "gpu_launches": int(2 * kt_n + kg_n + kn_n),
            "roofline": {"bound": "fp64",
                         "kernel": "track_batch_kernel (timed with its prefilter_kernel, 1.4 % of it)" if nS == 1 else
                         f"track_batch_kernel of {nS} concurrent streams (kernels overlap: time = the step)",
                         "achieved": achieved,
                         "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved / fp64_peak if fp64_peak else None,
                         "traffic": traffic, "traffic_unit": "DRAM bytes per launch (ncu --set full, profiles/r1_track_batch_kernel_ncu_metrics.json)",
                         "launches_per_step": launches_step,
                         "algorithmic_flops_per_launch": flops_step / launches_step if launches_step else None,
                         "kernel_ms_per_launch": kt_s * 1e3 / launches_step if launches_step else None,
                         "peak_source": "measured in this run: dependent-chain DFMA microbenchmark "
                         "(ebt_measure_fma_peak); FP32 non-tensor peak measured %.1f TFLOP/s" % fp32_peak,
                         "algorithmic_flops_per_step": flops_step, "kernel_ms_per_step": kt_s * 1e3,
                         "kernel_share_of_step": kt_s / (secs / args.steps)},
            "roofline_hbm": {"bound": "hbm", "kernel": "track_batch_kernel (event scan + integration side)",
                             "achieved": hbm_bytes / kt_s / 1e9 if kt_s > 0 else 0.0, "peak": hbm_peak,
                             "unit": "GB/s", "frac": hbm_bytes / kt_s / 1e9 / hbm_peak if kt_s > 0 else None,
                             "algorithmic_bytes_per_step": hbm_bytes},
            "other_kernels_ms_per_step": {"grad_kernel": kg_ms / args.steps, "num_events_kernel": kn_ms / args.steps},
            "clocks": clocks,
            "wall_s_timed_region": wall,
        }
        if not args.no_cpu_baseline:
            cf = min(args.cpu_frames, nF)
            n, ne, dt, kind = cpu_run(S, feat, cf, 1)
            line["cpu_baseline"] = {
                "value": n / dt, "unit": "optimisations/s", "cores": 1, "kind": kind,
                "events_per_s": ne / dt,
                "sample": f"first {cf} frames + event batches of the same stream ({ne} events, {n} optimisations, "
                          f"{dt:.1f} s), single thread as the reference runs it"}
        emit(line)
    for L in lanes:
        L.tr.close()
    if world > 1:
        dist.destroy_process_group()


def emit(line):
    """The JSON line is the only thing written to the real stdout."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


if __name__ == "__main__":
    # libraries (NCCL's version banner, torchrun notices) must not share stdout with the
    # one JSON line the contract asks for: everything else goes to stderr
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_ours(a)
