"""Scratch probe: per-CTA load of the stream kernel (ebt_debug_feature_load).  For every push of one
c2/c3 step: launch span, heaviest feature's CTA span and optimisation count, mean CTA span, the
balanced bound (sum of CTA spans / resident slots) and ns per optimisation for heavy and light CTAs."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from event_based_odomety_b200 import synth
from event_based_odomety_b200.cabi import Tracker
import bench

WL = sys.argv[1] if len(sys.argv) > 1 else "c2"
w, h, feat, rate, seed = bench.WORKLOADS[WL]
S = synth.make_stream(w, h, 0.30, rate, seed, max_corners=feat)
nF = len(S.frames)
t = S.events["t_us"]
bounds = [(int(np.searchsorted(t, S.frame_ts[k], "left")),
           int(np.searchsorted(t, S.frame_ts[k + 1], "left")) if k + 1 < nF else len(S.events)) for k in range(nF)]
tr = Tracker(w, h, feat, record_capacity=1 << 20)
MODE = int(sys.argv[2]) if len(sys.argv) > 2 else 2
tr.set_stream_kernel(MODE)
SLOTS = 148 * 3 if MODE == 0 else 148 * 4  # resident feature slots: 3 CTAs/SM, or 2 CTAs x 2 leaders
flows = bench.FlowCache(S)
d_events = torch.from_numpy(S.events.view(np.uint8).reshape(-1).copy()).cuda()
d_frames = [torch.from_numpy(f.copy()).cuda() for f in S.frames]
rows = []
raw = []
pre = []


def step(collect):
    tr.reset()
    for k in range(nF):
        tr.new_frame_begin_device(d_frames[k].data_ptr(), S.frame_ts[k], S.corners[k])
        ids, xy = tr.get_uninit()
        if len(ids):
            found, nxt = flows.get(k, ids, xy)
            tr.set_flows(ids, found, nxt)
        tr.new_frame_end()
        lo, hi = bounds[k]
        before = tr.tracks() if collect else None
        tr.push_events_device(d_events.data_ptr() + lo * 16, hi - lo)
        if collect:
            pre.append(before)
            ids, nopt, t0, t1, sm = tr.feature_load()
            ran = t1 > 0
            raw.append(dict(push=k, ids=ids.copy(), nopt=nopt.copy(), t0=t0.copy(), t1=t1.copy(), sm=sm.copy()))
            rows.append((k, nopt[ran].copy(), t0[ran].astype(np.int64), t1[ran].astype(np.int64), sm[ran].copy()))


for _ in range(3):
    step(False)
step(True)
out = []
for k, nopt, t0, t1, sm in rows:
    if not len(nopt):
        continue
    span = (t1.max() - t0.min()) / 1e3
    dur = (t1 - t0) / 1e3
    heavy = np.argmax(dur)
    order = np.argsort(-nopt)
    top = order[:max(1, len(order) // 10)]
    bot = order[len(order) // 2:]
    per_sm = np.bincount(sm, minlength=148)
    late = t0 > t0.min() + 100_000  # CTAs that had to wait for a free slot
    r = dict(push=k, ctas=int(len(nopt)), opts=int(nopt.sum()), span_us=float(span), mean_cta_us=float(dur.mean()),
             max_cta_us=float(dur.max()), max_cta_opts=int(nopt[heavy]), max_opts=int(nopt.max()),
             mean_opts=float(nopt.mean()), balanced_bound_us=float(dur.sum() / SLOTS), busy_fraction=float(dur.sum() / (min(SLOTS, len(dur)) * span)),
             opts_p50=float(np.median(nopt)), opts_p90=float(np.percentile(nopt, 90)), opts_p99=float(np.percentile(nopt, 99)),
             us_per_opt_top10pct=float(dur[top].sum() / max(nopt[top].sum(), 1)),
             us_per_opt_bottom_half=float(dur[bot].sum() / max(nopt[bot].sum(), 1)),
             late_start_us=float((t0.max() - t0.min()) / 1e3), ctas_per_sm_max=int(per_sm.max()),
             late_ctas=int(late.sum()), late_mean_opts=float(nopt[late].mean()) if late.any() else 0.0,
             late_mean_us=float(dur[late].mean()) if late.any() else 0.0)
    out.append(r)
    print(json.dumps(r))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open(f"gpurun_out/load_profile_{WL}.json", "w"), indent=1)
np.savez(f"gpurun_out/load_raw_{WL}.npz", **{f"{key}{r['push']}": r[key] for r in raw for key in ("ids", "nopt", "t0", "t1", "sm")})
np.savez(f"gpurun_out/load_pre_{WL}.npz", **{f"t{i}": b for i, b in enumerate(pre)})
# how well last push's count predicts this push's (the work order is built from it)
for (k0, n0, *_), (k1, n1, *_) in zip(rows[1:-1], rows[2:]):
    if len(n0) == len(n1):
        print(f"push {k1}: corr(prev, this) = {np.corrcoef(n0, n1)[0, 1]:.3f}, mean |diff| = {np.abs(n1 - n0).mean():.2f}")
