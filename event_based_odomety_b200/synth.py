"""Seeded synthetic inputs for the tracker hot path.

Two generators, both host-side numpy, both deterministic for a given seed:

* :func:`make_problems` — F independent feature-patch problems (the kernel
  micro-benchmark and the per-call parity cases): a band-limited random frame,
  a true SE(2) warp + flow angle per feature, an event window drawn from the
  brightness-increment model, and a perturbed initial state.
* :func:`make_stream` — a shapes_6dof-format stream (frames at 22.7 Hz and
  events ``(t_us, x, y, p)``) from a textured plane under smooth camera-plane
  motion, with per-frame corners and the "LK" flow seeds that the reference
  obtains from OpenCV (``goodFeaturesToTrack`` / ``calcOpticalFlowPyrLK``,
  feature_detector.cpp:568-583, flow_estimator.cpp:84-103) and that every
  implementation under test receives identically.
"""
from __future__ import annotations

import dataclasses
import os

import numpy as np

from .records import EVENT, PATCH_EXTENT, PATCH_SIDE, PROBLEM

FRAME_PERIOD_US = 44065  # tools/dataset_reader/test/test_data/images.txt spacing


def _gauss_blur(img: np.ndarray, sigma: float) -> np.ndarray:
    r = int(3 * sigma + 0.5)
    k = np.exp(-0.5 * (np.arange(-r, r + 1) / sigma) ** 2)
    k /= k.sum()
    pad = np.pad(img, r, mode="reflect")
    tmp = np.zeros((pad.shape[0], img.shape[1]))
    for i, kv in enumerate(k):
        tmp += kv * pad[:, i:i + img.shape[1]]
    out = np.zeros_like(img, dtype=np.float64)
    for i, kv in enumerate(k):
        out += kv * tmp[i:i + img.shape[0], :]
    return out


def make_texture(w: int, h: int, seed: int, sigma: float = 2.0) -> np.ndarray:
    """Band-limited random uint8 frame (Gaussian-blurred white noise)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    img = _gauss_blur(rng.standard_normal((h, w)), sigma)
    img = (img - img.min()) / (img.max() - img.min())
    return np.clip(20 + 215 * img, 0, 255).astype(np.uint8)


def log_sobel(gray: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """log(I/255+0.1)/8 -> Sobel 3x3, reflect-101 (feature_detector.cpp:714-732).

    numpy restatement used only to *generate* data; the checked versions live
    in oracle/ and in the CUDA kernel."""
    L = np.log(gray.astype(np.float64) * (1.0 / 255.0) + 10e-2) * (1.0 / 8)
    P = np.pad(L, 1, mode="reflect")
    dif = P[:, 2:] - P[:, :-2]
    smo = P[:, :-2] + 2 * P[:, 1:-1] + P[:, 2:]
    gx = dif[:-2] + 2 * dif[1:-1] + dif[2:]
    gy = smo[2:] - smo[:-2]
    return gx, gy


def _bilinear(img: np.ndarray, x: np.ndarray, y: np.ndarray) -> np.ndarray:
    h, w = img.shape
    x = np.clip(x, 0, w - 1.001)
    y = np.clip(y, 0, h - 1.001)
    x0 = np.floor(x).astype(np.int64)
    y0 = np.floor(y).astype(np.int64)
    fx = x - x0
    fy = y - y0
    return ((1 - fy) * ((1 - fx) * img[y0, x0] + fx * img[y0, x0 + 1])
            + fy * ((1 - fx) * img[y0 + 1, x0] + fx * img[y0 + 1, x0 + 1]))


@dataclasses.dataclass
class ProblemSet:
    width: int
    height: int
    frames: list            # uint8 (h, w) template frames, slot order
    problems: np.ndarray    # PROBLEM records
    events: np.ndarray      # EVENT records, windows newest-first
    true_warp: np.ndarray   # (F, 4) cos sin tx ty
    true_flow: np.ndarray   # (F,)


def make_problems(n: int, width: int = 240, height: int = 180, seed: int = 3000,
                  n_frames: int = 1, min_events: int = 100, max_events: int = 300,
                  border: int = 30) -> ProblemSet:
    """F independent patch problems (SURVEY 8d "kernel micro-benchmarks")."""
    rng = np.random.Generator(np.random.PCG64(seed))
    frames = [make_texture(width, height, seed * 7 + 13 * s + 1) for s in range(n_frames)]
    grads = [log_sobel(f) for f in frames]
    problems = np.zeros(n, PROBLEM)
    true_warp = np.zeros((n, 4))
    true_flow = np.zeros(n)
    n_ev = rng.integers(min_events, max_events + 1, size=n)
    offsets = np.concatenate([[0], np.cumsum(n_ev)])
    events = np.zeros(int(offsets[-1]), EVENT)
    ii, jj = np.meshgrid(np.arange(PATCH_SIDE), np.arange(PATCH_SIDE))  # ii = x idx, jj = y idx
    for k in range(n):
        slot = int(rng.integers(0, n_frames))
        gx, gy = grads[slot]
        ix = rng.uniform(border, width - border)
        iy = rng.uniform(border, height - border)
        th_t = np.deg2rad(rng.uniform(-3, 3))
        # warp maps current-frame pixels to template pixels: W = R p + t, and the
        # patch centre is warp^-1 * initPoint.  Choose the true centre offset.
        off = rng.uniform(-3, 3, size=2)
        c_t, s_t = np.cos(th_t), np.sin(th_t)
        ctr_t = np.array([ix, iy]) + off
        t_t = np.array([ix, iy]) - np.array([c_t * ctr_t[0] - s_t * ctr_t[1],
                                             s_t * ctr_t[0] + c_t * ctr_t[1]])
        phi_t = rng.uniform(-np.pi, np.pi)
        # initial (perturbed) state
        th_0 = th_t + np.deg2rad(rng.uniform(-1.0, 1.0))
        ctr_0 = ctr_t + rng.uniform(-1.5, 1.5, size=2)
        c_0, s_0 = np.cos(th_0), np.sin(th_0)
        t_0 = np.array([ix, iy]) - np.array([c_0 * ctr_0[0] - s_0 * ctr_0[1],
                                             s_0 * ctr_0[0] + c_0 * ctr_0[1]])
        phi_0 = phi_t + np.deg2rad(rng.normal(0, 10))
        # the rect the reference derives from the initial warp (patch.cpp:49-63)
        ncx = c_0 * ix + s_0 * iy + (-(c_0 * t_0[0] + s_0 * t_0[1]))
        ncy = -s_0 * ix + c_0 * iy + (s_0 * t_0[0] - c_0 * t_0[1])
        rx, ry = ncx - PATCH_EXTENT, ncy - PATCH_EXTENT
        ex = np.ceil(rx) + ii   # integer event coordinates inside the rect
        ey = np.ceil(ry) + jj
        # brightness increment model at the TRUE warp/flow
        wx = c_t * ex - s_t * ey + t_t[0]
        wy = s_t * ex + c_t * ey + t_t[1]
        u = _bilinear(gx, wx, wy) * np.cos(phi_t) + _bilinear(gy, wx, wy) * np.sin(phi_t)
        pr = np.abs(u).ravel()
        pr = pr / pr.sum() if pr.sum() > 0 else np.full(pr.size, 1.0 / pr.size)
        m = int(n_ev[k])
        pick = rng.choice(pr.size, size=m, p=pr)
        pol = -np.sign(u.ravel()[pick])
        pol[pol == 0] = 1
        flip = rng.random(m) < 0.10
        pol[flip] *= -1
        sl = slice(int(offsets[k]), int(offsets[k + 1]))
        events["x"][sl] = np.clip(ex.ravel()[pick], 0, width - 1).astype(np.uint16)
        events["y"][sl] = np.clip(ey.ravel()[pick], 0, height - 1).astype(np.uint16)
        events["polarity"][sl] = pol.astype(np.int8)
        events["t_us"][sl] = 1_000_000 + 10 * (m - np.arange(m))  # newest first
        p = problems[k]
        p["init_x"], p["init_y"] = ix, iy
        p["warp"] = (c_0, s_0, t_0[0], t_0[1])
        p["flow_dir"] = phi_0
        p["n_events"] = m
        p["event_offset"] = int(offsets[k])
        p["frame_slot"] = slot
        true_warp[k] = (c_t, s_t, t_t[0], t_t[1])
        true_flow[k] = phi_t
    return ProblemSet(width, height, frames, problems, events, true_warp, true_flow)


# ---------------------------------------------------------------------------
# Stream generator: frames + events + corner / LK-flow seeds
# ---------------------------------------------------------------------------
def harris_corners(gray: np.ndarray, max_corners: int, quality: float = 0.01,
                   min_distance: float = 10.0, inset: int = PATCH_EXTENT) -> np.ndarray:
    """Harris response + greedy min-distance selection, in the spirit of
    cv::goodFeaturesToTrack(..., useHarris, k=0.04) with the reference's mask
    inset (feature_detector.cpp:34-43, 568-583).  It only has to produce
    plausible corner seeds: the same seeds are injected into every
    implementation under test."""
    I = gray.astype(np.float64)
    P = np.pad(I, 1, mode="reflect")
    ix = (P[1:-1, 2:] - P[1:-1, :-2]) * 0.5
    iy = (P[2:, 1:-1] - P[:-2, 1:-1]) * 0.5

    def box3(a):
        Q = np.pad(a, 1, mode="reflect")
        return sum(Q[i:i + a.shape[0], j:j + a.shape[1]] for i in range(3) for j in range(3))

    sxx, syy, sxy = box3(ix * ix), box3(iy * iy), box3(ix * iy)
    R = sxx * syy - sxy * sxy - 0.04 * (sxx + syy) ** 2
    h, w = R.shape
    R[:inset, :] = R[h - inset:, :] = -np.inf
    R[:, :inset] = R[:, w - inset:] = -np.inf
    Q = np.pad(R, 1, mode="constant", constant_values=-np.inf)
    nb = np.max([Q[i:i + h, j:j + w] for i in range(3) for j in range(3) if (i, j) != (1, 1)], axis=0)
    cand = np.argwhere((R > nb) & (R > quality * R[np.isfinite(R)].max()))
    order = np.argsort(-R[cand[:, 0], cand[:, 1]], kind="stable")
    cell = max(int(min_distance), 1)
    grid = {}
    out = []
    for y, x in cand[order]:
        gy, gx = y // cell, x // cell
        ok = True
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                for (py, px) in grid.get((gy + dy, gx + dx), ()):
                    if (py - y) ** 2 + (px - x) ** 2 < min_distance ** 2:
                        ok = False
        if ok:
            grid.setdefault((gy, gx), []).append((y, x))
            out.append((float(x), float(y)))
            if len(out) >= max_corners:
                break
    return np.array(out, np.float64).reshape(-1, 2)


@dataclasses.dataclass
class Pose:
    """Camera-plane motion of the generator: t_s -> (theta, offx, offy).  A class, not
    a closure, so that a Stream can be pickled (streams generated by a process pool)."""
    amplitude: float
    theta_max: float

    def __call__(self, t):
        f1, f2, f3 = 0.5, 0.35, 0.25
        return (self.theta_max * np.sin(2 * np.pi * f3 * t), self.amplitude * np.sin(2 * np.pi * f1 * t),
                self.amplitude * (1 - np.cos(2 * np.pi * f2 * t)))


def make_streams(specs, processes=None):
    """make_stream(*spec) for every spec = (width, height, duration_s, rate, seed, max_corners),
    generated by worker processes (`python -m event_based_odomety_b200.synth`, no CUDA in them)."""
    import subprocess
    import sys
    import tempfile
    specs = [tuple(sp) for sp in specs]
    processes = max(1, min(processes or (os.cpu_count() or 1), len(specs)))
    if processes == 1 or len(specs) == 1:
        return [make_stream(*sp) for sp in specs]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = [None] * len(specs)
    with tempfile.TemporaryDirectory(prefix="ebt_streams_") as tmp:
        pending = list(enumerate(specs))
        running = []
        while pending or running:
            while pending and len(running) < processes:
                i, sp = pending.pop(0)
                path = os.path.join(tmp, f"s{i}.npz")
                cmd = [sys.executable, "-m", "event_based_odomety_b200.synth", path] + [repr(v) for v in sp]
                running.append((i, path, subprocess.Popen(cmd, cwd=root, stdout=subprocess.DEVNULL)))
            i, path, proc = running.pop(0)
            if proc.wait() != 0:
                raise RuntimeError(f"stream generator failed for spec {specs[i]}")
            out[i] = load_stream(path)
    return out


def save_stream(S, path):
    np.savez(path, width=S.width, height=S.height, frames=np.stack(S.frames), frame_ts=np.array(S.frame_ts, np.int64),
             events=S.events, corners=np.concatenate(S.corners) if S.corners else np.zeros((0, 2)),
             corner_counts=np.array([len(c) for c in S.corners], np.int64),
             pose=np.array([S.pose.amplitude, S.pose.theta_max]), margin=S.margin)


def load_stream(path):
    z = np.load(path)
    cuts = np.cumsum(z["corner_counts"])[:-1]
    return Stream(int(z["width"]), int(z["height"]), list(z["frames"]), [int(t) for t in z["frame_ts"]],
                  z["events"].view(EVENT) if z["events"].dtype != EVENT else z["events"],
                  [np.ascontiguousarray(c) for c in np.split(z["corners"], cuts)],
                  Pose(float(z["pose"][0]), float(z["pose"][1])), int(z["margin"]))


@dataclasses.dataclass
class Stream:
    width: int
    height: int
    frames: list          # uint8 (h, w)
    frame_ts: list        # int microseconds
    events: np.ndarray    # EVENT records sorted by time
    corners: list         # per frame (n, 2) float64 corner seeds
    pose: object          # callable t_s -> (theta, offx, offy)
    margin: int

    def to_texture(self, t_s, xy):
        th, ox, oy = self.pose(t_s)
        c, s = np.cos(th), np.sin(th)
        cx, cy = (self.width - 1) / 2.0, (self.height - 1) / 2.0
        x, y = xy[:, 0] - cx, xy[:, 1] - cy
        return np.stack([c * x - s * y + cx + self.margin + ox,
                         s * x + c * y + cy + self.margin + oy], axis=1)

    def from_texture(self, t_s, q):
        th, ox, oy = self.pose(t_s)
        c, s = np.cos(th), np.sin(th)
        cx, cy = (self.width - 1) / 2.0, (self.height - 1) / 2.0
        x, y = q[:, 0] - cx - self.margin - ox, q[:, 1] - cy - self.margin - oy
        return np.stack([c * x + s * y + cx, -s * x + c * y + cy], axis=1)

    def lk_next(self, k: int, xy: np.ndarray) -> np.ndarray:
        """What calcOpticalFlowPyrLK(prev=frame k-1, cur=frame k) would return for
        points xy (ground-truth image motion), as float32 like cv::Point2f."""
        q = self.to_texture(self.frame_ts[k - 1] * 1e-6, np.asarray(xy, np.float64).reshape(-1, 2))
        return self.from_texture(self.frame_ts[k] * 1e-6, q).astype(np.float32)

    def events_between(self, k: int) -> np.ndarray:
        """Events with frame_ts[k] <= t < frame_ts[k+1] (Replayer merge rule:
        an event goes first when t_event < t_image, replayer.cpp:73)."""
        t = self.events["t_us"]
        lo = np.searchsorted(t, self.frame_ts[k], side="left")
        hi = (np.searchsorted(t, self.frame_ts[k + 1], side="left")
              if k + 1 < len(self.frame_ts) else len(t))
        return self.events[lo:hi]


def make_stream(width: int, height: int, duration_s: float, rate_ev_s: float, seed: int,
                max_corners: int = 100, amplitude: float = 30.0, theta_max_deg: float = 5.0,
                noise_frac: float = 0.05, step_us: int = 1000) -> Stream:
    rng = np.random.Generator(np.random.PCG64(seed))
    margin = int(amplitude * 2 + 0.2 * max(width, height) + 8)
    tex = make_texture(width + 2 * margin, height + 2 * margin, seed * 31 + 7, sigma=2.5).astype(np.float64)
    pose = Pose(amplitude, np.deg2rad(theta_max_deg))

    S = Stream(width, height, [], [], np.zeros(0, EVENT), [], pose, margin)
    yy, xx = np.mgrid[0:height, 0:width]
    pix = np.stack([xx.ravel().astype(np.float64), yy.ravel().astype(np.float64)], axis=1)

    def render(t):
        q = S.to_texture(t, pix)
        return _bilinear(tex, q[:, 0], q[:, 1]).reshape(height, width)

    n_frames = int(duration_s * 1e6 // FRAME_PERIOD_US) + 1
    for k in range(n_frames):
        ts = k * FRAME_PERIOD_US
        fr = np.clip(np.rint(render(ts * 1e-6)), 0, 255).astype(np.uint8)
        S.frames.append(fr)
        S.frame_ts.append(ts)
        S.corners.append(harris_corners(fr, max_corners))

    def simulate(C, n_steps, rr, sub=1):
        """threshold-crossing event simulation; sub > 1 = every sub-th pixel in
        x and y only (rate pilot)"""
        sel = (slice(None, None, sub), slice(None, None, sub))
        ref = np.log(render(0.0)[sel] / 255.0 + 0.1)
        chunks = []
        for i in range(1, n_steps + 1):
            L = np.log(render(i * step_us * 1e-6)[sel] / 255.0 + 0.1)
            d = L - ref
            k = np.minimum(np.floor(np.abs(d) / C), 4).astype(np.int64)
            ys, xs = np.nonzero(k)
            if len(ys) == 0:
                continue
            kk = k[ys, xs]
            sg = np.sign(d[ys, xs])
            ref[ys, xs] += sg * kk * C
            rep = np.repeat(np.arange(len(ys)), kk)
            ev = np.zeros(len(rep), EVENT)
            ev["x"] = xs[rep] * sub
            ev["y"] = ys[rep] * sub
            ev["polarity"] = sg[rep].astype(np.int8)
            ev["t_us"] = (i - 1) * step_us + rr.integers(0, step_us, size=len(rep))
            chunks.append(ev)
        return np.concatenate(chunks) if chunks else np.zeros(0, EVENT)

    n_steps = int(duration_s * 1e6 // step_us)
    # contrast threshold from a pilot over the WHOLE duration on every 4th pixel
    # in x and y (two passes: the rate is not exactly proportional to 1/C)
    C = 0.15
    for _ in range(2):
        pilot = simulate(C, n_steps, np.random.Generator(np.random.PCG64(seed + 1)), sub=4)
        n_sub = len(range(0, height, 4)) * len(range(0, width, 4))
        rate0 = len(pilot) * (width * height / n_sub) / (n_steps * step_us * 1e-6)
        C = float(np.clip(C * rate0 / max(rate_ev_s * (1 - noise_frac), 1.0), 0.01, 2.0))
    ev = simulate(C, n_steps, np.random.Generator(np.random.PCG64(seed + 2)))
    n_noise = int(len(ev) * noise_frac / max(1 - noise_frac, 1e-9))
    nz = np.zeros(n_noise, EVENT)
    nz["x"] = rng.integers(0, width, n_noise)
    nz["y"] = rng.integers(0, height, n_noise)
    nz["polarity"] = rng.choice(np.array([-1, 1], np.int8), n_noise)
    nz["t_us"] = rng.integers(0, n_steps * step_us, n_noise)
    ev = np.concatenate([ev, nz])
    S.events = ev[np.argsort(ev["t_us"], kind="stable")]
    return S


def run_stream(tracker, stream: Stream, n_frames: int | None = None, on_batch=None):
    """Drive any tracker exposing the C-ABI call shapes (cabi.Tracker or the CPU
    checkers in tests/_oracle.py) through the stream the way Replayer +
    Evaluator would (replayer.cpp:56-98, evaluator.cpp:32-101): image k, then
    the events up to image k+1.  Returns the total number of optimisations."""
    n_frames = len(stream.frames) if n_frames is None else min(n_frames, len(stream.frames))
    total = 0
    for k in range(n_frames):
        tracker.new_frame_begin(stream.frames[k], stream.frame_ts[k], stream.corners[k])
        ids, xy = tracker.get_uninit()
        if len(ids):
            nxt = stream.lk_next(k, xy)
            for i, tid in enumerate(ids):
                inside = (0 <= nxt[i, 0] < stream.width) and (0 <= nxt[i, 1] < stream.height)
                tracker.set_flow(tid, int(inside), nxt[i, 0], nxt[i, 1])
        tracker.new_frame_end()
        ev = stream.events_between(k)
        if k == n_frames - 1 and n_frames < len(stream.frames):
            pass
        n = tracker.push_events(ev)
        total += n
        if on_batch:
            on_batch(k, ev, n)
    return total


if __name__ == "__main__":
    # worker of make_streams: argv = out.npz width height duration_s rate seed max_corners
    import sys
    _a = sys.argv[2:]
    save_stream(make_stream(int(_a[0]), int(_a[1]), float(_a[2]), float(_a[3]), int(_a[4]), int(_a[5])), sys.argv[1])
