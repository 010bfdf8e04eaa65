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
