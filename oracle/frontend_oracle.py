"""CPU restatement (numpy) of the per-frame re-arm the reference delegates to
OpenCV — TEST INFRASTRUCTURE ONLY (only tests/, smoke() and bench's CPU legs may
import this; the product path is event_based_odomety_b200/csrc).

SURVEY 8f n1:
* corner detection: FeatureDetector::detectFeatures
  (/root/reference/implementation/feature_tracker/src/feature_detector.cpp:568-583)
  = cv::goodFeaturesToTrack(gray, maxCorners = W*H/625, qualityLevel 0.01,
  minDistance 10, mask (patchExtent inset, :34-43), blockSize 3, useHarris, k 0.04)
* flow initialisation: FlowEstimator::getFlow (src/flow_estimator.cpp:84-103)
  = cv::calcOpticalFlowPyrLK(prev, cur, {pt}, winSize 21x21, maxLevel 3), default
  criteria (30 iterations, eps 0.01), flags 0, minEigThreshold 1e-4.

OpenCV is a system dependency of the reference (not vendored): the algorithm
restated here is OpenCV 4.x's imgproc/featureselect.cpp, imgproc/corner.cpp,
video/lkpyramid.cpp, pinned against golden vectors recorded from Python cv2
4.13 (tests/golden/frontend_cv2.npz, written by tests/golden/make_frontend_golden.py).
Float summation order inside OpenCV's SIMD loops is not reproduced; the pins are
"same corner list" and "LK points within 0.01 px, same status".
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


# ---------------------------------------------------------------------------
# goodFeaturesToTrack, Harris variant
# ---------------------------------------------------------------------------
def harris_response(gray: np.ndarray, block: int = 3, k: float = 0.04) -> np.ndarray:
    """cv::cornerHarris(gray, block, ksize=3, k), BORDER_REFLECT_101 (corner.cpp
    cornerEigenValsVecs + calcHarris): Sobel 3x3 scaled by 1/(4*block*255),
    products, un-normalised block x block box sum, det - k trace^2, all float32."""
    assert block == 3
    h, w = gray.shape
    scale = F32(1.0 / ((1 << 2) * block * 255.0))
    p = np.pad(gray.astype(F32), 1, mode="reflect")
    dxi = (p[1:-1, 2:] - p[1:-1, :-2]) * 2 + (p[:-2, 2:] - p[:-2, :-2]) + (p[2:, 2:] - p[2:, :-2])
    dyi = (p[2:, 1:-1] - p[:-2, 1:-1]) * 2 + (p[2:, :-2] - p[:-2, :-2]) + (p[2:, 2:] - p[:-2, 2:])
    dx = (dxi * scale).astype(F32)
    dy = (dyi * scale).astype(F32)

    def box(a):
        q = np.pad(a, 1, mode="reflect")
        s = np.zeros_like(a)
        for i in range(3):
            for j in range(3):
                s = (s + q[i:i + h, j:j + w]).astype(F32)
        return s

    a, b, c = box(dx * dx), box(dx * dy), box(dy * dy)
    return (a * c - b * b - F32(k) * (a + c) * (a + c)).astype(F32)


def select_corners(resp: np.ndarray, max_corners: int, quality: float, min_distance: float,
                   mask: np.ndarray) -> np.ndarray:
    """featureselect.cpp: threshold at quality * max, 3x3 non-maximum suppression
    on the interior, descending sort (ties: larger address first, greaterThanPtr
    in OpenCV 4.x; pinned by the "ties" golden case), greedy minimum
    distance on a grid of min_distance cells.  Returns (n, 2) float32 (x, y)."""
    h, w = resp.shape
    thr = F32(float(resp.max()) * float(quality))  # double product, cast to float by cv::threshold
    e = np.where(resp > thr, resp, F32(0))
    q = np.pad(e, 1, mode="constant", constant_values=-np.inf)
    d = np.max([q[i:i + h, j:j + w] for i in range(3) for j in range(3)], axis=0)
    ok = (e != 0) & (e == d) & (mask != 0)
    ok[0, :] = ok[-1, :] = False
    ok[:, 0] = ok[:, -1] = False
    ys, xs = np.nonzero(ok)
    vals = e[ys, xs]
    order = np.lexsort((-xs, -ys, -vals))
    ys, xs = ys[order], xs[order]
    cell = int(round(min_distance))
    gw, gh = (w + cell - 1) // cell, (h + cell - 1) // cell
    grid = [[[] for _ in range(gw)] for _ in range(gh)]
    md2 = min_distance * min_distance
    out = []
    for y, x in zip(ys.tolist(), xs.tolist()):
        xc, yc = x // cell, y // cell
        good = True
        for yy in range(max(0, yc - 1), min(gh - 1, yc + 1) + 1):
            for xx in range(max(0, xc - 1), min(gw - 1, xc + 1) + 1):
                for (px, py) in grid[yy][xx]:
                    if (x - px) * (x - px) + (y - py) * (y - py) < md2:
                        good = False
                        break
                if not good:
                    break
            if not good:
                break
        if good:
            grid[yc][xc].append((x, y))
            out.append((x, y))
            if max_corners > 0 and len(out) == max_corners:
                break
    return np.array(out, F32).reshape(-1, 2)


def detect_corners(gray: np.ndarray, patch_extent: int = 12, quality: float = 0.01,
                   min_distance: float = 10.0) -> np.ndarray:
    """FeatureDetector::detectFeatures with the reference's mask and maxCorners."""
    h, w = gray.shape
    mask = np.zeros((h, w), np.uint8)
    mask[patch_extent:h - patch_extent, patch_extent:w - patch_extent] = 1  # cv::rectangle, filled
    return select_corners(harris_response(gray), w * h // ((2 * patch_extent + 1) ** 2), quality,
                          min_distance, mask)


# ---------------------------------------------------------------------------
# calcOpticalFlowPyrLK (video/lkpyramid.cpp), 8-bit single channel
# ---------------------------------------------------------------------------
WIN = 21          # FlowEstimatorParams::window (flow_estimator.h)
MAX_LEVEL = 3     # FlowEstimatorParams::numLevels
W_BITS = 14
FLT_SCALE = F32(1.0 / (1 << 20))
MIN_EIG = F32(1e-4)
FLT_EPS = F32(np.finfo(np.float32).eps)


def pyr_down(img: np.ndarray) -> np.ndarray:
    """cv::pyrDown, CV_8U: separable [1 4 6 4 1], BORDER_REFLECT_101, (v + 128) >> 8."""
    h, w = img.shape
    oh, ow = (h + 1) // 2, (w + 1) // 2
    p = np.pad(img.astype(np.int32), 2, mode="reflect")
    k = (1, 4, 6, 4, 1)
    rows = sum(k[i] * p[:, i:i + 2 * ow:2][:, :ow] for i in range(5))  # horizontal, even columns
    out = sum(k[i] * rows[i:i + 2 * oh:2][:oh] for i in range(5))
    return ((out + 128) >> 8).astype(np.uint8)


def build_pyramid(img: np.ndarray, max_level: int = MAX_LEVEL, win: int = WIN):
    """buildOpticalFlowPyramid: stops before a level that is not larger than the window."""
    pyr = [img]
    h, w = img.shape
    for _ in range(max_level):
        w, h = (w + 1) // 2, (h + 1) // 2
        if w <= win or h <= win:
            break
        pyr.append(pyr_down(pyr[-1]))
    return pyr


def scharr_deriv(img: np.ndarray):
    """calcSharrDeriv: int16 (dx, dy), reflect-101 at the image border."""
    p = np.pad(img.astype(np.int32), 1, mode="reflect")
    t0 = (p[:-2, :] + p[2:, :]) * 3 + p[1:-1, :] * 10   # vertical smoothing, all padded columns
    t1 = p[2:, :] - p[:-2, :]                           # vertical derivative
    dx = t0[:, 2:] - t0[:, :-2]
    dy = (t1[:, 2:] + t1[:, :-2]) * 3 + t1[:, 1:-1] * 10
    return dx.astype(np.int16), dy.astype(np.int16)


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _weights(a, b):
    one = F32(1.0)
    s = F32(1 << W_BITS)
    iw00 = int(np.rint((one - a) * (one - b) * s))
    iw01 = int(np.rint(a * (one - b) * s))
    iw10 = int(np.rint((one - a) * b * s))
    return iw00, iw01, iw10, (1 << W_BITS) - iw00 - iw01 - iw10


def _interp(img_pad, x0, y0, w4, pad, n_bits):
    """(WIN x WIN) fixed-point bilinear patch with top-left integer corner (x0, y0)."""
    iw00, iw01, iw10, iw11 = w4
    ys, xs = y0 + pad, x0 + pad
    a = img_pad[ys:ys + WIN, xs:xs + WIN]
    b = img_pad[ys:ys + WIN, xs + 1:xs + 1 + WIN]
    c = img_pad[ys + 1:ys + 1 + WIN, xs:xs + WIN]
    d = img_pad[ys + 1:ys + 1 + WIN, xs + 1:xs + 1 + WIN]
    return _descale(a * iw00 + b * iw01 + c * iw10 + d * iw11, n_bits)


def lk_flow(prev: np.ndarray, cur: np.ndarray, pts: np.ndarray, max_count: int = 30, eps: float = 0.01):
    """Returns (next_pts float32 (n,2), status uint8 (n,))."""
    pts = np.asarray(pts, F32).reshape(-1, 2)
    n = len(pts)
    P, C = build_pyramid(prev), build_pyramid(cur)
    levels = len(P)
    pad = WIN + 1
    Ipad = [np.pad(p.astype(np.int32), pad, mode="reflect") for p in P]
    Jpad = [np.pad(c.astype(np.int32), pad, mode="reflect") for c in C]
    D = []
    for p in P:
        dx, dy = scharr_deriv(p)
        D.append((np.pad(dx.astype(np.int32), pad), np.pad(dy.astype(np.int32), pad)))  # constant 0 border
    nxt = np.zeros((n, 2), F32)
    status = np.ones(n, np.uint8)
    half = F32((WIN - 1) * 0.5)
    eps2 = eps * eps
    for level in range(levels - 1, -1, -1):
        h, w = P[level].shape
        for i in range(n):
            prev_pt = pts[i] * F32(1.0 / (1 << level))
            next_pt = prev_pt.copy() if level == levels - 1 else nxt[i] * F32(2.0)
            nxt[i] = next_pt
            prev_pt = prev_pt - half
            ix, iy = int(np.floor(prev_pt[0])), int(np.floor(prev_pt[1]))
            if ix < -WIN or ix >= w or iy < -WIN or iy >= h:
                if level == 0:
                    status[i] = 0
                continue
            a, b = F32(prev_pt[0] - F32(ix)), F32(prev_pt[1] - F32(iy))
            w4 = _weights(a, b)
            Iw = _interp(Ipad[level], ix, iy, w4, pad, W_BITS - 5)
            Ix = _interp(D[level][0], ix, iy, w4, pad, W_BITS)
            Iy = _interp(D[level][1], ix, iy, w4, pad, W_BITS)
            # (short) casts as in the reference
            Iw, Ix, Iy = Iw.astype(np.int16).astype(np.int64), Ix.astype(np.int16).astype(np.int64), \
                Iy.astype(np.int16).astype(np.int64)
            A11 = F32((Ix * Ix).astype(F32).sum(dtype=np.float64)) * FLT_SCALE
            A12 = F32((Ix * Iy).astype(F32).sum(dtype=np.float64)) * FLT_SCALE
            A22 = F32((Iy * Iy).astype(F32).sum(dtype=np.float64)) * FLT_SCALE
            Dd = A11 * A22 - A12 * A12
            min_eig = (A22 + A11 - np.sqrt((A11 - A22) * (A11 - A22) + F32(4.0) * A12 * A12)) / F32(2 * WIN * WIN)
            if min_eig < MIN_EIG or Dd < FLT_EPS:
                if level == 0:
                    status[i] = 0
                continue
            Dd = F32(1.0) / Dd
            next_pt = next_pt - half
            prev_delta = np.zeros(2, F32)
            for j in range(max_count):
                jx, jy = int(np.floor(next_pt[0])), int(np.floor(next_pt[1]))
                if jx < -WIN or jx >= w or jy < -WIN or jy >= h:
                    if level == 0:
                        status[i] = 0
                    break
                a, b = F32(next_pt[0] - F32(jx)), F32(next_pt[1] - F32(jy))
                Jw = _interp(Jpad[level], jx, jy, _weights(a, b), pad, W_BITS - 5)
                diff = Jw - Iw
                b1 = F32((diff * Ix).astype(F32).sum(dtype=np.float64)) * FLT_SCALE
                b2 = F32((diff * Iy).astype(F32).sum(dtype=np.float64)) * FLT_SCALE
                delta = np.array([(A12 * b2 - A22 * b1) * Dd, (A12 * b1 - A11 * b2) * Dd], F32)
                next_pt = next_pt + delta
                nxt[i] = next_pt + half
                if float(delta[0]) * float(delta[0]) + float(delta[1]) * float(delta[1]) <= eps2:
                    break
                if j > 0 and abs(delta[0] + prev_delta[0]) < 0.01 and abs(delta[1] + prev_delta[1]) < 0.01:
                    nxt[i] = nxt[i] - delta * F32(0.5)
                    break
                prev_delta = delta
    return nxt, status
