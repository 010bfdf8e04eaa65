"""Records golden vectors for the per-frame re-arm from Python
OpenCV (cv2 4.13 in the authoring container): cv2.goodFeaturesToTrack with the
reference's arguments (feature_detector.cpp:568-583, mask :34-43) and
cv2.calcOpticalFlowPyrLK with the reference's (flow_estimator.cpp:84-103), on
seeded synthetic frames.  Run:  python tests/golden/make_frontend_golden.py
Writes tests/golden/frontend_cv2.npz (frames are regenerated from the seeds by
the tests, only cv2's outputs are stored)."""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from event_based_odomety_b200 import synth  # noqa: E402
from tests.golden.tie_image import tie_image  # noqa: E402

CASES = [(240, 180, 1001), (346, 260, 1002), (640, 480, 1003)]


def frames_of(w, h, seed):
    S = synth.make_stream(w, h, 0.12, 500e3, seed=seed, max_corners=10)
    return S.frames[:3]


def main():
    out = {"cv2_version": np.array(cv2.__version__)}
    img = tie_image()
    mask = np.zeros(img.shape, np.uint8)
    cv2.rectangle(mask, (12, 12, 200 - 24, 100 - 24), 1, cv2.FILLED)
    out["corners_ties"] = cv2.goodFeaturesToTrack(img, 200 * 100 // 625, 0.01, 10, mask=mask, blockSize=3,
                                                  useHarrisDetector=True, k=0.04).reshape(-1, 2)
    for w, h, seed in CASES:
        fr = frames_of(w, h, seed)
        mask = np.zeros((h, w), np.uint8)
        cv2.rectangle(mask, (12, 12, w - 24, h - 24), 1, cv2.FILLED)
        for k, img in enumerate(fr):
            c = cv2.goodFeaturesToTrack(img, w * h // 625, 0.01, 10, mask=mask, blockSize=3,
                                        useHarrisDetector=True, k=0.04)
            c = np.zeros((0, 2), np.float32) if c is None else c.reshape(-1, 2)
            out[f"corners_{w}_{k}"] = c
            out[f"harris_{w}_{k}"] = cv2.cornerHarris(img, 3, 3, 0.04)[::7, ::5].copy()  # subsampled response
            if k + 1 < len(fr):
                # points the tracker would ask about: corners plus sub-pixel offsets and a few near the border
                rng = np.random.default_rng(seed + k)
                pts = np.concatenate([c + rng.uniform(-0.5, 0.5, c.shape).astype(np.float32),
                                      np.array([[1.5, 2.5], [w - 2.25, h - 3.5], [w * 0.5, 0.25]], np.float32)])
                nxt, st, _ = cv2.calcOpticalFlowPyrLK(img, fr[k + 1], pts.reshape(-1, 1, 2), None,
                                                      winSize=(21, 21), maxLevel=3)
                out[f"lk_pts_{w}_{k}"] = pts
                out[f"lk_next_{w}_{k}"] = nxt.reshape(-1, 2)
                out[f"lk_status_{w}_{k}"] = st.ravel()
    path = os.path.join(ROOT, "tests", "golden", "frontend_cv2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items() if k != "cv2_version"})


if __name__ == "__main__":
    main()
