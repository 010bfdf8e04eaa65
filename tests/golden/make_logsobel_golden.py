"""Golden vectors for FeatureDetector::getLogImage + getGradients
(/root/reference/implementation/feature_tracker/src/feature_detector.cpp:714-732) made with the
real OpenCV (Python cv2, authoring container only):

    image.convertTo(CV_64F, 1/255);  cv::log(normalized + 10e-2);  cv::Sobel(log / 8, CV_64F, dx, dy, 3)

cv::log on CV_64F uses OpenCV's own table-based routine, which is not bit-identical to libm's log; the
restatements (oracle/ebt_oracle.c, oracle/ref_api.cpp, csrc grad_kernel) use libm.  The fixture pins them
to OpenCV within a stated tolerance (tests/test_logsobel.py).

    python tests/golden/make_logsobel_golden.py   ->  tests/golden/logsobel_cv2.npz
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from event_based_odomety_b200 import synth  # noqa: E402


def reference_gradients(gray):
    normalized = gray.astype(np.float64) * (1.0 / 255.0)   # convertTo(CV_64F, 1.0 / 255.0)
    log_image = cv2.log(normalized + 10e-2)
    gx = cv2.Sobel(log_image / 8, cv2.CV_64F, 1, 0, ksize=3)
    gy = cv2.Sobel(log_image / 8, cv2.CV_64F, 0, 1, ksize=3)
    return log_image, gx, gy


def main():
    rng = np.random.Generator(np.random.PCG64(77))
    frames = [np.ascontiguousarray(synth.make_stream(120, 90, 0.05, 50e3, seed=1201, max_corners=4).frames[0][:48, :64]),
              rng.integers(0, 256, (48, 64)).astype(np.uint8),             # every grey level, white noise
              np.tile(np.arange(64, dtype=np.uint8) * 4, (48, 1))]         # a ramp: exact zeros in gy
    frames[2][:, :8] = 0
    frames[2][:, -8:] = 255
    out = {"frames": np.stack(frames), "cv2_version": cv2.__version__}
    gxs, gys, logs = [], [], []
    for f in frames:
        L, gx, gy = reference_gradients(f)
        logs.append(L)
        gxs.append(gx)
        gys.append(gy)
    out["gx"] = np.stack(gxs)
    out["gy"] = np.stack(gys)
    # the 256 values cv::log takes on an 8-bit input, as OpenCV computes them
    lut_in = (np.arange(256, dtype=np.float64) * (1.0 / 255.0) + 10e-2).reshape(1, 256)
    out["log_lut"] = cv2.log(lut_in).reshape(256)
    np.savez_compressed(os.path.join(HERE, "logsobel_cv2.npz"), **out)
    lib = np.log(lut_in).reshape(256)
    print("cv2", cv2.__version__, "max |cv::log - libm log| over the 256 inputs:", np.abs(out["log_lut"] - lib).max())


if __name__ == "__main__":
    main()
