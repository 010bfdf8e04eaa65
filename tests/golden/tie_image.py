"""Test image with repeated corner responses (no cv2 import, usable on the GPU box)."""
import numpy as np


def tie_image():
    """Five identical dark squares: every corner response occurs several times, which pins the
    order goodFeaturesToTrack gives to equal responses."""
    img = np.full((100, 200), 200, np.uint8)
    for (x, y) in [(30, 30), (90, 30), (150, 30), (30, 70), (150, 70)]:
        img[y:y + 12, x:x + 12] = 20
    return img
