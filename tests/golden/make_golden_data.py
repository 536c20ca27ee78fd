#!/usr/bin/env python
"""Golden vectors for the data path (data_reader.py:45-64) from the real cv2 in this image
(opencv-python-headless): tests/golden/data.npz.  Run in the build container:

    python tests/golden/make_golden_data.py

Each case stores the two uint8 BGR source frames, the target size, the flip flag, cv2.resize's uint8 outputs and
the float32 (H, W, 6) item exactly as read_item + read_batch produce it (the reference's own lines, inlined below
because data_reader.py reads files from disk and draws the flip from np.random)."""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

image_means = np.array([123.0, 117.0, 104.0])        # data_reader.py:7-9
image_means /= 255.0
image_means = np.reshape(image_means, [1, 1, 3])


def read_item_like(img1, img2, img_height, img_width, flip):
    if flip:                                          # data_reader.py:46-51
        img1, img2 = img2, img1
    img1 = cv2.resize(img1, (img_width, img_height))  # :55-56
    img2 = cv2.resize(img2, (img_width, img_height))
    r1, r2 = img1.copy(), img2.copy()
    img1 = img1.astype(np.float32) / 255.0            # :58-59
    img2 = img2.astype(np.float32) / 255.0
    img1 = img1 - image_means                         # :61-62
    img2 = img2 - image_means
    item = np.zeros((img_height, img_width, 6), np.float32)   # :36,:40-41
    item[:, :, :3] = img1
    item[:, :, 3:] = img2
    return r1, r2, item


if __name__ == "__main__":
    r = np.random.default_rng(20260102)
    out = {}
    cases = [(37, 123, 40, 128, 0), (47, 155, 48, 160, 1), (75, 248, 32, 104, 0), (20, 30, 60, 90, 1), (32, 32, 32, 32, 0),
             (50, 60, 25, 30, 1), (9, 200, 31, 57, 0)]
    for k, (h, w, oh, ow, flip) in enumerate(cases):
        a = r.integers(0, 256, (h, w, 3), dtype=np.uint8)
        b = np.clip(a.astype(np.int32) + r.integers(-20, 21, a.shape), 0, 255).astype(np.uint8)
        r1, r2, item = read_item_like(a, b, oh, ow, flip)
        out["c%d_img1" % k], out["c%d_img2" % k] = a, b
        out["c%d_meta" % k] = np.array([oh, ow, flip], np.int32)
        out["c%d_resized1" % k], out["c%d_resized2" % k], out["c%d_item" % k] = r1, r2, item
    out["ncases"] = np.int32(len(cases))
    np.savez_compressed(os.path.join(HERE, "data.npz"), **out)
    print("data.npz:", len(cases), "cases, cv2", cv2.__version__)
