"""ORACLE (test infrastructure only - never imported by the product path).

CPU restatement of the reference's per-item preprocessing, data_reader.py:45-64 (read_item):

    img = cv2.resize(img, (img_width, img_height))        # uint8 BGR, INTER_LINEAR          (:55-56)
    img = img.astype(np.float32) / 255.0                  #                                   (:58-59)
    img = img - image_means                               # float64 [123,117,104]/255 (:7-9), (:61-62)
    batch_imgs_np[i, :, :, :3] = image1 ; [..., 3:] = image2     # float32 batch              (:40-41)

with the random pair-order swap of :46-51 as an explicit ``flip`` flag.

cv2.resize on uint8 with INTER_LINEAR is OpenCV's fixed-point bilinear (opencv/modules/imgproc/src/resize.cpp,
resizeGeneric_ / HResizeLinear / VResizeLinear<uchar,int,short>), a third-party dependency that is not in
/root/reference; its published algorithm is restated here and pinned bit-exactly against the cv2 wheel
in this image (opencv-python-headless 4.13) by tests/golden/make_golden_data.py -> tests/golden/data.npz:

    scale = src / dst ;  f = (d + 0.5) * scale - 0.5 ;  s = floor(f) ;  f -= s
    x axis: s < 0 -> (s, f) = (0, 0) ;  s >= src - 1 -> (s, f) = (src - 1, 0)
    y axis: weights kept, row indices s and s + 1 clipped to [0, src - 1] when fetched
    coefficients: short(round_half_even((1 - f) * 2048)), short(round_half_even(f * 2048))   (11 bits)
    horizontal:   row[x]  = S[sx] * a0 + S[sx + 1] * a1                                       (int32)
    vertical:     dst     = ((b0 * (row0 >> 4) >> 16) + (b1 * (row1 >> 4) >> 16) + 2) >> 2    (uint8)

(the float coefficient product is computed in float32, as OpenCV does).
"""
import numpy as np

IMAGE_MEANS = np.array([123.0, 117.0, 104.0]) / 255.0           # data_reader.py:7-8 (applied to B, G, R as read)
COEF_BITS = 11
COEF_SCALE = 1 << COEF_BITS


def _axis_tables(src, dst, vertical=False):
    """Source index pair and 11-bit coefficients for one axis (resize.cpp: the INTER_LINEAR branch).
    The two axes differ at the borders: the x loop resets (sx, fx) to (0, 0) / (src - 1, 0) outside the image,
    while the y loop keeps its fractional weights and only the ROW INDICES are clipped when the rows are fetched
    (so the first / last output rows of an upscale blend the border row with itself, truncating twice)."""
    scale = np.float64(1.0) / (np.float64(dst) / np.float64(src))    # resize.cpp: scale_x = 1. / inv_scale_x
    d = np.arange(dst)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)              # float fx = (float)((dx + 0.5) * scale_x - 0.5)
    s = np.floor(f).astype(np.int64)
    f = (f - s.astype(np.float32)).astype(np.float32)
    if vertical:
        c0 = np.rint((np.float32(1.0) - f) * np.float32(COEF_SCALE)).astype(np.int32)
        c1 = np.rint(f * np.float32(COEF_SCALE)).astype(np.int32)
        return np.clip(s, 0, src - 1), np.clip(s + 1, 0, src - 1), c0, c1
    lo = s < 0
    s[lo], f[lo] = 0, 0.0
    hi = s >= src - 1
    s[hi], f[hi] = src - 1, 0.0
    c0 = np.rint((np.float32(1.0) - f) * np.float32(COEF_SCALE)).astype(np.int64)   # cvRound: half to even
    c1 = np.rint(f * np.float32(COEF_SCALE)).astype(np.int64)
    s1 = np.minimum(s + 1, src - 1)                               # the second tap of a clamped pixel has weight 0
    return s, s1, c0.astype(np.int32), c1.astype(np.int32)


def resize_linear_u8(img, out_h, out_w):
    """cv2.resize(img, (out_w, out_h)) for uint8 HxWxC input, INTER_LINEAR - bit-exact."""
    assert img.dtype == np.uint8 and img.ndim == 3
    H, W, _ = img.shape
    if (H, W) == (out_h, out_w):
        return img.copy()
    sx0, sx1, a0, a1 = _axis_tables(W, out_w)
    sy0, sy1, b0, b1 = _axis_tables(H, out_h, vertical=True)
    src = img.astype(np.int32)
    rows = src[:, sx0, :] * a0[None, :, None] + src[:, sx1, :] * a1[None, :, None]    # (H, out_w, C) int32
    r0, r1 = rows[sy0], rows[sy1]
    v = (((b0[:, None, None] * (r0 >> 4)) >> 16) + ((b1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(v, 0, 255).astype(np.uint8)


def normalize(img_u8):
    """data_reader.py:58-62: float32(u8) / 255.0 (float32), minus the float64 means, stored as float32."""
    x = img_u8.astype(np.float32) / np.float32(255.0)
    return (x.astype(np.float64) - IMAGE_MEANS.reshape(1, 1, 3)).astype(np.float32)


def read_item(img1_u8, img2_u8, out_h, out_w, flip=False):
    """data_reader.py:45-64 on already-decoded frames; ``flip`` = the branch of :46-51 that swaps the pair."""
    if flip:
        img1_u8, img2_u8 = img2_u8, img1_u8
    a = normalize(resize_linear_u8(img1_u8, out_h, out_w))
    b = normalize(resize_linear_u8(img2_u8, out_h, out_w))
    return np.concatenate([a, b], axis=-1)                         # data_reader.py:40-41
