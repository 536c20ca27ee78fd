"""Data path (data_reader.py:35-64): the oracle against vectors from the real cv2 (CPU), the CUDA
kernel against both (GPU).  Bit-exact: uint8 resize results and the float32 items."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_data as D  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "data.npz"))
CASES = list(range(int(G["ncases"])))


def case(k):
    oh, ow, flip = [int(v) for v in G["c%d_meta" % k]]
    return G["c%d_img1" % k], G["c%d_img2" % k], oh, ow, bool(flip)


@pytest.mark.parametrize("k", CASES)
def test_oracle_resize_and_item_match_cv2(k):
    a, b, oh, ow, flip = case(k)
    x, y = (b, a) if flip else (a, b)
    assert np.array_equal(D.resize_linear_u8(x, oh, ow), G["c%d_resized1" % k])
    assert np.array_equal(D.resize_linear_u8(y, oh, ow), G["c%d_resized2" % k])
    assert np.array_equal(D.read_item(a, b, oh, ow, flip), G["c%d_item" % k])


def test_oracle_edge_cases():
    r = np.random.default_rng(0)
    img = r.integers(0, 256, (1, 1, 3), dtype=np.uint8)                  # a single pixel replicates
    assert (D.resize_linear_u8(img, 5, 7) == img[0, 0]).all()
    img = r.integers(0, 256, (6, 9, 3), dtype=np.uint8)
    assert np.array_equal(D.resize_linear_u8(img, 6, 9), img)             # identity size = copy
    item = D.read_item(img, img, 6, 9)
    assert item.dtype == np.float32 and item.shape == (6, 9, 6)
    assert np.array_equal(item[..., :3], item[..., 3:])
    lo = (np.float32(0) / np.float32(255) - D.IMAGE_MEANS).astype(np.float32)
    hi = (np.float32(255) / np.float32(255) - D.IMAGE_MEANS).astype(np.float32)
    assert (item[..., :3] >= lo).all() and (item[..., :3] <= hi).all()


@pytest.mark.gpu
@pytest.mark.parametrize("k", CASES)
def test_cuda_read_item_matches_cv2_and_oracle(cuda, k):
    from optical_flow_b200 import data_reader as R
    a, b, oh, ow, flip = case(k)
    opts = R.ReaderOpts(batch_size=1, img_height=oh, img_width=ow)
    got = R.read_item(a, b, opts, flip=flip, device=cuda).cpu().numpy()
    assert np.array_equal(got, G["c%d_item" % k])                          # the reference's own lines over real cv2
    assert np.array_equal(got, D.read_item(a, b, oh, ow, flip))


@pytest.mark.gpu
def test_cuda_read_batch_full_size(cuda):
    """KITTI-sized frames (375 x 1242) to the bench resolution 384 x 1280, batch of 2, against the oracle."""
    from optical_flow_b200 import data_reader as R
    r = np.random.default_rng(5)
    pairs = [(r.integers(0, 256, (375, 1242, 3), dtype=np.uint8), r.integers(0, 256, (375, 1242, 3), dtype=np.uint8))
             for _ in range(2)]
    opts = R.ReaderOpts(batch_size=2, img_height=384, img_width=1280)
    batch = R.read_batch(pairs, opts, flips=[False, True], device=cuda).cpu().numpy()
    for i, (f1, f2) in enumerate(pairs):
        assert np.array_equal(batch[i], D.read_item(f1, f2, 384, 1280, flip=bool(i)))


@pytest.mark.gpu
def test_cuda_read_batch_frames_one_launch(cuda):
    """of_preprocess_batch (one H2D copy of the uint8 pairs + one launch) == the per-item path, with and without resize."""
    import torch
    from optical_flow_b200 import data_reader as R
    r = np.random.default_rng(6)
    for (sh, sw, oh, ow) in ((375, 1242, 384, 1280), (48, 64, 48, 64)):
        frames = r.integers(0, 256, (3, 2, sh, sw, 3), dtype=np.uint8)
        flips = [False, True, False]
        opts = R.ReaderOpts(batch_size=3, img_height=oh, img_width=ow)
        got = R.read_batch_frames(torch.from_numpy(frames).to(cuda), opts, flips=flips).cpu().numpy()
        for i in range(3):
            assert np.array_equal(got[i], D.read_item(frames[i, 0], frames[i, 1], oh, ow, flip=flips[i]))
