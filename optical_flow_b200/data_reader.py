"""Device-side mirror of the reference's data_reader.py item / batch assembly.

The reference decodes two frames with cv2.imread, resizes them with cv2.resize, scales to [0, 1], subtracts the
BGR means and stacks the pair on the channel axis (data_reader.py:35-64) in 6 worker processes.  Here the decoded
uint8 frames are handed to the GPU (pinned host memory -> device copy) and ONE kernel per item does the rest,
bit-exactly (csrc/data.cu; oracle/ref_data.py; tests/golden/data.npz from the real cv2).  File decoding and the
KITTI directory walk (read_kitti, AsyncReader) stay on the host and are out of scope.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L

image_means = np.array([123.0, 117.0, 104.0]) / 255.0             # data_reader.py:7-9


class ReaderOpts:
    """data_reader.py:71-77 (the fields the item path uses)."""

    def __init__(self, kitti_path=None, batch_size=4, img_height=192, img_width=640, nworkers=0):
        self.kitti_path, self.batch_size = kitti_path, batch_size
        self.img_height, self.img_width, self.nworkers = img_height, img_width, nworkers


def _as_device_u8(img, device):
    if isinstance(img, np.ndarray):
        img = torch.from_numpy(np.ascontiguousarray(img))
    assert img.dtype == torch.uint8 and img.dim() == 3 and img.shape[2] == 3, "frames are (h, w, 3) uint8 BGR"
    return img.to(device, non_blocking=True).contiguous()


def read_item(img1, img2, opts, out=None, flip=None, device=None):
    """data_reader.py:45-64 on decoded frames.  Returns the (H, W, 6) float32 device tensor (or fills ``out``).
    ``flip``: None draws np.random.rand() < 0.5 like the reference (:46), else the explicit choice
    (True = the reference's else-branch: the pair is swapped)."""
    device = device or (out.device if out is not None else torch.device("cuda", torch.cuda.current_device()))
    if flip is None:
        flip = not (np.random.rand() < 0.5)
    a, b = _as_device_u8(img1, device), _as_device_u8(img2, device)
    assert tuple(a.shape) == tuple(b.shape), "the two frames of a pair have the same size"
    if out is None:
        out = torch.empty((opts.img_height, opts.img_width, 6), dtype=torch.float32, device=device)
    assert out.is_contiguous() and tuple(out.shape) == (opts.img_height, opts.img_width, 6)
    L.check(L.load().of_preprocess_pair(L.ptr(a), L.ptr(b), a.shape[0], a.shape[1], L.ptr(out), opts.img_height,
                                        opts.img_width, 1 if flip else 0, L.stream_ptr()), "of_preprocess_pair")
    out._keepalive = (a, b)                                        # the kernel is asynchronous
    return out


def read_batch(frame_pairs, opts, flips=None, device=None):
    """data_reader.py:35-42: (batch_size, H, W, 6) float32 from a list of decoded (frame1, frame2) pairs."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    batch = torch.zeros((opts.batch_size, opts.img_height, opts.img_width, 6), dtype=torch.float32, device=device)
    keep = []
    for i, (f1, f2) in enumerate(frame_pairs):
        item = read_item(f1, f2, opts, out=batch[i], flip=None if flips is None else bool(flips[i]), device=device)
        keep.append(item._keepalive)
    batch._keepalive = keep
    return batch


def read_batch_frames(frames, opts, flips=None, out=None):
    """The same batch assembly from ONE device tensor of decoded pairs, frames = (B, 2, h, w, 3) uint8 (a single
    pinned host -> device copy of 6 bytes per pixel pair instead of 24 bytes of float32), one launch for the whole
    batch (of_preprocess_batch).  ``flips``: None = no swaps, else B booleans (data_reader.py:46-51 drawn by the caller)."""
    assert frames.dtype == torch.uint8 and frames.dim() == 5 and frames.shape[1] == 2 and frames.shape[4] == 3
    assert frames.is_cuda and frames.is_contiguous()
    B, _, h, w, _ = frames.shape
    if out is None:
        out = torch.empty((B, opts.img_height, opts.img_width, 6), dtype=torch.float32, device=frames.device)
    assert out.is_contiguous() and tuple(out.shape) == (B, opts.img_height, opts.img_width, 6)
    fl = None
    if flips is not None:
        import ctypes
        fl = (ctypes.c_uint8 * B)(*[1 if f else 0 for f in flips])
    L.check(L.load().of_preprocess_batch(L.ptr(frames), B, h, w, L.ptr(out), opts.img_height, opts.img_width, fl,
                                         L.stream_ptr()), "of_preprocess_batch")
    return out
