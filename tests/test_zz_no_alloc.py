"""of_bn_train_stats_static: the allocation-free entry point (scratch in a file-scope __device__ array inside the
library, no counterpart in the reference) against of_bn_train_stats with a caller workspace, through the C ABI, on two
calls in a row.  Both run the same deterministic kernel, so the sums must be bit-identical (the BN tests' tolerance
against the fp64 reference is 1e-4; this check is exact).  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _stats(lib, L, fn, x, shift, ws=None):
    C = x.shape[-1]
    npix = x.numel() // C
    s = torch.full((C,), float("nan"), device=x.device)
    q = torch.full((C,), float("nan"), device=x.device)
    args = [L.ptr(x), L.ptr(shift) if shift is not None else None, L.ptr(s), L.ptr(q), npix, C, L.dtype_code(x)]
    if ws is not None:
        args.append(L.ptr(ws))
    args.append(L.stream_ptr())
    L.check(fn(*args), fn.__name__)
    return s, q


@pytest.mark.parametrize("shape,dtype", [((2, 48, 160, 64), torch.float32), ((3, 17, 9, 256), torch.float32),
                                         ((4, 96, 320, 128), torch.bfloat16)])
def test_bn_stats_static_matches_workspace_variant(cuda, shape, dtype):
    from optical_flow_b200 import _lib as L
    lib = L.load()
    C = shape[-1]
    g = torch.Generator(device="cpu").manual_seed(sum(shape))
    ws = torch.zeros(lib.of_bn_train_workspace_bytes(C), dtype=torch.uint8, device=cuda)
    for call in range(2):                      # two calls in a row: the shared array keeps the previous call's contents
        x = (torch.randn(shape, generator=g) * (1.5 + call) + 0.3 * call).to(dtype).to(cuda)
        shift = (0.1 * torch.randn(C, generator=g)).to(cuda) if call else None
        e_s, e_q = _stats(lib, L, lib.of_bn_train_stats, x, shift, ws)
        s, q = _stats(lib, L, lib.of_bn_train_stats_static, x, shift)
        torch.cuda.synchronize()
        assert torch.equal(s, e_s) and torch.equal(q, e_q), (call, (s - e_s).abs().max(), (q - e_q).abs().max())
        # and both agree with a float64 sum of the same values within the BN tests' tolerance
        xs = x.double().reshape(-1, C) - (shift.double() if shift is not None else 0.0)
        ref_s, ref_q = xs.sum(0), (xs * xs).sum(0)
        assert (s.double() - ref_s).abs().max() <= 1e-4 * max(ref_s.abs().max().item(), xs.abs().sum(0).max().item() * 1e-3)
        assert (q.double() - ref_q).abs().max() <= 1e-4 * ref_q.abs().max()


def test_bn_stats_static_rejects_wide_c(cuda):
    from optical_flow_b200 import _lib as L
    lib = L.load()
    x = torch.zeros((1, 2, 2, 512), device=cuda)
    s = torch.zeros(512, device=cuda); q = torch.zeros(512, device=cuda)
    rc = lib.of_bn_train_stats_static(L.ptr(x), None, L.ptr(s), L.ptr(q), 4, 512, L.dtype_code(x), L.stream_ptr())
    assert rc == -2 and b"exceeds the static scratch" in lib.of_last_error()
