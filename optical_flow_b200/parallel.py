"""Data-parallel host logic of the train step: one process per GPU, the batch sharded
by samples, ONE sum-allreduce of the flat fp32 gradient per step, 1/N folded into Adam.

The reference is single-device (train.py:47-61); this is the exchange step that makes N replicas
equal one replica on the concatenated batch: every op of the path is per-sample and each loss term
is a mean over B*h*w*3 (loss.py:28), so with equal shards  grad(global) = (1/N) * sum_r grad(shard r).
BatchNorm runs in inference mode: no cross-replica statistics.

Two transports, same semantics:
  * ``CAbiComm``  - the library's own NCCL communicator (of_comm_*), used on GPUs;
  * ``TorchDistComm`` - torch.distributed (any backend); with ``gloo`` it lets the host logic be
    tested on CPU at world_size 2 (tests/test_c1_dp_gloo.py).
"""
from __future__ import annotations

import ctypes

import torch


def shard_range(global_batch: int, rank: int, world_size: int):
    """Samples [start, stop) of the global batch that rank ``rank`` processes."""
    assert world_size >= 1 and 0 <= rank < world_size, "bad rank / world_size"
    assert global_batch % world_size == 0, \
        "global batch %d must be divisible by the number of replicas %d (equal shards keep the " \
        "mean-of-means equal to the global mean)" % (global_batch, world_size)
    per = global_batch // world_size
    return rank * per, (rank + 1) * per


def shard_batch(batch_imgs, rank: int, world_size: int):
    """The rank's slice of a (B, H, W, 6) global batch (a view; sample order is preserved)."""
    a, b = shard_range(batch_imgs.shape[0], rank, world_size)
    return batch_imgs[a:b]


class TorchDistComm:
    """Gradient exchange over an initialised torch.distributed process group."""

    def __init__(self, group=None):
        import torch.distributed as dist
        assert dist.is_initialized(), "torch.distributed is not initialised"
        self.dist = dist
        self.group = group
        self.world_size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def allreduce_sum_(self, flat):
        self.dist.all_reduce(flat, op=self.dist.ReduceOp.SUM, group=self.group)
        return flat

    def destroy(self):
        pass


class CAbiComm:
    """The library's NCCL communicator (csrc/comm.cu).  The 128-byte unique id is created on rank 0
    and handed to the other ranks by ``broadcast_bytes(bytes_or_None) -> bytes`` (any side channel:
    torch.distributed, MPI, a file)."""

    def __init__(self, rank: int, world_size: int, broadcast_bytes):
        from . import _lib as L
        self.L = L
        self.rank, self.world_size = rank, world_size
        raw = None
        if rank == 0:
            buf = (ctypes.c_char * 128)()
            L.check(L.load().of_comm_unique_id(buf), "of_comm_unique_id")
            raw = bytes(buf.raw)
        raw = broadcast_bytes(raw)
        assert isinstance(raw, (bytes, bytearray)) and len(raw) == 128
        out = ctypes.c_void_p()
        L.check(L.load().of_comm_init(bytes(raw), world_size, rank, ctypes.byref(out)), "of_comm_init")
        self.handle = out

    @classmethod
    def from_torch_dist(cls, device):
        """Bootstrap over an existing torch.distributed group (as bench.py does under torchrun)."""
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()

        def bcast(raw):
            t = torch.zeros(128, dtype=torch.uint8)
            if raw is not None:
                t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).clone()
            t = t.to(device)
            dist.broadcast(t, 0)
            return t.cpu().numpy().tobytes()

        return cls(rank, world, bcast)

    def allreduce_sum_(self, flat):
        L = self.L
        assert flat.dtype == torch.float32 and flat.is_contiguous()
        L.check(L.load().of_comm_allreduce_sum_f32(self.handle, L.ptr(flat), flat.numel(), L.stream_ptr()),
                "of_comm_allreduce_sum_f32")
        return flat

    def destroy(self):
        if self.handle is not None:
            self.L.check(self.L.load().of_comm_destroy(self.handle), "of_comm_destroy")
            self.handle = None


def exchange_gradients(flat_grad, comm):
    """Sum the replicas' flat gradients in place; returns the scale (1/N) the optimizer must apply so
    the update uses the global-batch mean gradient.  ``comm`` None = single replica."""
    if comm is None or comm.world_size == 1:
        return 1.0
    comm.allreduce_sum_(flat_grad)
    return 1.0 / comm.world_size


def exchange_gradients_in_buckets(flat_grad, comm, boundaries):
    """The same exchange as ``exchange_gradients`` issued bucket by bucket, LAST bucket first (the order in which the
    backward pass finishes them: decoder, then encoder - Trainer launches the first of these on a side stream under
    the encoder backward).  ``boundaries``: increasing element offsets that split the flat gradient."""
    if comm is None or comm.world_size == 1:
        return 1.0
    edges = [0] + [int(b) for b in boundaries] + [flat_grad.numel()]
    assert all(a < b for a, b in zip(edges, edges[1:])), "bucket boundaries must be strictly increasing and inside"
    for a, b in reversed(list(zip(edges, edges[1:]))):
        comm.allreduce_sum_(flat_grad[a:b])
    return 1.0 / comm.world_size
