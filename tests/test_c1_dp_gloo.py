"""The N > 1 path on CPU: world_size-2 ``gloo`` processes run the data-parallel host logic
(optical_flow_b200/parallel.py: equal sample shards, one sum-allreduce of the flat gradient, 1/N
folded into Adam) with the oracle as the per-replica compute (SURVEY 8c known-answer test 12):

    N-replica averaged gradient  ==  single-replica gradient on the concatenated batch,

and after the Adam step every replica holds the same weights as the single-replica run.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from optical_flow_b200 import parallel  # noqa: E402
from oracle import ref_torch as T  # noqa: E402

B_GLOBAL, H, W = 2, 64, 128


def _flat(grads):
    return torch.cat([g.reshape(-1) for g in grads.values()]).contiguous()


def _single_replica():
    params, stats = T.init_params(seed=42, randomize_bn=True)
    net = T.FlowNetRef(params, stats)
    imgs = torch.tensor(T.synthetic_pairs(B_GLOBAL, H, W, seed=3))
    loss, _, grads = net.loss_and_grads(imgs)
    return loss, _flat(grads), _flat(net.p).detach().clone()


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = parallel.TorchDistComm()
        params, stats = T.init_params(seed=42, randomize_bn=True)       # same seed on every replica
        net = T.FlowNetRef(params, stats)
        global_batch = torch.tensor(T.synthetic_pairs(B_GLOBAL, H, W, seed=3))
        mine = parallel.shard_batch(global_batch, rank, world)
        assert mine.shape[0] == B_GLOBAL // world
        loss, _, grads = net.loss_and_grads(mine)
        flat = _flat(grads)
        bucketed = flat.clone()
        scale = parallel.exchange_gradients(flat, comm)                  # in-place sum over replicas
        assert scale == 1.0 / world
        # two buckets (decoder first, like Trainer's overlapped exchange) give the same sums
        assert parallel.exchange_gradients_in_buckets(bucketed, comm, [flat.numel() * 9 // 16]) == scale
        assert torch.equal(bucketed, flat)
        loss_sum = loss.clone().reshape(1)
        dist.all_reduce(loss_sum)
        w = _flat(net.p).detach().clone()
        new_w, _, _ = T.adam_step(w, flat * scale, torch.zeros_like(w), torch.zeros_like(w), 1)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), grad=(flat * scale).numpy(), w=new_w.numpy(),
                 loss=(loss_sum / world).numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_range_and_preconditions():
    assert parallel.shard_range(64, 0, 8) == (0, 8)
    assert parallel.shard_range(64, 7, 8) == (56, 64)
    covered = [i for r in range(4) for i in range(*parallel.shard_range(32, r, 4))]
    assert covered == list(range(32))
    with pytest.raises(AssertionError):
        parallel.shard_range(10, 0, 4)                 # unequal shards would bias the mean
    with pytest.raises(AssertionError):
        parallel.shard_range(8, 4, 4)
    assert parallel.exchange_gradients(torch.ones(3), None) == 1.0


@pytest.mark.timeout(600)
def test_two_replicas_equal_one_replica_on_the_concatenated_batch(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    loss1, grad1, w0 = _single_replica()
    new_w1, _, _ = T.adam_step(w0, grad1, torch.zeros_like(w0), torch.zeros_like(w0), 1)
    r0 = np.load(tmp_path / "rank0.npz")
    r1 = np.load(tmp_path / "rank1.npz")
    # replicas agree bit for bit with each other (same reduction result, same update)
    assert np.array_equal(r0["grad"], r1["grad"]) and np.array_equal(r0["w"], r1["w"])
    # and with the single replica on the whole batch up to fp32 effects: summation order, the conv
    # library's batch-dependent blocking, and the few pixels whose |diff| ~ 0 flips sign(diff) (L1 loss)
    g1 = grad1.numpy()
    assert np.abs(r0["grad"] - g1).max() <= 5e-3 * np.abs(g1).max()
    assert np.linalg.norm(r0["grad"] - g1) <= 1e-3 * np.linalg.norm(g1)
    assert abs(float(r0["loss"][0]) - float(loss1)) <= 1e-6 * abs(float(loss1))
    # Adam's first step is +-lr wherever |g| >> eps; compare where the gradient is not tiny
    big = np.abs(g1) > 1e-2 * np.abs(g1).max()
    assert np.abs(r0["w"] - new_w1.numpy())[big].max() <= 2e-6
