"""Mirror of the reference's train.py: ``train_step(batch_imgs, step_count)`` ->
``(loss_value, flows)`` (train.py:47-61) plus a thin epoch driver.

The reference is a module-level script (constants at import time, train.py:15-27);
here the same constants are defaults of ``Trainer`` / ``main()``.  Data-parallel
replicas (one process per GPU) all-reduce the flat gradient with NCCL before Adam.
"""
from __future__ import annotations

import argparse
import os
import time

import torch

from . import _lib as L
from . import functional as F
from . import parallel
from .loss import LossLayer
from .model import build_flow_net

img_height = 192            # train.py:15
img_width = 640             # train.py:16
batch_size = 4              # train.py:22
nepochs = 20                # train.py:25
learning_rate = 1e-4        # train.py:34


def regularized_parameters(specs):
    """Names of the tensors the reference attaches l2(l2_reg) to: kernel and bias of the stem (model.py:12-13) and of
    every residual block's conv1 / conv2 (resnet/models.py:17-23); not the shortcuts, BN or the decoder."""
    return [n for n in specs
            if (n.startswith("conv1/") or n.split("/")[0].endswith(("_conv1", "_conv2"))) and not n.startswith("flow")
            and n.endswith(("/kernel", "/bias"))]


class Trainer:
    def __init__(self, height=img_height, width=img_width, pretrained_weights_path=None, lr=learning_rate,
                 conv_mode="bf16", warp_mode="ref_ij", comm=None, seed=42, apply_l2=False, l2_reg=1e-4,
                 bn_mode="inference", fuse_residual_epilogues=True):
        """``comm``: None (single replica) or a parallel.CAbiComm / parallel.TorchDistComm; every
        replica must be built with the same ``seed`` so the weights start identical.
        ``apply_l2`` (opt-in, default off = the reference's behaviour): add the L2 penalty the reference
        DECLARES on the encoder convs' kernels and biases (model.py:7,13; resnet/models.py:5,18,23) but never adds to
        its loss (train.py:50-55 ignores model.losses).
        ``bn_mode`` (opt-in, default "inference" = the reference): "train" runs every BatchNormalization on
        batch statistics (old/train.py:59 ``training=True`` semantics) with cross-replica sums over ``comm``."""
        self.flow_net = build_flow_net(height, width, pretrained_weights_path, conv_mode=conv_mode,
                                       warp_mode=warp_mode, seed=seed, bn_mode=bn_mode,
                                       bn_comm=comm if bn_mode == "train" else None,
                                       fuse_residual_epilogues=fuse_residual_epilogues)
        self.loss_layer = LossLayer(warp_mode=warp_mode)
        self._lr = float(lr)
        self.step = 0
        self._graph = None              # (input shape, static input, CUDAGraph, loss, flows) once enable_graph() ran
        self._opt_state = None          # device of_adam_state {int step; float lr} (graph mode)
        self.comm = comm
        self.world_size = comm.world_size if comm is not None else 1
        # data-parallel gradient exchange in two buckets: the decoder's 2 153 736 gradients are final
        # before the encoder's backward starts, so their all-reduce runs on a side stream under it; the encoder
        # bucket follows on the compute stream
        self._dec_off = None
        self._comm_stream = None
        if comm is not None and self.world_size > 1:
            st = self.flow_net.store
            first = min((st.g[n].data_ptr() for n in st.specs if n.startswith("flow")))
            self._dec_off = (first - st.grad.data_ptr()) // 4
            assert all((st.g[n].data_ptr() - st.grad.data_ptr()) // 4 >= self._dec_off
                       for n in st.specs if n.startswith("flow"))
            assert all((st.g[n].data_ptr() - st.grad.data_ptr()) // 4 < self._dec_off
                       for n in st.specs if not n.startswith("flow"))
            self._comm_stream = torch.cuda.Stream(device=self.flow_net.device)
            self._dec_ready = torch.cuda.Event()
            self._dec_done = torch.cuda.Event()
        self._dec_in_flight = False
        self.apply_l2, self.l2_reg = bool(apply_l2), float(l2_reg)

    @property
    def trainable_weights(self):
        return self.flow_net.trainable_weights

    @property
    def lr(self):
        return self._lr

    @lr.setter
    def lr(self, value):
        """train.py:69-71 changes the learning rate mid-run; in graph mode it lives in device memory."""
        self._lr = float(value)
        if self._opt_state is not None:
            self._write_opt_state()

    def _write_opt_state(self):
        import struct
        raw = torch.frombuffer(bytearray(struct.pack("<if", int(self.step), self._lr)), dtype=torch.uint8)
        self._opt_state.copy_(raw.to(self._opt_state.device))

    def forward_backward(self, batch_imgs):
        """tape.gradient(loss, trainable_weights) of train.py:50-55: loss + flat gradient."""
        flows = self.flow_net(batch_imgs)
        loss_value = self.loss_layer(batch_imgs, flows)            # also yields d loss / d flow_s
        self.flow_net.backward(self.loss_layer.d_flows,
                               after_decoder=self._exchange_decoder_bucket if self._comm_stream is not None else None)
        if self.apply_l2:
            st = self.flow_net.store
            for name in regularized_parameters(st.specs):
                F.l2_regularizer(st.p[name], st.g[name], self.l2_reg, loss_value)
        return loss_value, flows

    def _exchange_decoder_bucket(self):
        cur = torch.cuda.current_stream(self.flow_net.device)
        self._dec_ready.record(cur)
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(self._dec_ready)
            self.comm.allreduce_sum_(self.flow_net.store.grad[self._dec_off:])
            self._dec_done.record(self._comm_stream)
        self._dec_in_flight = True

    def _exchange_gradients(self):
        """Sum the replicas' gradients; returns the 1/N the optimizer applies.  With the decoder bucket already in
        flight only the encoder bucket is exchanged here, then the compute stream joins the side stream."""
        st = self.flow_net.store
        if self._dec_in_flight:
            self._dec_in_flight = False
            self.comm.allreduce_sum_(st.grad[:self._dec_off])
            torch.cuda.current_stream(self.flow_net.device).wait_event(self._dec_done)
            return 1.0 / self.world_size
        return parallel.exchange_gradients(st.grad, self.comm)

    def apply_gradients(self):
        """optimizer.apply_gradients (train.py:56): Adam, Keras defaults (eps 1e-7)."""
        st = self.flow_net.store
        scale = self._exchange_gradients()                         # sum over replicas; 1/N goes into Adam
        self.step += 1
        if self._opt_state is not None:                            # graph mode: the device holds (step, lr)
            F.adam_step_dev(st.flat, st.grad, st.adam_m, st.adam_v, self._opt_state, grad_scale=scale)
        else:
            F.adam_step(st.flat, st.grad, st.adam_m, st.adam_v, self.step, lr=self.lr, grad_scale=scale)

    def train_step(self, batch_imgs, step_count=None):
        """train.py:47-61.  Returns (loss_value [0-d device tensor], flows).  After ``enable_graph`` a batch of the
        captured shape is ONE graph launch (train.py:47's @tf.function): forward, loss, backward, gradient exchange
        and Adam; the returned tensors are the graph's static outputs (overwritten by the next step)."""
        g = self._graph
        if g is not None and tuple(batch_imgs.shape) == g[0]:
            if batch_imgs.data_ptr() != g[1].data_ptr():
                g[1].copy_(batch_imgs, non_blocking=True)
            g[2].replay()
            self.step += 1
            return g[3], g[4]
        loss_value, flows = self.forward_backward(batch_imgs)
        self.apply_gradients()
        return loss_value, flows

    def enable_graph(self, sample_batch, warmup=2):
        """Capture the whole step for batches shaped like ``sample_batch`` (B, H, W, 6) into a CUDA graph.  Everything
        that changes from step to step is device state: the input buffer, the weights / Adam moments, and the
        optimiser's (step, lr) pair (of_adam_step_dev), so replaying the graph IS the next step.  ``warmup`` eager
        forward + backward passes over the sample run first on a side stream (lazy allocations, tensor-map encodes,
        function attributes); they compute gradients but apply nothing."""
        dev = self.flow_net.device
        st = self.flow_net.store
        self._opt_state = torch.zeros(8, dtype=torch.uint8, device=dev)
        static_in = sample_batch.detach().clone().contiguous()

        def step_dev():
            loss_value, flows = self.forward_backward(static_in)
            scale = self._exchange_gradients()
            F.adam_step_dev(st.flat, st.grad, st.adam_m, st.adam_v, self._opt_state, grad_scale=scale)
            return loss_value, flows

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            self._write_opt_state()
            for _ in range(max(warmup, 1)):
                self.forward_backward(static_in)
                self._exchange_gradients()
            graph = torch.cuda.CUDAGraph()
            n0 = L.load().of_launch_count()
            with torch.cuda.graph(graph, stream=side):
                loss_value, flows = step_dev()
            self.graph_kernels = int(L.load().of_launch_count() - n0)      # this library's kernels in one replay
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._graph = (tuple(static_in.shape), static_in, graph, loss_value, flows)
        return static_in

    def graph_input(self):
        """The captured step's static input buffer (fill it in place to skip the device-to-device copy)."""
        return self._graph[1] if self._graph is not None else None


class HostBatchPrefetcher:
    """Feeds ``Trainer.train_step`` from HOST batches (what data_reader.py:35-64 yields): two device
    buffers and a copy stream, so the pinned host -> device copy of batch i+1 runs under the compute
    of batch i.  ``batches`` is any iterable of pinned (B,H,W,6) float32 CPU tensors."""

    def __init__(self, batches, device):
        self.it = iter(batches)
        self.device = device
        self.copy_stream = torch.cuda.Stream(device=device)
        self.bufs = [None, None]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.k = 0
        self.pending = None
        self._issue()

    def _issue(self):
        try:
            h = next(self.it)
        except StopIteration:
            self.pending = None
            return
        slot = self.k % 2
        if self.bufs[slot] is None or self.bufs[slot].shape != h.shape:
            self.bufs[slot] = torch.empty(h.shape, dtype=h.dtype, device=self.device)
            self.free[slot].record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.free[slot])       # the step that last read this buffer is done
            self.bufs[slot].copy_(h, non_blocking=True)
            self.ready[slot].record(self.copy_stream)
        self.pending = slot
        self.k += 1

    def __iter__(self):
        return self

    def __next__(self):
        """Returns (device_batch, release): call ``release()`` after launching the step that reads it."""
        if self.pending is None:
            raise StopIteration
        slot = self.pending
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.ready[slot])
        self._issue()                                          # next copy goes out before this step's kernels
        return self.bufs[slot], (lambda: self.free[slot].record(cur))


class HostFramePrefetcher(HostBatchPrefetcher):
    """The data path of data_reader.py:35-64 feeding ``Trainer.train_step``: ``batches`` yields pinned uint8 CPU
    tensors (B, 2, h, w, 3) of decoded frame pairs (6 bytes per pixel pair over PCIe instead of 24 bytes of float32);
    the copy of batch i+1 runs on a copy stream under the compute of batch i, and one of_preprocess_batch launch on
    the compute stream turns the frames into the (B, H, W, 6) float32 network input (resize, / 255, BGR means)."""

    def __init__(self, batches, device, height, width, out=None):
        """``out``: write every batch into this (B, H, W, 6) tensor instead of two rotating buffers - pass
        ``Trainer.graph_input()`` so a graph-captured step reads its input where the preprocessing kernel left it."""
        from .data_reader import ReaderOpts
        self.opts = ReaderOpts(img_height=height, img_width=width)
        self.net_in = [out, out] if out is not None else [None, None]
        super().__init__(batches, device)

    def __next__(self):
        from .data_reader import read_batch_frames
        slot = self.pending                                   # the buffer pair this batch uses (None -> StopIteration)
        frames, release = super().__next__()
        B = frames.shape[0]
        if self.net_in[slot] is None or self.net_in[slot].shape[0] != B:
            self.net_in[slot] = torch.empty((B, self.opts.img_height, self.opts.img_width, 6), dtype=torch.float32,
                                            device=self.device)
        x = read_batch_frames(frames, self.opts, out=self.net_in[slot])
        release()                                             # the uint8 buffer is free once the kernel is queued
        return x, (lambda: None)


_default_trainer = None


def train_step(batch_imgs, step_count=None):
    """Module-level entry point with the reference's signature (train.py:48)."""
    global _default_trainer
    if _default_trainer is None:
        _default_trainer = Trainer(batch_imgs.shape[1], batch_imgs.shape[2])
    return _default_trainer.train_step(batch_imgs, step_count)


def synthetic_batch(batch, height, width, seed, device):
    """Seeded synthetic image pairs in the value range data_reader.py:53-63 produces
    (pixels/255 minus the BGR mean); frame 2 is frame 1 shifted by a few pixels."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    yy = torch.arange(height, dtype=torch.float32).view(1, height, 1, 1)
    xx = torch.arange(width, dtype=torch.float32).view(1, 1, width, 1)
    im1 = torch.zeros(batch, height, width, 3)
    im2 = torch.zeros(batch, height, width, 3)
    shift = (torch.rand(batch, 2, generator=g) * 16 - 8).view(batch, 1, 1, 2)
    for _ in range(8):
        fx = 0.005 + 0.145 * torch.rand(batch, 1, 1, 3, generator=g)
        fy = 0.005 + 0.145 * torch.rand(batch, 1, 1, 3, generator=g)
        ph = 6.2831853 * torch.rand(batch, 1, 1, 3, generator=g)
        amp = 0.03 + 0.09 * torch.rand(batch, 1, 1, 3, generator=g)
        im1 += amp * torch.sin(6.2831853 * (fx * xx + fy * yy) + ph)
        im2 += amp * torch.sin(6.2831853 * (fx * (xx + shift[..., :1]) + fy * (yy + shift[..., 1:])) + ph)
    means = torch.tensor([104.0, 117.0, 123.0]).view(1, 1, 1, 3) / 255.0
    im1 = (im1 + 0.5 + 0.02 * torch.randn(im1.shape, generator=g)).clamp(0, 1) - means
    im2 = (im2 + 0.5 + 0.01 * torch.randn(im2.shape, generator=g)).clamp(0, 1) - means
    return torch.cat([im1, im2], dim=-1).contiguous().to(device)


def main(argv=None):
    ap = argparse.ArgumentParser(description="optical-flow training (synthetic pairs; mirrors train.py:64-88)")
    ap.add_argument("--height", type=int, default=img_height)
    ap.add_argument("--width", type=int, default=img_width)
    ap.add_argument("--batch-size", type=int, default=batch_size)
    ap.add_argument("--epochs", type=int, default=nepochs)
    ap.add_argument("--batches-per-epoch", type=int, default=50)
    ap.add_argument("--conv-mode", default="bf16", choices=["bf16", "fp32", "bf16x3"])
    ap.add_argument("--warp-mode", default="ref_ij", choices=["ref_ij", "xy"])
    ap.add_argument("--save-dir", default="ckpts")
    args = ap.parse_args(argv)
    dev = torch.device("cuda", 0)
    trainer = Trainer(args.height, args.width, None, conv_mode=args.conv_mode, warp_mode=args.warp_mode)
    trainer.flow_net.summary()
    step_count = 0
    for epoch in range(args.epochs):
        t0 = time.time()
        if epoch == 15:                                   # train.py:69-71
            trainer.lr = 1e-5
        for batch_idx in range(args.batches_per_epoch):
            batch = synthetic_batch(args.batch_size, args.height, args.width, 1234 + step_count, dev)
            loss_value, _ = trainer.train_step(batch, step_count)
            if (batch_idx + 1) % 10 == 0:                 # loss stays on device between reads
                print("epoch %d batch %d/%d loss %.4e" % (epoch + 1, batch_idx + 1, args.batches_per_epoch,
                                                          loss_value.item()))
            step_count += 1
        print("Epoch computed in %.1fs" % (time.time() - t0))
        os.makedirs(args.save_dir, exist_ok=True)
        trainer.flow_net.save_weights(os.path.join(args.save_dir, "flow_net_%d" % epoch))   # train.py:88


if __name__ == "__main__":
    main()
