#!/usr/bin/env python
"""bench.py - headline benchmark of the optical-flow training hot path on B200.

  python bench.py --gpus N --steps K --warmup W [--workload train|loss] [--impl reference]

Workloads (BASELINE.json configs):
  train  configs[2]/[3]: full unsupervised train step (encoder/decoder fwd+bwd, fused
         warp+L1 pyramid loss, Adam), batch 8 per GPU at 384x1280 (KITTI shape), synthetic
         pairs, random-init weights.  metric = image-pairs/s (whole job).
  loss   configs[1]: fused warp + photometric pyramid loss kernel only (fwd + d_flow),
         batch 8 at 384x1280.  metric = image-pairs/s through the kernel; roofline = HBM.

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for the definitions.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H_FULL, W_FULL, B_PER_GPU = 384, 1280, 8


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


# --------------------------------------------------------------- algorithmic work --
def loss_algorithmic_bytes(B, H, W, ns=4):
    """read image pair once + read flows + write d_flows + loss scalar."""
    px = sum(B * int(H / 2 ** (s + 1)) * int(W / 2 ** (s + 1)) for s in range(ns))
    return 4 * (6 * B * H * W + 2 * px + 2 * px) + 4


def conv_layer_table(B, H, W):
    """(name, N, Ho, Wo, Cin, k, Cout, stride) of the 39 convs."""
    L = [("conv1", 2 * B, H // 2, W // 2, 3, 7, 64, 2)]
    cin, h, w = 64, H // 4, W // 4
    for layer_idx, down in ((2, False), (3, True), (4, True)):
        for blk in (1, 2):
            ds = down and blk == 1
            cout = 2 * cin if ds else cin
            if ds:
                h, w = h // 2, w // 2
                L.append(("layer%d_block_%d_shortcut" % (layer_idx, blk), 2 * B, h, w, cin, 1, cout, 2))
            L.append(("layer%d_block_%d_conv1" % (layer_idx, blk), 2 * B, h, w, cin, 3, cout, 2 if ds else 1))
            L.append(("layer%d_block_%d_conv2" % (layer_idx, blk), 2 * B, h, w, cout, 3, cout, 1))
            cin = cout
    feat = (256, 128, 64, 64)
    for lvl in range(4):
        h, w = H // (16 >> lvl), W // (16 >> lvl)
        c = feat[lvl] + 49 + (2 if lvl else 0)
        for ci, cout in enumerate((128, 128, 96, 64, 32, 2)):
            L.append(("flow%d_conv%d" % (lvl, ci + 1), B, h, w, c, 3, cout, 1))
            c = cout
    return L


def conv_flops(B, H, W, tc_only=False):
    """Algorithmic conv FLOPs of one train step: fprop + dgrad + wgrad (no dgrad for conv1)."""
    tot = 0.0
    for name, n, ho, wo, cin, k, cout, s in conv_layer_table(B, H, W):
        if tc_only and (cin < 32 or cout < 32):
            continue
        f = 2.0 * n * ho * wo * cin * cout * k * k
        tot += f * (2 if name == "conv1" else 3)
    return tot


# ------------------------------------------------------------------------ clocks --
class ClockSampler:
    """nvidia-smi sampled DURING the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def mark(self):
        """Start of the timed region: the sampler was started during the warm-up (nvidia-smi needs a few hundred
        ms before its first line - longer than a 0.2 s timed region); only rows from here on are reported."""
        self.rows = []

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 3 + k and r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm),
                "window": "timed region + the e2e timed region that follows it (same step, under load)"}


# ------------------------------------------------------------------ CPU baselines --
def cpu_loss_baseline(B, budget_s=15.0):
    """Oracle port (torch-CPU, all host threads) of the warp+loss fwd+bwd on a bounded
    sample: the same 384x1280 shape, as many batch-1 passes as fit the budget."""
    import numpy as np
    import torch
    from oracle import ref_torch as T
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    imgs = torch.tensor(T.synthetic_pairs(1, H_FULL, W_FULL, seed=1))
    rng = np.random.default_rng(0)
    flows = [torch.tensor((rng.standard_normal((1, H_FULL >> (s + 1), W_FULL >> (s + 1), 2)) * (4.0 / 2 ** s)).astype(np.float32),
                          requires_grad=True) for s in range(4)]

    def one():
        for f in flows:
            f.grad = None
        T.loss_layer(imgs, flows).backward()

    one()
    t0, n = time.time(), 0
    while True:
        one()
        n += 1
        if time.time() - t0 > budget_s or n >= 50:
            break
    dt = (time.time() - t0) / n
    return {"value": 1.0 / dt, "unit": "image-pairs/s", "cores": cores, "kind": "port",
            "sample": "oracle/ref_torch.loss_layer fwd+bwd, batch 1 at %dx%d, %d passes" % (H_FULL, W_FULL, n)}


def cpu_train_baseline(budget_s=25.0, height=H_FULL, width=W_FULL, max_steps=20, warmup=1, sample_batch=1):
    """Oracle port (torch-CPU fp32 autograd, all host threads) of the full train step - forward, loss, backward AND the
    Adam update of all 4.9 M parameters (train.py:47-56) - on a bounded sample of the SAME workload: ``sample_batch``
    pairs per step at the bench's resolution (a batch-8 tape at 384x1280 needs tens of GB of host memory), as many
    steps as fit the time budget.  value = pairs per second."""
    import torch
    from oracle import ref_torch as T
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    params, stats = T.init_params(seed=42)
    net = T.FlowNetRef(params, stats)
    imgs = torch.tensor(T.synthetic_pairs(sample_batch, height, width, seed=1))
    m = {k: torch.zeros_like(v) for k, v in net.p.items()}
    v = {k: torch.zeros_like(p) for k, p in net.p.items()}
    state = {"step": 0}

    def one():
        _, _, grads = net.loss_and_grads(imgs)
        state["step"] += 1
        with torch.no_grad():
            for k, p in net.p.items():
                pn, m[k], v[k] = T.adam_step(p.detach(), grads[k], m[k], v[k], state["step"])
                p.copy_(pn)

    for _ in range(max(warmup, 1)):
        one()
    t0, n = time.time(), 0
    while True:
        one()
        n += 1
        if time.time() - t0 > budget_s or n >= max_steps:
            break
    dt = (time.time() - t0) / n
    return {"value": sample_batch / dt, "unit": "image-pairs/s", "cores": cores, "kind": "port",
            "sample": "oracle/ref_torch.FlowNetRef fwd+loss+bwd+Adam (torch-CPU fp32, %d threads), %d pair(s) per step at "
                      "%dx%d, %d steps of %.2f s" % (cores, sample_batch, height, width, n, dt)}


def train_metric(H, W):
    return "image-pairs/s per train step at %dx%d" % (H, W)


def per_gpu_batch(args, world):
    """weak scaling: --batch pairs per GPU; strong scaling (BASELINE configs[3] read literally): --global-batch
    pairs over all GPUs (64 -> 32 / 16 / 8 per GPU at 2 / 4 / 8)."""
    if args.scaling == "strong":
        assert args.global_batch % world == 0, "global batch must divide by the number of GPUs"
        return args.global_batch // world
    return args.batch


def train_config(B, H, W, world, conv_mode, scaling="weak"):
    tag = "configs[2]" if scaling == "weak" else "configs[3], fixed global batch"
    if (H, W) == (1024, 2048):
        tag = "configs[4] shape" + ("" if B == 4 else ", batch differs from its 4 per GPU")
    elif (H, W) != (H_FULL, W_FULL) or (scaling == "weak" and B != B_PER_GPU):
        tag = "not a BASELINE config"
    return {"workload": "full train step fwd+bwd+Adam, ResNet encoder + cost-volume decoder, batch %d/GPU at %dx%d (%s)" % (B, H, W, tag),
            "global_batch": B * world, "parallelism": "dp%d" % world, "conv_mode": conv_mode,
            "l2": "activations per step (GBs) far exceed the 126 MB L2; 2 rotating input batches"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  TensorFlow is not
    installable here (no wheel, no network), so this is the oracle port (torch-CPU fp32, all host
    threads) on the same metric / config keys; a step = a bounded sample of that workload (2 pairs of the
    batch at the same resolution, Adam included), value normalised to pairs per second; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(args.steps, 1)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    B = per_gpu_batch(args, world)
    if args.workload == "loss":
        base = cpu_loss_baseline(B, budget_s=min(60.0, 4.0 * steps))
        metric = "image-pairs/s through the fused warp+loss (fwd+bwd) at %dx%d" % (H_FULL, W_FULL)
        cfg = {"workload": "fused warp + L1 pyramid loss fwd+bwd, batch %d/GPU at %dx%d (configs[1])" % (B, args.height, args.width)}
    else:
        base = cpu_train_baseline(budget_s=150.0, height=args.height, width=args.width, max_steps=steps,
                                  warmup=min(args.warmup, 2), sample_batch=min(B, 2))
        metric = train_metric(args.height, args.width)
        cfg = train_config(B, args.height, args.width, world, "bf16", args.scaling)
    line = {"impl": "reference", "metric": metric, "value": base["value"], "unit": "image-pairs/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * B / base["value"],       # normalised to the config's batch (the sample is smaller)
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "image-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# -------------------------------------------------------------------- GPU arms ----
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="train", choices=["train", "loss", "kernels"])
    ap.add_argument("--conv-mode", default="bf16", choices=["bf16", "fp32", "bf16x3"])
    ap.add_argument("--batch", type=int, default=B_PER_GPU)
    ap.add_argument("--height", type=int, default=H_FULL)
    ap.add_argument("--width", type=int, default=W_FULL)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --batch pairs per GPU (default); strong: --global-batch pairs over all GPUs (configs[3])")
    ap.add_argument("--global-batch", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="train workload: skip the distribution / timeline / loss-kernel legs (scaling sweeps)")
    ap.add_argument("--flows", default="smooth", choices=["smooth", "iid"],
                    help="loss workload: flow field statistics (SURVEY 8d: realistic smooth / worst-case i.i.d.)")
    ap.add_argument("--no-graph", action="store_true", help="time direct launches instead of CUDA graph replay")
    ap.add_argument("--graph", default="on", choices=["on", "off"],
                    help="train workload: replay the whole step from one CUDA graph (Trainer.enable_graph) or launch eagerly")
    ap.add_argument("--dump-timeline", default=None, metavar="PATH",
                    help="train workload: write the per-launch records (kernel, us, FLOPs) of the timeline pass as JSON")
    ap.add_argument("--no-fuse-residual", action="store_true",
                    help="FlowNet(fuse_residual_epilogues=False): separate add_relu / act_bwd / accumulate launches (A/B)")
    ap.add_argument("--option", action="append", default=[], metavar="NAME=VALUE",
                    help="kernel-variant option passed to of_set_option (A/B runs; recorded in config.options)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from optical_flow_b200 import build
    if rank == 0:
        build.build_library()
    if world > 1:
        dist.barrier()
    if args.option:
        from optical_flow_b200 import _lib as L
        for kv in args.option:
            name, _, val = kv.partition("=")
            L.check(L.load().of_set_option(name.encode(), (val or "1").encode()), "of_set_option")
    if args.workload == "loss":
        line = bench_loss(args, dev, world, rank)
    elif args.workload == "kernels":
        line = bench_kernels(args, dev, world, rank)
    else:
        line = bench_train(args, dev, world, rank)
    if args.option:
        line["config"]["options"] = list(args.option)
    if rank == 0:
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = (cpu_loss_baseline(args.batch) if args.workload in ("loss", "kernels") else
                                    cpu_train_baseline(budget_s=20.0, height=args.height, width=args.width, max_steps=12,
                                                       sample_batch=min(per_gpu_batch(args, world), 2)))
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _timed(torch, dist, world, fn_steps, steps):
    """barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn_steps(steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def make_loss_sets(torch, dev, B, H, W, flows_kind, rank, nsets=4):
    """nsets x (images, 4 flows, 4 d_flows, loss): 4 x 115 MB of inputs / outputs rotate through the 126 MB L2."""
    from optical_flow_b200.train import synthetic_batch
    sets = []
    for k in range(nsets):
        imgs = synthetic_batch(B, H, W, 1234 + rank * 100 + k, dev)
        g = torch.Generator(device="cpu").manual_seed(77 + k)
        flows = []
        for s in range(4):
            h, w = H >> (s + 1), W >> (s + 1)
            if flows_kind == "iid":
                # SURVEY 8d worst case: i.i.d. N(0, (4/2^s)^2) px per pixel (every tap a different line)
                f = torch.randn((B, h, w, 2), generator=g) * (4.0 / 2 ** s)
            else:
                # SURVEY 8d realistic variant: the same amplitude, low-pass filtered (what a flow
                # decoder emits): N(0,1) on a 1/8-resolution grid, bilinearly upsampled
                c = torch.randn((B, 2, max(h // 8, 2), max(w // 8, 2)), generator=g)
                f = torch.nn.functional.interpolate(c, size=(h, w), mode="bilinear", align_corners=False)
                f = (f * (4.0 / 2 ** s) / f.std()).permute(0, 2, 3, 1).contiguous()
            flows.append(f.to(dev))
        d_flows = [torch.empty_like(f) for f in flows]
        sets.append((imgs, flows, d_flows, torch.empty(1, device=dev)))
    return sets


def loss_runner(torch, sets, use_graph=True):
    """Returns (run(n), launch mode): the kernel (~25-50 us) is shorter than a Python + ctypes launch, so the
    len(sets) launches are replayed from a CUDA graph - the device, not the host, is what is timed."""
    from optical_flow_b200 import functional as F
    nsets = len(sets)

    def run(n):
        for i in range(n):
            imgs, flows, d_flows, lo = sets[i % nsets]
            F.loss_l1_pyramid(imgs, flows, need_grads=True, loss_out=lo, d_flows=d_flows)

    run(max(3, nsets))
    graph = None
    if use_graph:
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                run(nsets)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=side):
                    run(nsets)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
        except Exception as e:          # noqa: BLE001 - fall back to direct launches, say so
            graph = None
            sys.stderr.write("CUDA graph capture failed (%s); timing direct launches\n" % e)
    if graph is None:
        return run, "direct"

    def run_graph(n):
        for _ in range(max(n // nsets, 1)):
            graph.replay()

    return run_graph, "cuda graph replay"


def loss_kernel_leg(torch, dev, B, H, W, rank, peaks, iters=200):
    """configs[1] inside the default line: the fused warp+loss kernel alone, both flow statistics, CUDA events around
    `iters` graph-replayed launches over 4 rotating input sets (> L2)."""
    out = {}
    alg = loss_algorithmic_bytes(B, H, W)
    for kind in ("smooth", "iid"):
        sets = make_loss_sets(torch, dev, B, H, W, kind, rank)
        run, mode = loss_runner(torch, sets)
        n = max(iters // len(sets), 1) * len(sets)
        run(len(sets) * 3)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(n)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / n * 1e3
        ach = alg / (us * 1e-6) / 1e9
        out[kind] = {"us": us, "achieved": ach, "frac": ach / peaks["hbm"], "launches": n, "launch": mode}
        del sets
    return {"bound": "hbm", "unit": "GB/s", "peak": peaks["hbm"], "peak_source": peaks["src"],
            "algorithmic_bytes": alg, "kernel": "loss_fused_kernel (warp + L1 pyramid loss fwd + d_flow, one launch)",
            "workload": "batch %d at %dx%d (configs[1]), 4 rotating input sets (> 126 MB L2)" % (B, H, W),
            "flows": out, "achieved": out["smooth"]["achieved"], "frac": out["smooth"]["frac"],
            "note": "headline = low-pass (decoder-like) flows; iid = SURVEY 8d worst-case scatter"}


def bench_loss(args, dev, world, rank):
    import torch
    import torch.distributed as dist
    from optical_flow_b200 import functional as F
    B, H, W = args.batch, args.height, args.width
    peaks = measured_peaks()
    sets = make_loss_sets(torch, dev, B, H, W, args.flows, rank)
    nsets = len(sets)
    run_timed, mode = loss_runner(torch, sets, use_graph=not args.no_graph)
    steps = max(args.steps // nsets, 1) * nsets
    clocks = ClockSampler(dev.index or 0)
    clocks.start()
    run_timed(nsets * 3)
    clocks.mark()
    ms = _timed(torch, dist, world, run_timed, steps)
    per = ms / steps
    value = world * B / (per * 1e-3)
    alg = loss_algorithmic_bytes(B, H, W)
    ach = alg / (per * 1e-3) / 1e9
    # end to end: host (pinned) images + flows -> device, kernel, loss + d_flows back to host
    h_imgs = torch.empty((B, H, W, 6), dtype=torch.float32).pin_memory()
    h_imgs.copy_(sets[0][0].cpu())
    h_flows = [f.cpu().pin_memory() for f in sets[0][1]]
    h_dflows = [torch.empty_like(f).pin_memory() for f in h_flows]
    h_loss = torch.empty(1).pin_memory()
    d_imgs, d_fl, d_dfl, d_lo = sets[1]

    def run_e2e(n):
        for _ in range(n):
            d_imgs.copy_(h_imgs, non_blocking=True)
            for a, b in zip(d_fl, h_flows):
                a.copy_(b, non_blocking=True)
            F.loss_l1_pyramid(d_imgs, d_fl, need_grads=True, loss_out=d_lo, d_flows=d_dfl)
            h_loss.copy_(d_lo, non_blocking=True)
            for a, b in zip(h_dflows, d_dfl):
                a.copy_(b, non_blocking=True)
            torch.cuda.current_stream().synchronize()

    run_e2e(2)
    ms_e = _timed(torch, dist, world, run_e2e, args.steps)
    ck = clocks.stop()
    fl_bytes = sum(f.numel() * 4 for f in h_flows)
    e2e = {"value": world * B / (ms_e / args.steps * 1e-3), "unit": "image-pairs/s",
           "h2d_bytes_per_step": h_imgs.numel() * 4 + fl_bytes, "d2h_bytes_per_step": fl_bytes + 4}
    return {"metric": "image-pairs/s through fused warp+loss (fwd + d_flow), %dx%d" % (H, W), "value": value,
            "unit": "image-pairs/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": per, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "fused warp + L1 pyramid loss fwd+bwd, batch %d/GPU at %dx%d (configs[1])" % (B, H, W),
                       "l2": "4 rotating input/output sets (4 x %.0f MB > 126 MB L2)" % (alg / 1e6),
                       "launch": mode,
                       "flows": args.flows + (" (low-pass N(0,(4/2^s)^2) px)" if args.flows == "smooth" else " (i.i.d. N(0,(4/2^s)^2) px, worst-case scatter)")},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peaks["hbm"], "unit": "GB/s",
                         "frac": ach / peaks["hbm"], "traffic": None, "kernel": "loss_fused_kernel",
                         "algorithmic_bytes": alg, "peak_source": peaks["src"]},
            "e2e": e2e, "gpu_launches": steps, "clocks": ck}


def _time_us(torch, fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def bench_kernels(args, dev, world, rank):
    """Stand-alone HBM-bound kernels of SURVEY 8a at their C2 shapes, each against its algorithmic bytes:
    K2 / K3 feature warp fwd / bwd (model.py:55-73) at the three decoder levels that warp, fp32 and bf16 storage;
    K1 absolute-coordinate sampler (transformations.py:85-129); K4 UnFlow image warp (old/image_warp.py:7-78);
    K7 MyLoss photometric + smoothness (old/loss.py:40-105, 5 scales).  Inputs exceed or rotate through L2."""
    import torch
    from optical_flow_b200 import functional as F
    B, H, W = args.batch, args.height, args.width
    peaks = measured_peaks()
    iters = max(args.steps, 10)
    rows = {}
    g = torch.Generator(device="cpu").manual_seed(5)

    def smooth_flow(h, w, amp):
        c = torch.randn((B, 2, max(h // 8, 2), max(w // 8, 2)), generator=g)
        f = torch.nn.functional.interpolate(c, size=(h, w), mode="bilinear", align_corners=False)
        return (f * amp / f.std()).permute(0, 2, 3, 1).contiguous().to(dev)

    def add(name, us, nbytes):
        ach = nbytes / (us * 1e-6) / 1e9
        rows[name] = {"us": round(us, 2), "algorithmic_bytes": int(nbytes), "achieved": round(ach, 1),
                      "frac": round(ach / peaks["hbm"], 4)}

    for (lvl, h, w, C) in ((3, H // 2, W // 2, 64), (2, H // 4, W // 4, 64), (1, H // 8, W // 8, 128)):
        flow = smooth_flow(h, w, 4.0 / 2 ** (3 - lvl))
        for dt, es in ((torch.float32, 4), (torch.bfloat16, 2)):
            feats = torch.randn((B, h, w, C), device=dev).to(dt)
            d_out = torch.randn((B, h, w, C), device=dev).to(dt)
            tag = "level%d_%dx%dx%d_%s" % (lvl, h, w, C, "f32" if es == 4 else "bf16")
            us = _time_us(torch, lambda: F.warp_features_fwd(flow, feats), iters)
            add("K2_warp_features_fwd_" + tag, us, B * h * w * (2 * C * es + 8))
            us = _time_us(torch, lambda: F.warp_features_bwd(flow, feats, d_out), iters)
            add("K3_warp_features_bwd_" + tag, us, B * h * w * (2 * C * es + C * 4 + 16))   # d_features is f32
        del feats, d_out
    h, w, C = H // 2, W // 2, 64
    feats = torch.randn((B, h, w, C), device=dev)
    pts = (smooth_flow(h, w, 4.0) + torch.stack(torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy"), -1)
           .to(dev).float())
    add("K1_bilinear_sample_fwd_%dx%dx%d_f32" % (h, w, C), _time_us(torch, lambda: F.bilinear_sample_fwd(feats, pts), iters),
        B * h * w * 4 * (2 * C + 2))
    del feats, pts
    im = torch.rand((B, H, W, 3), device=dev)
    fl = smooth_flow(H, W, 8.0)
    add("K4_image_warp_fwd_%dx%dx3" % (H, W), _time_us(torch, lambda: F.image_warp_fwd(im, fl), iters), B * H * W * 4 * (2 * 3 + 2))
    d_im = torch.rand((B, H, W, 3), device=dev)
    add("K4_image_warp_bwd_%dx%dx3" % (H, W), _time_us(torch, lambda: F.image_warp_bwd(im, fl, d_im), iters),
        B * H * W * 4 * (3 * 3 + 4))
    del im, fl, d_im
    # K7: MyLoss, 5 scales coarse to fine (old/loss.py:24-25), 4 rotating input sets
    from optical_flow_b200.train import synthetic_batch
    sets = []
    for k in range(4):
        imgs = synthetic_batch(B, H, W, 99 + k, dev)
        flows = [smooth_flow(H // d, W // d, 8.0 / d) for d in (16, 8, 4, 2, 1)]
        sets.append((imgs, flows, [torch.empty_like(f) for f in flows], torch.empty(1, device=dev)))
    state = {"i": 0}

    def k7():
        imgs, flows, d_flows, lo = sets[state["i"] % 4]
        state["i"] += 1
        F.loss_charbonnier_smooth(imgs, flows, need_grads=True, loss_out=lo, d_flows=d_flows)

    px = sum(B * (H // d) * (W // d) for d in (16, 8, 4, 2, 1))
    add("K7_my_loss_fwdbwd_5_scales", _time_us(torch, k7, iters), 4 * (6 * B * H * W + 4 * px))
    worst = min(rows.values(), key=lambda r: r["frac"])
    k3 = rows["K3_warp_features_bwd_level3_%dx%dx64_bf16" % (H // 2, W // 2)]
    return {"metric": "achieved HBM GB/s of the stand-alone warp / loss kernels (K1-K4, K7) at the C2 shapes",
            "value": k3["achieved"], "unit": "GB/s", "n_gpus": world, "steps": iters, "warmup": 3,
            "ms_per_step": k3["us"] * 1e-3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "stand-alone kernels K1-K4, K7, batch %d at %dx%d (value = K3 level-3 bf16 backward)" % (B, H, W),
                       "l2": "each kernel's tensors exceed the 126 MB L2 at level 3; K7 rotates 4 input sets"},
            "roofline": {"bound": "hbm", "achieved": k3["achieved"], "peak": peaks["hbm"], "unit": "GB/s",
                         "frac": k3["frac"], "traffic": None, "kernel": "sample_bwd_kernel (K3, level 3, bf16)",
                         "peak_source": peaks["src"], "kernels": rows,
                         "note": "algorithmic bytes per SURVEY 8d; the scatter's zero-fill and read-modify-write are overhead"},
            "e2e": {"value": None, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "kernel-only leg; the end-to-end number is the train workload's"},
            "gpu_launches": len(rows) * iters, "clocks": None}


def time_dominant_conv(torch, dev, N, H, W, Cin, Cout, iters=20):
    """One 3x3 / stride-1 tensor-core conv layer through the C ABI (of_conv_tc_run), CUDA events on
    the launching stream.  Input + output (N*H*W*(Cin+Cout)*2 B = 503 MB at C2) exceed the 126 MB L2."""
    from optical_flow_b200 import conv_tc, _lib as L
    x = torch.randn((N, H, W, Cin), device=dev).bfloat16()
    w = torch.randn((3, 3, Cin, Cout), device=dev) * 0.05
    b = torch.randn((Cout,), device=dev)
    n_mma = (Cout + 15) // 16 * 16
    wp = conv_tc.pack_weights(w, [(0, Cin, 0)], n_mma, (Cin + 63) // 64 * 64, transpose=False)
    out = torch.empty((N, H, W, Cout), dtype=torch.bfloat16, device=dev)

    def f():
        conv_tc.tc_run([conv_tc.Source(x)], wp, n_mma, Cout, N, H, W, H, W, 3, 3, 1, 1, 1, +1, out, bias=b,
                       act=L.OF_ACT_LEAKY)

    for _ in range(3):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        f()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    fl = 2.0 * N * H * W * Cin * Cout * 9
    return {"us": us, "flops": fl, "tflops": fl / us / 1e6, "iters": iters}


TC_CONV_KERNELS = ("conv_tc3_kernel", "conv_tc_kernel", "conv_tc_wgrad_kernel")
TC_EXCLUDED_KERNELS = ("conv_tc_kernel<stem>", "conv_tc_wgrad_kernel<stem>", "conv_tc3_kernel<head>")


def step_timeline(tr, batches, nsteps=3, delay_ms=25.0):
    """Per-kernel device time of one train step, measured LIVE with one CUDA event after every launch of the library
    (of_prof_enable / of_prof_collect) over `nsteps` steps: {kernel: (us per step, launches per step)}."""
    from optical_flow_b200 import _lib as L

    import torch
    lib = L.load()
    delay_cycles = int(delay_ms * 1e-3 * 1.9e9)

    def run():
        for i in range(nsteps):
            # eager launches through Python are slower than the device at this step size, and an event interval would
            # then include the idle gap before its kernel: queue a GPU-side delay first, so every launch of the step
            # is already waiting in the stream when its predecessor ends (interval = the kernel's own duration)
            torch.cuda._sleep(delay_cycles)
            lib.of_prof_marker(b"_delay")
            tr.train_step(batches[i % len(batches)], i)

    L.timeline(run)                      # first pass creates the events (cudaEventCreate is slow); not used
    recs = [r for r in L.timeline(run) if r[0] != "_delay"]
    agg = {}
    for name, ms, work in recs:
        a = agg.setdefault(name, [0.0, 0, 0.0])
        a[0] += ms * 1e3 / nsteps
        a[1] += 1
        a[2] += work / nsteps
    return {k: {"us": v[0], "launches": v[1] / nsteps, "flops": v[2]} for k, v in agg.items()}, recs


def bench_train(args, dev, world, rank):
    import torch
    import torch.distributed as dist
    from optical_flow_b200 import _lib as L
    from optical_flow_b200.train import Trainer, synthetic_batch, HostFramePrefetcher
    B, H, W = per_gpu_batch(args, world), args.height, args.width
    peaks = measured_peaks()
    comm = None
    if world > 1:
        from optical_flow_b200 import parallel
        comm = parallel.CAbiComm.from_torch_dist(dev)          # the library's own NCCL communicator
    tr = Trainer(H, W, None, conv_mode=args.conv_mode, comm=comm, fuse_residual_epilogues=not args.no_fuse_residual)
    nsets = 2
    batches = [synthetic_batch(B, H, W, 1234 + rank * 100 + k, dev) for k in range(nsets)]
    launch_mode = "eager (one ctypes call per kernel)"
    if args.graph == "on":
        try:
            tr.enable_graph(batches[0])
            launch_mode = "one CUDA graph launch per step (+ a device-to-device copy of the batch into the graph's input)"
        except Exception as e:          # noqa: BLE001 - say so and time the eager path
            tr._graph = None
            sys.stderr.write("CUDA graph capture of the train step failed (%s); eager launches\n" % e)

    def run(n):
        for i in range(n):
            loss, _ = tr.train_step(batches[i % nsets], i)

    clocks = ClockSampler(dev.index or 0)
    clocks.start()
    run(max(args.warmup, 3))
    launches_before = L.load().of_launch_count()
    clocks.mark()
    ms = _timed(torch, dist, world, run, args.steps)
    launches = L.load().of_launch_count() - launches_before
    if tr._graph is not None:           # replayed launches are not counted by the library: kernels per captured step
        launches = args.steps * getattr(tr, "graph_kernels", 0)
    per = ms / args.steps
    value = world * B / (per * 1e-3)

    # distribution over >= 100 individually timed steps (SURVEY 8d protocol): median, p10, p90
    dist_stats = None
    if not args.no_extras:
        n_d = int(max(20, min(100, 4000.0 / max(per, 1e-3))))
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_d + 1)]
        torch.cuda.synchronize()
        evs[0].record()
        for i in range(n_d):
            tr.train_step(batches[i % nsets], i)
            evs[i + 1].record()
        torch.cuda.synchronize()
        t = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(n_d))
        dist_stats = {"steps": n_d, "median_ms": t[n_d // 2], "p10_ms": t[n_d // 10], "p90_ms": t[(n_d * 9) // 10],
                      "min_ms": t[0], "max_ms": t[-1]}

    # end to end through the public API from HOST frames (data_reader.py:35-64 -> train.py:47-61): pinned uint8 pairs
    # -> one H2D copy per batch (copy stream, under the previous step) -> of_preprocess_batch -> train_step, and a
    # device -> host read of the loss, every step
    means = torch.tensor([123.0, 117.0, 104.0] * 2, device=dev) / 255.0
    h_frames = []
    for b in batches:
        u8 = ((b + means) * 255.0).round().clamp(0, 255).to(torch.uint8)                   # (B,H,W,6)
        fr = u8.view(B, H, W, 2, 3).permute(0, 3, 1, 2, 4).contiguous()                    # (B,2,H,W,3)
        h_frames.append(fr.cpu().pin_memory())
    # every step's loss is copied to pinned host memory and read by the host; the read of step i happens after step
    # i + 1 has been enqueued (the way a training loop logs its loss), so the GPU does not idle while the host waits
    h_loss = torch.empty(2).pin_memory()
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    losses_seen = []

    def run_e2e(n):
        feed = HostFramePrefetcher((h_frames[i % nsets] for i in range(n)), dev, H, W, out=tr.graph_input())
        pending = None
        for i, (d_in, release) in enumerate(feed):
            loss, _ = tr.train_step(d_in, i)
            release()
            h_loss[i % 2: i % 2 + 1].copy_(loss.reshape(1), non_blocking=True)
            loss_ev[i % 2].record()
            if pending is not None:
                loss_ev[pending].synchronize()
                losses_seen.append(float(h_loss[pending]))
            pending = i % 2
        if pending is not None:
            loss_ev[pending].synchronize()
            losses_seen.append(float(h_loss[pending]))

    run_e2e(2)
    del losses_seen[:]
    ms_e = _timed(torch, dist, world, run_e2e, args.steps)
    assert len(losses_seen) == args.steps and all(math.isfinite(v) for v in losses_seen), "e2e: a step's loss was not read"

    ck = clocks.stop()
    e2e = {"value": world * B / (ms_e / args.steps * 1e-3), "unit": "image-pairs/s",
           "h2d_bytes_per_step": h_frames[0].numel(), "d2h_bytes_per_step": 4,
           "input": "pinned uint8 frame pairs (B,2,H,W,3) -> of_preprocess_batch -> train_step; every step's loss read "
                    "on the host one step behind (asynchronously)"}
    flops = conv_flops(B, H, W)
    tc_flops = conv_flops(B, H, W, tc_only=True)
    step_tf = (tc_flops if args.conv_mode == "bf16" else flops) / (per * 1e-3) / 1e12
    line_extra = {}
    if args.conv_mode == "bf16":
        roof = {"bound": "tensor", "unit": "TFLOP/s", "peak": peaks["tf_sust"],
                "peak_source": peaks["src"] + " (sustained bf16: kernels timed inside a long step)",
                "algorithmic_flops": tc_flops,
                "step": {"tensor_path_flops": tc_flops, "achieved": step_tf, "peak": peaks["tf_sust"],
                         "frac": step_tf / peaks["tf_sust"],
                         "note": "all tensor-core conv FLOPs of one step / whole step time (includes every non-conv kernel)"}}
        if not args.no_extras:
            # FLOP-weighted in-step figure: all tensor-core conv launches of the step, timed live by the library's
            # per-launch CUDA-event timeline over 3 steps
            graph_saved, tr._graph = tr._graph, None     # the per-launch events need eager launches
            agg, recs = step_timeline(tr, batches)
            tr._graph = graph_saved
            if args.dump_timeline and rank == 0:
                with open(args.dump_timeline, "w") as f:
                    json.dump([{"kernel": n, "us": ms * 1e3, "flops": w} for n, ms, w in recs], f)
            tc_us = sum(v["us"] for k, v in agg.items() if k in TC_CONV_KERNELS)
            tot_us = sum(v["us"] for v in agg.values())
            ach = tc_flops / (tc_us * 1e-6) / 1e12
            # the tensor-pipe figure is over the convs with Cin >= 64 and Cout >= 32; enc.conv1 (Cin = 3) and
            # the Cout = 2 heads run on the same kernels but are HBM-bound - their launches carry their own timeline
            # names, are left out of BOTH the FLOPs and the time above, and are listed here with the all-launch figure
            excl_us = sum(v["us"] for k, v in agg.items() if k in TC_EXCLUDED_KERNELS)
            excl_flops = sum(2.0 * n * ho * wo * cin * cout * k * k * (2 if name == "conv1" else 1)
                             for name, n, ho, wo, cin, k, cout, s_ in conv_layer_table(B, H, W) if cin < 32 or cout < 32)
            ach_all = (tc_flops + excl_flops) / ((tc_us + excl_us) * 1e-6) / 1e12
            roof.update({"achieved": ach, "frac": ach / peaks["tf_sust"],
                         "kernel": "conv_tc3_kernel + conv_tc_kernel + conv_tc_wgrad_kernel (tcgen05 implicit GEMM): all %d "
                                   "tensor-core conv launches (Cin >= 64, Cout >= 32) of one step, FLOP-weighted; %.0f us of a %.0f us step "
                                   "(CUDA events after every launch, 3 steps, launches queued behind a GPU-side delay)"
                                   % (round(sum(v["launches"] for k, v in agg.items() if k in TC_CONV_KERNELS)), tc_us, tot_us),
                         "conv_share_of_step": tc_us / tot_us,
                         "excluded_layers": {"what": "enc.conv1 fprop + wgrad (Cin = 3, 7x7/2) and the four Cout = 2 flow-head "
                                                     "fprops: same tcgen05 kernels, HBM-bound (SURVEY 8d lists them apart)",
                                             "us": excl_us, "flops": excl_flops,
                                             "kernels": {k: round(v["us"], 1) for k, v in agg.items() if k in TC_EXCLUDED_KERNELS}},
                         "all_tc_launches": {"achieved": ach_all, "frac": ach_all / peaks["tf_sust"],
                                             "us": tc_us + excl_us, "flops": tc_flops + excl_flops}})
            top = sorted(agg.items(), key=lambda kv: -kv[1]["us"])[:12]
            line_extra["kernels"] = {k: {"us": round(v["us"], 1), "launches": round(v["launches"], 1),
                                         "share": round(v["us"] / tot_us, 4)} for k, v in top}
            # best single layer, timed alone (burst peak): decoder level-3 128 -> 128
            dom = time_dominant_conv(torch, dev, B, H // 2, W // 2, 128, 128, iters=max(args.steps, 10))
            traffic = None                               # DRAM bytes per launch of that kernel, from the committed ncu capture
            tpath = os.path.join(ROOT, "profiles", "r02_roofline_traffic.json")
            if (B, H, W) == (8, 384, 1280) and os.path.exists(tpath):
                traffic = json.load(open(tpath))["conv_tc3_fprop_N8_192x640_128_128"]["traffic_bytes"]
            roof["traffic"] = traffic
            roof["best_layer"] = {"kernel": "conv_tc3_kernel, decoder level-3 conv 128->128 3x3 fprop: M=%d K=1152 N=128, timed alone"
                                            % (B * (H // 2) * (W // 2)),
                                  "us": dom["us"], "achieved": dom["tflops"], "peak": peaks["tf_burst"],
                                  "frac": dom["tflops"] / peaks["tf_burst"], "traffic": traffic,
                                  "peak_source": peaks["src"] + " (burst bf16: kernel timed alone)"}
            # BASELINE metric part 2 ("warp+loss HBM GB/s", configs[1]) in the same run
            roof["loss_fused"] = loss_kernel_leg(torch, dev, B, H, W, rank, peaks)
        else:
            roof.update({"achieved": step_tf, "frac": step_tf / peaks["tf_sust"], "traffic": None,
                         "kernel": "whole step (--no-extras: no per-kernel timeline)"})
    else:
        roof = {"bound": "tensor", "achieved": step_tf, "peak": peaks["tf_sust"], "unit": "TFLOP/s",
                "frac": step_tf / peaks["tf_sust"], "traffic": None,
                "kernel": "gemm_simt_kernel (fp32 SIMT parity mode, whole step)", "algorithmic_flops": flops,
                "peak_source": peaks["src"]}
    line = {"metric": train_metric(H, W), "value": value, "unit": "image-pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": per,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": {"bf16": "bf16", "bf16x3": "bf16x3 (fp32 storage; three bf16 tensor-core passes per conv)"}.get(args.conv_mode, "f32"),
            "data": "synthetic",
            "config": dict(train_config(B, H, W, world, args.conv_mode, args.scaling), launch=launch_mode),
            "roofline": roof, "e2e": e2e, "gpu_launches": launches if launches else None, "clocks": ck}
    if dist_stats is not None:
        line["distribution"] = dist_stats
    line.update(line_extra)
    return line


if __name__ == "__main__":
    main()
