"""Per-kernel breakdown of one train step in a given conv_mode (CUDA event after every library launch)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from optical_flow_b200.train import Trainer, synthetic_batch

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
B, H, W = 8, 384, 1280
dev = torch.device("cuda", 0)
tr = Trainer(H, W, None, conv_mode=mode)            # eager launches (no enable_graph): the per-launch events need them
batches = [synthetic_batch(B, H, W, k, dev) for k in range(2)]
for i in range(2): tr.train_step(batches[i % 2], i)
torch.cuda.synchronize()
agg, _ = bench.step_timeline(tr, batches, nsteps=2, delay_ms=100.0)
tot = sum(v["us"] for v in agg.values())
print("mode %s: %.1f us per step over %d launches" % (mode, tot, sum(v["launches"] for v in agg.values())))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1]["us"])[:25]:
    print("%-44s n=%5.1f %9.1f us %5.1f%%" % (k[:44], v["launches"], v["us"], 100 * v["us"] / tot))
