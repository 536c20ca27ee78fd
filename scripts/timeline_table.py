#!/usr/bin/env python
"""Per-launch table of one train step from bench.py --dump-timeline: kernel, us, GFLOP, TFLOP/s, sorted by time
(the last complete step of the dump, delimited by adam_kernel launches)."""
import json
import sys

recs = json.load(open(sys.argv[1]))
ends = [i for i, r in enumerate(recs) if r["kernel"] == "adam_kernel"]
step = recs[ends[-2] + 1:ends[-1] + 1] if len(ends) >= 2 else recs
tot = sum(r["us"] for r in step)
print("one step: %d launches, %.1f us" % (len(step), tot))
rows = []
for k, r in enumerate(step):
    tf = r["flops"] / r["us"] / 1e6 if r["flops"] and r["us"] > 0 else 0.0
    rows.append((r["us"], k, r["kernel"], r["flops"] / 1e9, tf))
if "--order" in sys.argv:
    rows.sort(key=lambda t: t[1])
else:
    rows.sort(reverse=True)
for us, k, name, gf, tf in rows[: int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 60]:
    print("%4d %-28s %8.1f us %8.1f GFLOP %7.1f TFLOP/s" % (k, name, us, gf, tf))
