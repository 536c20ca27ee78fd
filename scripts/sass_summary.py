#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-native SASS mnemonics in the built library (cuobjdump -sass):
UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG / UBLKCP = TMA, plus the legacy HMMA.
Usage: python scripts/sass_summary.py [lib.so] > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "optical_flow_b200", "liboflow_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCCP|HMMA|SYNCS|UTCBAR)\b")
counts = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        counts.setdefault(cur, collections.Counter())
        continue
    if cur is None:
        continue
    m = pat.search(line)
    if m and re.search(r"/\*[0-9a-f]{4,}\*/", line):
        counts[cur][m.group(1)] += 1
    elif re.search(r"/\*[0-9a-f]{4,}\*/", line):
        counts[cur]["(total)"] += 1
cols = ["UTCHMMA", "LDTM", "UTMALDG", "UBLKCP", "SYNCS", "HMMA"]
print("SASS evidence, %s (cuobjdump -sass; sm_100a)" % os.path.basename(lib))
print("UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA), SYNCS = mbarrier ops; HMMA = legacy mma.sync (none expected)")
print("%-64s %8s %6s %8s %7s %6s %5s %7s" % ("kernel", *cols, "instrs"))
tot = collections.Counter()
for k, c in counts.items():
    if not any(c[x] for x in cols[:4]):
        continue
    print("%-64s %8d %6d %8d %7d %6d %5d %7d" % (k[:64], *[c[x] for x in cols], sum(c.values())))
    for x in cols:
        tot[x] += c[x]
print("%-64s %8d %6d %8d %7d %6d %5d" % ("TOTAL (kernels with tensor / TMA instructions)", *[tot[x] for x in cols]))
print("kernels in the library: %d" % len(counts))
