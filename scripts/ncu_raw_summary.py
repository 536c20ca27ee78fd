"""Summarise an `ncu --page raw --csv` export: one line per launch with the metrics that matter."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {n: i for i, n in enumerate(hdr)}
def g(r, n, d=float('nan')):
    i = col.get(n)
    if i is None or r[i] in ('', 'n/a'): return d
    try: return float(r[i].replace(',', ''))
    except ValueError: return d
want = [("dur_us", "gpu__time_duration.sum", 1e-3), ("dramR_MB", "dram__bytes_read.sum", None), ("dramW_MB", "dram__bytes_write.sum", None),
        ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1), ("sm%", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1),
        ("l1%", "l1tex__throughput.avg.pct_of_peak_sustained_active", 1), ("l2%", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
        ("occ%", "sm__warps_active.avg.pct_of_peak_sustained_active", 1), ("regs", "launch__registers_per_thread", 1),
        ("ipc", "sm__inst_executed.avg.per_cycle_active", 1)]
print("%-58s %-14s" % ("kernel", "grid") + " ".join("%9s" % w[0] for w in want))
for r in data:
    name = r[col["Kernel Name"]][:58]
    grid = r[col["Grid Size"]] if "Grid Size" in col else ""
    vals = []
    for label, metric, scale in want:
        v = g(r, metric)
        if scale is None:
            u = units[col[metric]] if metric in col else ""
            v = v / 1e6 if u == "byte" else v * (1e-3 if u == "Kbyte" else 1 if u == "Mbyte" else 1e3 if u == "Gbyte" else 1)
        else:
            v *= scale
            if label == "dur_us":
                u = units[col[metric]]
                v = g(r, metric) * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)
        vals.append(v)
    print("%-58s %-14s" % (name, grid) + " ".join("%9.1f" % v for v in vals))
