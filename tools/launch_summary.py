"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
Usage: python tools/launch_summary.py gpurun_out/launches.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
agg = collections.defaultdict(list)
for r in rows[1:]:
    try:
        agg[r[ki]].append(float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0))
    except ValueError:
        pass
tot = sum(sum(v) for v in agg.values())
print(f"{'kernel':72s} {'n':>5s} {'total ms':>10s} {'avg ms':>9s} {'share':>7s}")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k[:72]:72s} {len(v):5d} {sum(v):10.3f} {sum(v) / len(v):9.4f} {100 * sum(v) / tot:6.1f}%")
