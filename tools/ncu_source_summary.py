"""Summarise `ncu --page source --csv` output: instruction mix by executed count and the top
stall sites.  Usage: ncu -i X.ncu-rep --page source --csv | python tools/ncu_source_summary.py"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(sys.stdin))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
ops, samples_by_op = Counter(), Counter()
total_exec = total_samples = 0
for r in body:
    toks = r[col["Source"]].split()
    op = toks[0] if not toks[0].startswith("@") else toks[1]
    op = op.split(".")[0]
    ex = int(r[col["Instructions Executed"]] or 0)
    sm = int(r[col["# Samples"]] or 0)
    ops[op] += ex
    samples_by_op[op] += sm
    total_exec += ex
    total_samples += sm
print(f"warp instructions executed: {total_exec:,}   stall samples: {total_samples:,}")
print("op            executed      share   samples share")
for op, ex in ops.most_common(18):
    print(f"{op:10s} {ex:14,d} {100 * ex / total_exec:8.2f}% {100 * samples_by_op[op] / max(total_samples, 1):8.2f}%")
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("\nstall reasons (all samples):")
tot = Counter()
for r in body:
    for h in stall_cols:
        tot[h] += int(r[col[h]] or 0)
for h, v in tot.most_common(10):
    print(f"  {h:24s} {v:9,d} {100 * v / max(total_samples, 1):6.2f}%")
print("\ntop 25 instructions by samples:")
top = sorted(body, key=lambda r: -int(r[col["# Samples"]] or 0))[:25]
for r in top:
    reasons = sorted(((int(r[col[h]] or 0), h) for h in stall_cols), reverse=True)[:2]
    print(f"  {r[col['Address']][-5:]} {int(r[col['# Samples']]):7d} ex={int(r[col['Instructions Executed']]):11,d} "
          f"{r[col['Source']].strip()[:70]:70s} {reasons[0][1]}={reasons[0][0]} {reasons[1][1]}={reasons[1][0]}")
