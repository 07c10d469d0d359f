"""Per-CUDA-line attribution of an ncu capture.
Usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass | python tools/ncu_line_summary.py [top_n]
Prints, per source file, the lines with the most executed warp instructions and stall samples."""
import csv
import sys
from collections import defaultdict

def _int(v):
    try:
        return int(v)
    except ValueError:
        return 0


top_n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = list(csv.reader(sys.stdin))
fname = "?"
col = None
cur_line, cur_src = None, ""
ex = defaultdict(int)
sm = defaultdict(int)
text = {}
for r in rows:
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        col = {}
        for i, h in enumerate(r):
            col.setdefault(h, i)
        continue
    if col is None or len(r) < len(col):
        continue
    if r[0]:
        cur_line, cur_src = int(r[0]), r[1]
    key = (fname, cur_line)
    text[key] = cur_src.strip()
    if not r[col["Address"]]:
        continue   # the CUDA line's own roll-up row
    ex[key] += _int(r[col["Instructions Executed"]])
    sm[key] += _int(r[col["# Samples"]])
tot_ex, tot_sm = sum(ex.values()), sum(sm.values())
print(f"warp instructions {tot_ex:,}  samples {tot_sm:,}")
per_file = defaultdict(lambda: [0, 0])
for k in ex:
    per_file[k[0]][0] += ex[k]
    per_file[k[0]][1] += sm[k]
for f, (e, s) in sorted(per_file.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:28s} inst {100 * e / tot_ex:6.2f}%  samples {100 * s / max(tot_sm, 1):6.2f}%")
print(f"\ntop {top_n} lines by samples:")
for k in sorted(sm, key=lambda k: -sm[k])[:top_n]:
    print(f"  {k[0]}:{k[1]:<5d} inst {100 * ex[k] / tot_ex:5.2f}% samples {100 * sm[k] / max(tot_sm, 1):5.2f}%  {text[k][:110]}")
