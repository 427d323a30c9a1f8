"""Summarise an `ncu --page source --csv` export: stall-reason totals, and the SASS instructions that collect the
most warp-stall samples.  Usage: ncu -i X.ncu-rep --page source --csv > src.csv; python tools/ncu_src_summary.py src.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
data = [r for r in rows[2:] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
S = ci["# Samples"]
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[S] or 0) for r in data)
print(len(data), "SASS instructions, total samples", tot)
for j in stall:
    s = sum(int(r[j] or 0) for r in data)
    if s > tot * 0.005:
        print(f"  {hdr[j]:28s} {s:7d} {100 * s / tot:5.1f}%")
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
marks = set(i for i, r in enumerate(data) if "SYNCS" in r[1] or "BAR" in r[1] or "EXIT" in r[1])
top = set(sorted(range(len(data)), key=lambda i: -int(data[i][S] or 0))[:top_n])
cum = 0
for i, r in enumerate(data):
    cum += int(r[S] or 0)
    if i in top or i in marks:
        reasons = {hdr[j]: int(r[j] or 0) for j in stall if int(r[j] or 0) > 0}
        main = sorted(reasons.items(), key=lambda kv: -kv[1])[:2]
        print(f"{i:5d} cum={cum:6d} n={r[S]:>5s} {r[1].strip()[:72]:72s} {main}")
