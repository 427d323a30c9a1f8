#!/usr/bin/env python
"""Per-source-line stall summary of an ncu report (needs -lineinfo and --import-source on).
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > x.csv; ncu_lines.py x.csv [min_samples]
Prints, per source file, the lines that collect at least min_samples warp-stall samples with their top stall reasons."""
import csv, sys, collections
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 100
rows = list(csv.reader(open(sys.argv[1])))
cur = None; hdr = None; tot = 0
out = collections.defaultdict(list)
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No': hdr = r; S = hdr.index('# Samples'); stall = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]; continue
    if hdr is None or len(r) != len(hdr) or not r[0]: continue
    n = int(r[S] or 0); tot += n
    if n >= thr:
        reasons = sorted(((hdr[j][6:], int(r[j] or 0)) for j in stall if int(r[j] or 0) > 0), key=lambda kv: -kv[1])[:3]
        out[cur].append((int(r[0]), n, r[1].strip()[:100], reasons))
print('total samples', tot)
for f, ls in out.items():
    print('==', f, sum(l[1] for l in ls))
    for l in sorted(ls): print(f'  {l[0]:5d} {l[1]:6d}  {l[2]:100s} {l[3]}')
