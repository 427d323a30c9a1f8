#!/usr/bin/env python
"""Per-role stall summary of an `ncu --page source --csv` export of forward_tc_kernel (tcgen05 build).
usage: ncu_roles.py src.csv [min_samples]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 150
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr)]
S = hdr.index('# Samples'); I = hdr.index('Instructions Executed')
stall = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
print(len(data), 'SASS lines,', sum(int(r[S] or 0) for r in data), 'samples')
hm = [i for i, r in enumerate(data) if 'HMMA.16816' in r[1]]; ut = [i for i, r in enumerate(data) if 'UTCHMMA' in r[1]]
st = [i for i, r in enumerate(data) if 'STTM' in r[1]]
def region(a, b, label):
    tot = 0; agg = collections.Counter()
    for i in range(a, b):
        r = data[i]; tot += int(r[S] or 0)
        for j in stall:
            v = int(r[j] or 0)
            if v: agg[hdr[j][6:]] += v
    print(label, a, b, 'samples', tot, agg.most_common(6), 'instr/img', round(sum(int(r[I] or 0) for r in data[a:b]) / 65536, 1))
if ut and st:
    region(st[0] - 250, hm[0] - 120, 'workers')
    region(hm[0] - 120, ut[0] - 200, 'conv2+fc')
    region(ut[0] - 200, len(data), 'issuers')
for i, r in enumerate(data):
    n = int(r[S] or 0)
    if n >= thr:
        reasons = {hdr[j][6:]: int(r[j] or 0) for j in stall if int(r[j] or 0) > 0}
        print(i, n, r[1].strip()[:80], sorted(reasons.items(), key=lambda kv: -kv[1])[:2])
