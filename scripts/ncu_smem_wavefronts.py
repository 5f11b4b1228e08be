#!/usr/bin/env python
"""Shared-memory wavefronts per SASS instruction of one kernel, against their conflict-free
('ideal') count, from `ncu --page source --csv`: finds bank-conflict hot spots a stall profile
does not show (the bucket join's staging stores were found this way).
usage: ncu -i X.ncu-rep --page source --csv > src.csv; ncu_smem_wavefronts.py src.csv [N]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
c, ci = "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal"
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
tot = 0.0
for r in data:
    w = float(r[ix[c]] or 0)
    if not w:
        continue
    tot += w
    src = r[ix["Source"]].strip()
    op = src.split()[1] if src.startswith("@") else src.split()[0]
    agg[op][0] += 1
    agg[op][1] += w
    agg[op][2] += float(r[ix[ci]] or 0)
print("shared-memory wavefronts: %.2f M" % (tot / 1e6))
for op, (n, w, i) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-14s instrs %3d  wavefronts %8.2f M  ideal %8.2f M" % (op, n, w / 1e6, i / 1e6))
for r in sorted(data, key=lambda r: -float(r[ix[c]] or 0))[:topn]:
    print("%12.0f ideal %11.0f  %5d %s" % (float(r[ix[c]]), float(r[ix[ci]] or 0), data.index(r), r[ix["Source"]].strip()[:70]))
