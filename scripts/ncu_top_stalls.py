#!/usr/bin/env python
"""Top stalled SASS instructions of one kernel from `ncu --page source --csv` output.
usage: ncu -i X.ncu-rep --page source --csv --kernel-id :::1 > src.csv; ncu_top_stalls.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"]
tot = sum(int(r[ix['# Samples']]) for r in data)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {s: sum(int(r[ix[s]]) for r in data) for s in stalls}
print('total samples', tot)
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
for r in sorted(data, key=lambda r: -int(r[ix['# Samples']]))[:topn]:
    st = {s[6:]: int(r[ix[s]]) for s in stalls if int(r[ix[s]])}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"{int(r[ix['# Samples']]):6d} {data.index(r):5d} {r[ix['Source']].strip()[:64]:64s} {st}")
