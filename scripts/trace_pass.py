#!/usr/bin/env python
"""Per-tile timeline of the last scatter pass of one sort (instrumented build: `make trace`).
usage: HMJ_LIB_PATH=build/trace/libhmj.so python scripts/trace_pass.py [n] [cfg]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("HMJ_LIB_PATH", os.path.join(ROOT, "build", "trace", "libhmj.so"))
import numpy as np
import torch
import hashmergejoin_b200 as hmj

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256 * 2 ** 20
if len(sys.argv) > 2:
    os.environ["HMJ_SORT_CFG"] = sys.argv[2]
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev)
g.manual_seed(7)
keys = torch.randint(-2 ** 63, 2 ** 63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
ctx = hmj.Context(0)
lib = hmj.load()
tiles = (n + 4095) // 4096
buf = torch.zeros(tiles * 12, dtype=torch.int64, device=dev)
out_h = torch.empty_like(keys)
out_r = torch.empty(n, dtype=torch.int32, device=dev)
ctx.sort_pairs(keys, None, out_hash=out_h, out_rowid=out_r)
torch.cuda.synchronize()
lib.hmj_debug_trace.argtypes = [C.c_void_p, C.c_uint64]
assert lib.hmj_debug_trace(buf.data_ptr(), tiles) == 0
ctx.sort_pairs(keys, None, out_hash=out_h, out_rowid=out_r)
torch.cuda.synchronize()
t = buf.cpu().numpy().reshape(tiles, 12)
t = t[t[:, 0] != 0]
t0 = t[:, 0].min()
names = ["load", "rank", "partial+scan", "lookback", "scatter", "writeout"]
print(f"tiles {len(t)}  pass span {(t[:, 6].max() - t0) / 1e3:.1f} us")
for i, nm in enumerate(names):
    d = (t[:, i + 1] - t[:, i]) / 1e3
    print(f"{nm:14s} mean {d.mean():7.2f} us  p50 {np.median(d):7.2f}  p90 {np.percentile(d, 90):7.2f}  max {d.max():7.2f}")
life = (t[:, 6] - t[:, 0]) / 1e3
print(f"{'lifetime':14s} mean {life.mean():7.2f} us  p50 {np.median(life):7.2f}  p90 {np.percentile(life, 90):7.2f}")
print("look-back round trips: mean %.2f p50 %d p90 %d max %d; not-ready polls mean %.2f p90 %d max %d" % (
    t[:, 7].mean(), np.median(t[:, 7]), np.percentile(t[:, 7], 90), t[:, 7].max(),
    t[:, 8].mean(), np.percentile(t[:, 8], 90), t[:, 8].max()))
# how far ahead of the tile order do starts run?  (start time vs tile id)
order = np.argsort(t[:, 0], kind="stable")
print("tiles started out of id order: %.1f %%" % (100.0 * np.mean(np.diff(order) < 0)))
# concurrency: tiles whose [start,end) covers the midpoint of the pass
mid = (t0 + t[:, 6].max()) // 2
print("tiles in flight at mid-pass:", int(np.sum((t[:, 0] <= mid) & (t[:, 6] > mid))))
sm = t[:, 9]
per_sm = np.bincount(sm.astype(np.int64))
print("tiles per SM: min %d max %d" % (per_sm[per_sm > 0].min(), per_sm.max()))
ctx.close()
