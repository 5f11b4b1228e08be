#!/usr/bin/env python
"""Timeline of host-buffer joins issued from several host threads (one context each)."""
import os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 28
lanes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
rk = datagen.unique_u64_torch(rows, seed=1, device="cuda", perm_mult=0x9E3779B1)
sk = datagen.unique_u64_torch(rows, seed=1, device="cuda", perm_mult=0x85EBCA6B)
h_r = torch.empty(rows, dtype=torch.int64).pin_memory(); h_r.copy_(rk)
h_s = torch.empty(rows, dtype=torch.int64).pin_memory(); h_s.copy_(sk)
del rk, sk
a = [h_r.numpy().view(np.uint64), h_s.numpy().view(np.uint64)]
ctxs, outs = [], []
for i in range(lanes):
    ctxs.append(hmj.Context(0, use_torch_stream=False))
    o = [torch.empty(rows, dtype=torch.int32).pin_memory() for _ in range(2)]
    outs.append([x.numpy().view(np.uint32) for x in o])
    ctxs[i].join_u64_host(a[0], a[1], *outs[i])
t0 = time.perf_counter()
ctxs[0].join_u64_host(a[0], a[1], *outs[0])
single = time.perf_counter() - t0
print("single call %.1f ms" % (single * 1e3))
log = []
ctxs[0].set_profiling(len(sys.argv) > 3)
def lane(i):
    time.sleep(i * single / lanes)
    for k in range(4):
        s = time.perf_counter()
        ctxs[i].join_u64_host(a[0], a[1], *outs[i])
        log.append((i, k, (s - t0) * 1e3, (time.perf_counter() - t0) * 1e3))
        if i == 0 and len(sys.argv) > 3:
            tm = ctxs[0].timings()
            print("lane 0 call %d kernel_ms %s launches %s stage %s" % (k, [round(x, 2) for x in tm["kernel_ms"]], tm["kernel_launches"],
                  {q: round(tm[q], 2) for q in ("hash_ms", "sort_ms", "join_ms", "total_ms")}), flush=True)
th = [threading.Thread(target=lane, args=(i,)) for i in range(lanes)]
t0 = time.perf_counter()
[t.start() for t in th]; [t.join() for t in th]
tot = time.perf_counter() - t0
for e in sorted(log, key=lambda e: e[2]):
    print("lane %d call %d: %.1f -> %.1f ms (%.1f)" % (e[0], e[1], e[2], e[3], e[3] - e[2]))
print("%d joins in %.1f ms = %.1f ms per join" % (4 * lanes, tot * 1e3, tot * 1e3 / (4 * lanes)))
