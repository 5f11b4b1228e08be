#!/usr/bin/env python
"""One warm-up sort + one sort of n uniform random (u64, row-id) pairs: the ncu target for the
sort kernels.  usage: python scripts/one_sort.py [n]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hashmergejoin_b200 as hmj

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64 * 2 ** 20
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev)
g.manual_seed(7)
keys = torch.randint(-2 ** 63, 2 ** 63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
ctx = hmj.Context(0)
out_h = torch.empty_like(keys)
out_r = torch.empty(n, dtype=torch.int32, device=dev)
for _ in range(2):
    ctx.sort_pairs(keys, None, out_hash=out_h, out_rowid=out_r)
torch.cuda.synchronize()
ctx.close()
