#!/usr/bin/env python
"""One warm-up + one string join of C5's shape (build keys twice, Zipf(1.0) probe) generated on the
device: the ncu target for hash_strings_kernel, the stable passes and join_tile_kernel.
usage: python scripts/one_c5.py [rows per side]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 16 * 2 ** 20
ctx = hmj.Context(0)
gen = datagen.DeviceStrgen(ctx, rows // 2)
r_ids = gen.shuffled_ids(rows, seed=11, copies=2)
cdf = torch.cumsum(1.0 / torch.arange(1, rows // 2 + 1, dtype=torch.float64, device="cuda"), 0)
cdf /= cdf[-1].clone()
cdf[-1] = 1.0
s_ids = gen.zipf_ids(cdf, rows, seed=5, key_seed=13)
(roff, rchars), (soff, schars) = gen.keys(r_ids), gen.keys(s_ids)
for _ in range(2):
    m = ctx.join_strings(roff, rchars, soff, schars)
torch.cuda.synchronize()
assert m == 2 * rows, m
print("pairs", m, ctx.timings())
ctx.close()
