#!/usr/bin/env python
"""Three whole-operator joins of the scaled C5 shape (build keys twice, Zipf(1.0) probe, strgen
strings): the ncu target for join_tile_kernel on skewed string input.
usage: python scripts/one_join_zipf.py [rows]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4 * 2 ** 20
off, chars, _, _ = datagen.strgen(n // 2, seed=13, shuffle_seed=4)
r = datagen.take_rows(off, chars, np.random.default_rng(5).permutation(np.repeat(np.arange(n // 2), 2)))
s = datagen.take_rows(off, chars, datagen.zipf_indices(n, n // 2, 1.0, seed=7))
dev = lambda a: torch.from_numpy(a.view({8: np.int64, 1: np.uint8}[a.dtype.itemsize])).cuda()
d = [dev(x) for x in (*r, *s)]
ctx = hmj.Context(0)
for _ in range(3):
    m = ctx.join_strings(*d)
torch.cuda.synchronize()
tm = ctx.timings()
print("pairs", m, {k: round(tm[k], 3) for k in ("hash_ms", "sort_ms", "join_ms")})
assert m == 2 * n, m
ctx.close()
