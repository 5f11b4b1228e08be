#!/usr/bin/env python
"""One warm-up + one whole-operator join of n x n unique u64 keys (1:1): the ncu target for the
join kernels.  usage: python scripts/one_join.py [n]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64 * 2 ** 20
dev = torch.device("cuda", 0)
rk = datagen.unique_u64_torch(n, seed=3, device=dev, perm_mult=0x9E3779B1)
sk = datagen.unique_u64_torch(n, seed=3, device=dev, perm_mult=0x85EBCA6B)
ctx = hmj.Context(0)
for _ in range(2):
    m = ctx.join_u64(rk, sk)
torch.cuda.synchronize()
assert m == n, m
ctx.close()
