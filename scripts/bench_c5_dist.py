#!/usr/bin/env python
"""C5 (scaled): many-to-many string join over N B200s -- build side every key twice, probe side
Zipf(1.0) draws of the same dictionary; keys travel with their rows in the one exchange step.
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29544 scripts/bench_c5_dist.py [--rows 33554432] [--steps 3]
Rank 0 prints one JSON line: whole-job tuples/s, device-timed (CUDA events between barriers, max
over ranks), per-rank row counts after the exchange (the head key's owner gets more)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen
from hashmergejoin_b200.dist import DistJoin


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=32 * 2 ** 20, help="rows per relation, whole job")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    stdout = os.dup(1)
    os.dup2(2, 1)  # library banners go to stderr; the JSON line alone goes to stdout
    ctx = hmj.Context(torch.cuda.current_device())
    n_keys, n = args.rows // 2, args.rows // world
    off, chars, _, _ = datagen.strgen(n_keys, seed=13, shuffle_seed=4)
    r_rows = np.random.default_rng(5).permutation(np.repeat(np.arange(n_keys), 2))[rank * n:(rank + 1) * n]
    w = 1.0 / np.arange(1, n_keys + 1, dtype=np.float64)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    u = np.random.default_rng(7).random(args.rows)[rank * n:(rank + 1) * n]
    s_rows = np.minimum(np.searchsorted(cdf, u), n_keys - 1)
    dev = lambda a: torch.from_numpy(a.view({8: np.int64, 1: np.uint8}[a.dtype.itemsize])).cuda()
    r = [dev(x) for x in datagen.take_rows(off, chars, r_rows)]
    s = [dev(x) for x in datagen.take_rows(off, chars, s_rows)]
    key_bytes = float(chars.size) / n_keys
    del off, chars, cdf, u, w
    joiner = DistJoin(ctx, rank, world)
    for _ in range(args.warmup):
        m = joiner.join_strings(r[0], r[1], s[0], s[1])
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        m = joiner.join_strings(r[0], r[1], s[0], s[1])
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot = torch.tensor([m], dtype=torch.int64, device="cuda")
    dist.all_reduce(tot)
    info = [None] * world
    dist.all_gather_object(info, {"rows": joiner.local_rows, "pairs": int(m), "sent": int(joiner.exchange_bytes),
                                  "stage_ms": {k: round(ctx.timings()[k], 3) for k in ("hash_ms", "sort_ms", "join_ms")}})
    if rank == 0:
        ms = float(t.item())
        line = {"config": "C5 scaled: %d x %d strgen strings (mean key %.1f B), build keys twice, Zipf(1.0) probe" % (args.rows, args.rows, key_bytes),
                "n_gpus": world, "ms_per_join": ms, "tuples_per_s": 2 * args.rows / ms * 1e3,
                "pairs": int(tot.item()), "expected_pairs": 2 * args.rows, "steps": args.steps, "warmup": args.warmup,
                "per_rank": info, "exchange": "NCCL all_to_all_single: row ids, key lengths, key bytes; hashes recomputed by the receiver",
                "timing": "CUDA events between barriers, max over ranks; includes the exchange and its host-side count planning"}
        os.dup2(stdout, 1)
        print(json.dumps(line), flush=True)
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
