#!/usr/bin/env python
"""Bench lines of the string configurations of BASELINE.json (C1, C2, a scaled C5) on one B200.
One JSON line per configuration on stdout: device-resident GPU time (CUDA events, median of
`reps`), the unmodified reference on the host cores (1 thread and all threads, on the stated
sample), and the parity check that was made on the way.  bench.py stays the C3/C4 headline.
usage: python scripts/bench_configs.py [--c2-rows 67108864] [--c5-rows 16777216] [--reps 5]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen
from tests import oracle_binding as ob


def dev(a):
    return torch.from_numpy(a.view({8: np.int64, 4: np.int32, 1: np.uint8}[a.dtype.itemsize])).cuda()


def gpu_time(fn, reps):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--c2-rows", type=int, default=64 * 2 ** 20)
    ap.add_argument("--c5-rows", type=int, default=16 * 2 ** 20)
    ap.add_argument("--ref-rows", type=int, default=4 * 2 ** 20)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    ctx = hmj.Context(0)
    ref = ob.load_ref()
    cores = ref.cores()

    # ---- C1: hashjoin_bench 1M x 1M strgen strings, unique keys, two shuffles of one key set
    n = 10 ** 6
    off, chars, _, _ = datagen.strgen(n, seed=11, shuffle_seed=1)
    off2, chars2, _, _ = datagen.strgen(n, seed=11, shuffle_seed=2)
    d = [dev(x) for x in (off, chars, off2, chars2)]
    ms = gpu_time(lambda: ctx.join_strings(*d), args.reps)
    m = ctx.join_strings(*d)
    tm = ctx.timings()
    out_r, out_s = ctx.join_result_to_host(m)
    r_r, r_s, _, t1 = ref.join(ob.Keys(offsets=off, chars=chars), ob.Keys(offsets=off2, chars=chars2), threads=1)
    _, _, _, tp = ref.join(ob.Keys(offsets=off, chars=chars), ob.Keys(offsets=off2, chars=chars2), threads=cores)
    o = np.lexsort((out_s, out_r))
    q = np.lexsort((r_s, r_r))
    same = bool(np.array_equal(out_r[o], r_r[q]) and np.array_equal(out_s[o], r_s[q]))
    print(json.dumps({"config": "C1 hashjoin_bench 1M x 1M strgen strings (mean key %.1f B)" % (chars.size / n),
                      "gpu_ms": ms, "gpu_tuples_per_s": 2 * n / ms * 1e3, "pairs": int(m),
                      "uniform_path": tm["bucket_join"] == 1,
                      "stage_ms": {k: tm[k] for k in ("hash_ms", "sort_ms", "join_ms")},
                      "reference_1thread_s": t1, "reference_%dthreads_s" % cores: tp,
                      "reference_tuples_per_s_1t": 2 * n / t1, "pairs_equal_reference": same}), flush=True)
    del d

    # ---- C2: radix_hash_bench hash + sort of (hash,rowid) pairs, strgen strings and u64 keys
    n2 = args.c2_rows
    t0 = time.time()
    off, chars, _, _ = datagen.strgen(n2, seed=12, shuffle_seed=3)
    gen_s = time.time() - t0
    d_off, d_chars = dev(off), dev(chars)
    oh = torch.empty(n2, dtype=torch.int64, device="cuda")
    orow = torch.empty(n2, dtype=torch.int32, device="cuda")
    ms = gpu_time(lambda: ctx.hash_sort_strings(d_off, d_chars, out_hash=oh, out_rowid=orow), args.reps)
    sorted_ok = bool((oh[1:].view(torch.int64) ^ (-2 ** 63) >= oh[:-1].view(torch.int64) ^ (-2 ** 63)).all())
    nr = min(args.ref_rows, n2)
    soff, schars = datagen.take_rows(off, chars, np.arange(nr))
    rh, _, ts1 = ref.sort(ob.Keys(offsets=soff, chars=schars), threads=1)
    _, _, tsp = ref.sort(ob.Keys(offsets=soff, chars=schars), threads=cores)
    gh, _ = ctx.hash_sort_strings(dev(soff), dev(schars))
    same = bool(np.array_equal(gh.cpu().numpy().view(np.uint64), rh))
    bytes_per_key = chars.size / n2 + 8 + 8 + 20 + 24 + 24
    print(json.dumps({"config": "C2 radix_hash_bench hash+sort of %d strgen strings (mean key %.1f B)" % (n2, chars.size / n2),
                      "gpu_ms": ms, "gpu_keys_per_s": n2 / ms * 1e3, "sorted": sorted_ok,
                      "algorithmic_bytes_per_key": bytes_per_key,
                      "achieved_gbps": n2 * bytes_per_key / ms / 1e6,
                      "reference_sample_rows": nr, "reference_1thread_keys_per_s": nr / ts1,
                      "reference_%dthreads_keys_per_s" % cores: nr / tsp,
                      "sample_hash_column_equals_reference": same, "host_generation_s": gen_s}), flush=True)
    del d_off, d_chars
    keys = datagen.unique_u64_torch(n2, seed=5, device="cuda", perm_mult=0x9E3779B1)
    ms = gpu_time(lambda: ctx.hash_sort_u64(keys, out_hash=oh, out_rowid=orow), args.reps)
    print(json.dumps({"config": "C2 (u64 variant) hash+sort of %d uniform u64 keys" % n2, "gpu_ms": ms,
                      "gpu_keys_per_s": n2 / ms * 1e3, "achieved_gbps": n2 * 68.0 / ms / 1e6,
                      "algorithmic_bytes_per_key": 68.0}), flush=True)
    del keys, oh, orow
    torch.cuda.empty_cache()

    # ---- C5 (scaled to one GPU): many-to-many strgen strings, build keys twice, Zipf(1.0) probe
    n5 = args.c5_rows
    dkeys = n5 // 2
    off, chars, _, _ = datagen.strgen(dkeys, seed=13, shuffle_seed=4)
    rng = np.random.default_rng(5)
    roff, rchars = datagen.take_rows(off, chars, rng.permutation(np.repeat(np.arange(dkeys), 2)))
    soff, schars = datagen.take_rows(off, chars, datagen.zipf_indices(n5, dkeys, 1.0, seed=7))
    d = [dev(x) for x in (roff, rchars, soff, schars)]
    ms = gpu_time(lambda: ctx.join_strings(*d), max(2, args.reps // 2))
    m = ctx.join_strings(*d)
    tm = ctx.timings()
    print(json.dumps({"config": "C5 scaled: %d x %d strgen strings, build keys twice, Zipf(1.0) probe, 1 B200" % (n5, n5),
                      "gpu_ms": ms, "gpu_tuples_per_s": 2 * n5 / ms * 1e3, "pairs": int(m), "expected_pairs": 2 * n5,
                      "uniform_path": tm["bucket_join"] == 1, "sort_passes": tm["sort_passes"],
                      "stage_ms": {k: tm[k] for k in ("hash_ms", "sort_ms", "join_ms")}}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
