#!/usr/bin/env python
"""Sweep HMJ_HOST_CHUNKS for the single-call host-buffer join (hmj_join_u64_host) at the C3 size.

  python scripts/e2e_chunks.py [--rows N] [--chunks 4,8,12,16,24,32] [--calls 4]

Prints one line per chunk count: min / median wall time of `calls` joins (after one warm-up) with
pinned host buffers, and a final JSON line.  A tuning aid, not a bench value."""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=256 * 2 ** 20)
    ap.add_argument("--chunks", default="4,8,12,16,24,32")
    ap.add_argument("--calls", type=int, default=4)
    args = ap.parse_args()
    import numpy as np
    import torch
    import bench
    import hashmergejoin_b200 as hmj
    from hashmergejoin_b200 import datagen

    rows = args.rows
    dev = torch.device("cuda:0")
    ctx = hmj.Context(0, use_torch_stream=True)
    ar = torch.arange(0, rows, dtype=torch.int64, device=dev)
    h = []
    for mult in (bench.MULT_R, bench.MULT_S):
        k = bench.keys_of_rows(torch, datagen, ar, mult, rows, False)
        t = torch.empty(rows, dtype=torch.int64).pin_memory()
        t.copy_(k)
        h.append(t)
        del k
    del ar
    torch.cuda.empty_cache()
    o_r = torch.empty(rows, dtype=torch.int32).pin_memory()
    o_s = torch.empty(rows, dtype=torch.int32).pin_memory()
    a = [x.numpy().view(np.uint64) for x in h] + [x.numpy().view(np.uint32) for x in (o_r, o_s)]
    out = {}
    for c in [int(x) for x in args.chunks.split(",")]:
        os.environ["HMJ_HOST_CHUNKS"] = str(c)
        ctx.join_u64_host(*a)
        torch.cuda.synchronize()
        ts = []
        for _ in range(args.calls):
            t0 = time.perf_counter()
            m = ctx.join_u64_host(*a)
            ts.append((time.perf_counter() - t0) * 1e3)
        assert m == rows
        assert bool((a[0][a[2][:1 << 20]] == a[1][a[3][:1 << 20]]).all())
        tm = ctx.timings()
        out[c] = {"min_ms": min(ts), "median_ms": statistics.median(ts), "bucket_join": bool(tm["bucket_join"])}
        print("chunks %2d: min %.2f ms  median %.2f ms  bucket_join=%s" % (c, min(ts), statistics.median(ts), tm["bucket_join"]),
              flush=True)
    print(json.dumps({"rows": rows, "calls": args.calls, "by_chunks": out}))


if __name__ == "__main__":
    main()
