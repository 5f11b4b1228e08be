#!/usr/bin/env python
"""Sweep of the onesweep-pass tuning configs (HMJ_SORT_CFG) on one GPU.
For every config: sort n random (u64,u32) pairs through hmj_sort_pairs, check against
torch.sort (stable) once, then time `reps` sorts with CUDA events.  Prints one JSON line per
config: ms per sort, ms per pass, GB/s per pass (24 B per pair per pass).
usage: python scripts/sort_sweep.py [--n 67108864] [--cfgs 0,1,2] [--reps 5] [--env K=V ...]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=64 * 2 ** 20)
    ap.add_argument("--cfgs", default="0,5")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--var", default="HMJ_SORT_CFG")
    ap.add_argument("--dist", default="uniform", choices=["uniform", "seq", "dup16", "zipf"])
    args = ap.parse_args()
    import torch
    import hashmergejoin_b200 as hmj

    dev = torch.device("cuda", 0)
    ctx = hmj.Context(0)
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    n = args.n
    if args.dist == "uniform":
        keys = torch.randint(-2 ** 63, 2 ** 63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
    elif args.dist == "seq":
        keys = torch.randperm(n, device=dev, generator=g).to(torch.int64)
    elif args.dist == "dup16":
        keys = torch.randint(-2 ** 63, 2 ** 63 - 1, (n // 16,), dtype=torch.int64, device=dev, generator=g)
        keys = keys.repeat(16)[torch.randperm(n, device=dev, generator=g)]
    else:
        u = torch.rand(n, device=dev, generator=g, dtype=torch.float64)
        ranks = torch.exp(u * torch.log(torch.tensor(float(n), dtype=torch.float64))).to(torch.int64)
        keys = ranks * -7046029254386353131  # wraps: a bijective mix of the rank
    # expected order: unsigned 64-bit ascending, stable
    ukey = keys ^ (-2 ** 63)
    exp_sorted, exp_idx = torch.sort(ukey, stable=True)
    exp_sorted = exp_sorted ^ (-2 ** 63)
    out_h = torch.empty_like(keys)
    out_r = torch.empty(n, dtype=torch.int32, device=dev)
    for cfg in args.cfgs.split(","):
        os.environ[args.var] = cfg
        ctx.sort_pairs(keys, None, out_hash=out_h, out_rowid=out_r)
        torch.cuda.synchronize()
        ok = bool(torch.equal(out_h, exp_sorted)) and bool(torch.equal(out_r.to(torch.int64) & 0xFFFFFFFF, exp_idx))
        ctx.set_profiling(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            ctx.sort_pairs(keys, None, out_hash=out_h, out_rowid=out_r)
        e1.record()
        torch.cuda.synchronize()
        tm = ctx.timings()
        ctx.set_profiling(False)
        ms = e0.elapsed_time(e1) / args.reps
        pl = max(tm["pass_launches"], 1)
        pms = tm["pass_ms_sum"] / pl
        print(json.dumps({"var": args.var, "cfg": cfg, "dist": args.dist, "n": n, "ok": ok,
                          "ms_per_sort": round(ms, 4), "passes_per_sort": pl / args.reps,
                          "ms_per_pass": round(pms, 4), "gbps_per_pass": round(n * 24 / pms / 1e6, 1) if pms > 0 else None,
                          "mkeys_per_s": round(n / ms / 1e3, 1), "finishes": tm["sort_finishes"],
                          "fallbacks": tm["sort_fallbacks"], "uniform": tm["sort_uniform"],
                          "kernel_ms_per_sort": [round(x / args.reps, 3) for x in tm["kernel_ms"][:5]]}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
