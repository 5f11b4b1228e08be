#!/usr/bin/env python
"""Width sweep of the bucket-scatter levels on a B200 -- the counterpart of the reference's
find_k_bench.cc:31,116-129, which sorts 10M rows with every partition width 1..15 to pick the
radix width.  Here the sort of n (u64 hash, row id) pairs (hmj_hash_sort_u64: scatter levels +
in-shared-memory finish) is timed for every split of the plan's prefix bits between the two
levels (HMJ_UNIFORM_BITS0), beside the general path (histogram-planned stable passes,
HMJ_SORT_UNIFORM=0).  One JSON line per (n, split).
usage: python scripts/sweep_bits.py [n ...]   (default 10M, 64M, 256M rows)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import json, os, sys, torch
sys.path.insert(0, %r)
import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen
n = int(sys.argv[1])
keys = datagen.splitmix64_torch(torch.arange(n, dtype=torch.int64, device="cuda") + (7 << 40))
ctx = hmj.Context(0)
oh = torch.empty(n, dtype=torch.int64, device="cuda"); ov = torch.empty(n, dtype=torch.int32, device="cuda")
for _ in range(3): ctx.hash_sort_u64(keys, oh, ov)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ctx.set_profiling(True)
e0.record()
reps = 10
for _ in range(reps): ctx.hash_sort_u64(keys, oh, ov)
e1.record(); torch.cuda.synchronize()
tm = ctx.timings()
ok = bool(((oh[1:] ^ (-2**63)) >= (oh[:-1] ^ (-2**63))).all().item())
print(json.dumps({"ms": e0.elapsed_time(e1) / reps, "sorted": ok, "uniform_path": tm["sort_uniform"] > 0,
                  "scatter_launches_per_sort": tm["kernel_launches"][1] / reps, "stable_passes_per_sort": tm["kernel_launches"][0] / reps,
                  "scatter_ms_per_sort": tm["kernel_ms"][1] / reps, "finish_ms_per_sort": tm["kernel_ms"][2] / reps}))
''' % ROOT


def run(n, env):
    e = dict(os.environ)
    e.update(env)
    out = subprocess.run([sys.executable, "-c", CHILD, str(n)], env=e, capture_output=True, text=True)
    if out.returncode != 0:
        return {"error": out.stderr.strip().splitlines()[-1][:200] if out.stderr.strip() else "failed"}
    return json.loads(out.stdout.strip().splitlines()[-1])


def main():
    sizes = [int(x) for x in sys.argv[1:]] or [10_000_000, 64 * 2 ** 20, 256 * 2 ** 20]
    for n in sizes:
        qb = 0
        while ((n + (1 << qb) - 1) >> qb) > 4096:
            qb += 1
        rows = []
        if qb > 8:
            for b0 in range(max(1, qb - 9), min(9, qb - 1) + 1):
                r = run(n, {"HMJ_UNIFORM_BITS0": str(b0)})
                r.update({"n": n, "prefix_bits": qb, "level_bits": [b0, qb - b0],
                          "default": b0 == (qb + 1) // 2})
                rows.append(r)
        else:
            r = run(n, {})
            r.update({"n": n, "prefix_bits": qb, "level_bits": [qb], "default": True})
            rows.append(r)
        r = run(n, {"HMJ_SORT_UNIFORM": "0"})
        r.update({"n": n, "plan": "histogram-planned stable passes (general path)"})
        rows.append(r)
        for r in rows:
            if "ms" in r:
                r["keys_per_s"] = n / (r["ms"] * 1e-3)
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
