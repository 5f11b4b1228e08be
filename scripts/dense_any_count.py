#!/usr/bin/env python
"""Dense ids 0..N-1 with N not a power of two through hmj_join_u64: the bucket plan with the
density the sample measures, beside the general path (HMJ_JOIN_UNIFORM=0) on the same inputs.
  python scripts/dense_any_count.py [N]        (default 200,000,000 rows per side)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import hashmergejoin_b200 as hmj
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000_000
    ctx = hmj.Context(0, use_torch_stream=True)
    rk = torch.randperm(n, device="cuda")
    sk = torch.randperm(n, device="cuda")
    out = {"rows_per_side": n}
    for name, env in (("bucket_plan_with_density", None), ("general_path", "0")):
        if env is None:
            os.environ.pop("HMJ_JOIN_UNIFORM", None)
        else:
            os.environ["HMJ_JOIN_UNIFORM"] = env
        ts = []
        for i in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            m = ctx.join_u64(rk, sk)
            e1.record()
            torch.cuda.synchronize()
            assert m == n
            if i:
                ts.append(e0.elapsed_time(e1))
        tm = ctx.timings()
        out[name] = {"ms": min(ts), "tuples_per_s": 2 * n / (min(ts) * 1e-3), "bucket_join": tm["bucket_join"],
                     "scatter_levels": tm["scatter_levels"], "sort_passes": tm["sort_passes"]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
