#!/usr/bin/env python
"""Multi-GPU parity check of DistJoin (fused NVLink scatter-exchange and the NCCL path).
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 \
        --master-port 29533 scripts/dist_check.py [rows_total]
Every rank owns a block of both relations; the union of the ranks' (build_row, probe_row) pairs
must equal the join computed with numpy on rank 0 (unique keys, plus a few duplicates)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen
from hashmergejoin_b200.dist import DistJoin


def expected_pairs(rk, sk):
    order_s = np.argsort(sk, kind="stable")
    ss = sk[order_s]
    lo = np.searchsorted(ss, rk, side="left")
    cnt = np.searchsorted(ss, rk, side="right") - lo
    r_idx = np.repeat(np.arange(rk.size), cnt)
    # position of every pair inside its run: running index minus the run's first index
    first = np.repeat(np.cumsum(cnt) - cnt, cnt)
    offs = np.repeat(lo, cnt) + (np.arange(r_idx.size) - first)
    return r_idx.astype(np.uint32), order_s[offs].astype(np.uint32)


def canon(a, b):
    o = np.lexsort((b, a))
    return a[o], b[o]


def main():
    rows_total = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    ctx = hmj.Context(torch.cuda.current_device())
    keys = datagen.unique_u64(rows_total, seed=17)
    rk_all, sk_all = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2).copy()
    rk_all[:300] = rk_all[300:600]  # duplicates on the build side (1:N)
    n = rows_total // world
    dev = lambda a: torch.from_numpy(a.view(np.int64)).cuda()
    rk, sk = dev(rk_all[rank * n:(rank + 1) * n].copy()), dev(sk_all[rank * n:(rank + 1) * n].copy())
    for fused in ("1", "0"):
        os.environ["HMJ_DIST_FUSED"] = fused
        if ctx.lib.hmj_dist_teardown(ctx.h) != 0:
            raise RuntimeError("teardown")
        joiner = DistJoin(ctx, rank, world, rows_total)
        m = joiner.join_u64(rk, sk)
        out_r, out_s = ctx.join_result_to_host(m) if (joiner.fused or world > 1) and m else (np.empty(0, np.uint32),) * 2
        if not joiner.fused:
            # the NCCL path returns tensors from merge_join / join_pairs into the context buffers too
            out_r, out_s = ctx.join_result_to_host(m)
        gathered = [None] * world
        dist.all_gather_object(gathered, (out_r, out_s))
        if rank == 0:
            a = np.concatenate([g[0] for g in gathered])
            b = np.concatenate([g[1] for g in gathered])
            e_r, e_s = expected_pairs(rk_all, sk_all)
            ok = all(np.array_equal(x, y) for x, y in zip(canon(a, b), canon(e_r, e_s)))
            print(f"world {world} rows {rows_total} fused_requested={fused} fused_used={joiner.fused} "
                  f"fused_joins={joiner.fused_joins} pairs={a.size} expected={e_r.size} "
                  f"{'OK' if ok else 'MISMATCH'}", flush=True)
            assert ok
    # string keys (C5's shape): build side every key twice, probe side Zipf(1.0); keys travel with
    # their rows over NCCL, the local join verifies them
    n_keys = min(max(rows_total // 16, 1024), 1 << 18)  # host-generated: keep this part in seconds
    off, chars, _, _ = datagen.strgen(n_keys, seed=31, shuffle_seed=3)
    rng = np.random.default_rng(17)
    r_rows = rng.permutation(np.repeat(np.arange(n_keys), 2))
    s_rows = datagen.zipf_indices(2 * n_keys, n_keys, 1.0, seed=5)
    n = 2 * n_keys // world
    devb = lambda a: torch.from_numpy(a.view({8: np.int64, 1: np.uint8}[a.dtype.itemsize])).cuda()
    blk = [datagen.take_rows(off, chars, rows[rank * n:(rank + 1) * n]) for rows in (r_rows, s_rows)]
    joiner = DistJoin(ctx, rank, world)
    m, (g_r, g_s) = joiner.join_strings(devb(blk[0][0]), devb(blk[0][1]), devb(blk[1][0]), devb(blk[1][1]),
                                        want_pairs=True)
    gathered = [None] * world
    dist.all_gather_object(gathered, (g_r.cpu().numpy().view(np.uint32), g_s.cpu().numpy().view(np.uint32),
                                      joiner.local_rows, joiner.exchange_bytes))
    if rank == 0:
        a = np.concatenate([g[0] for g in gathered])
        b = np.concatenate([g[1] for g in gathered])
        e_r, e_s = expected_pairs(r_rows.astype(np.uint64), s_rows.astype(np.uint64))  # equal dictionary row = equal key
        ok = all(np.array_equal(x, y) for x, y in zip(canon(a, b), canon(e_r, e_s)))
        print(f"world {world} strings: {2 * n_keys} x {2 * n_keys} rows, pairs={a.size} expected={e_r.size} "
              f"rows per rank {[g[2] for g in gathered]} bytes sent {[g[3] for g in gathered]} "
              f"{'OK' if ok else 'MISMATCH'}", flush=True)
        assert ok
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
