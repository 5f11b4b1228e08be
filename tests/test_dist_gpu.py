"""Multi-GPU parity of DistJoin on real GPUs (needs >= 2 of them; `-m gpu`): the union of the
ranks' (build_row, probe_row) pairs equals the join computed with numpy, for the fused NVLink path
(hmj_dist_join_u64: level 1 of the bucket scatter written into the owners' memory, relations
pipelined, flag-synchronised), the NCCL path (partition pass + all_to_all_single) and the string
join.  Ranks are spawned with torch.multiprocessing, one per GPU; every failure travels back to
the pytest process through a queue."""
import os
import socket
import sys
import traceback

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def expected_pairs(rk, sk):
    order_s = np.argsort(sk, kind="stable")
    ss = sk[order_s]
    lo = np.searchsorted(ss, rk, side="left")
    cnt = np.searchsorted(ss, rk, side="right") - lo
    r_idx = np.repeat(np.arange(rk.size), cnt)
    first = np.repeat(np.cumsum(cnt) - cnt, cnt)
    offs = np.repeat(lo, cnt) + (np.arange(r_idx.size) - first)
    return r_idx.astype(np.uint32), order_s[offs].astype(np.uint32)


def canon(a, b):
    """the pair list in canonical order, as one sorted array of (build_row << 32 | probe_row)"""
    w = np.sort((a.astype(np.uint64) << np.uint64(32)) | b.astype(np.uint64))
    return w, w


def _rank_main(rank, world, port, rows_total, q):
    try:
        sys.path.insert(0, ROOT)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch.distributed as dist
        import hashmergejoin_b200 as hmj
        from hashmergejoin_b200 import datagen
        from hashmergejoin_b200.dist import DistJoin
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
        ctx = hmj.Context(rank)
        report = []
        keys = datagen.unique_u64(rows_total, seed=17)
        rk_all, sk_all = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2).copy()
        rk_all[:300] = rk_all[300:600]  # duplicates on the build side (1:N)
        n = rows_total // world
        dev = lambda a: torch.from_numpy(a.view(np.int64)).cuda()
        rk, sk = dev(rk_all[rank * n:(rank + 1) * n].copy()), dev(sk_all[rank * n:(rank + 1) * n].copy())
        e_r, e_s = canon(*expected_pairs(rk_all, sk_all)) if rank == 0 else (None, None)

        def gather_and_check(tag, out_r, out_s, extra=""):
            gathered = [None] * world
            dist.all_gather_object(gathered, (out_r, out_s))
            if rank == 0:
                a, b = canon(np.concatenate([g[0] for g in gathered]), np.concatenate([g[1] for g in gathered]))
                ok = np.array_equal(a, e_r) and np.array_equal(b, e_s)
                report.append(f"{tag}: world {world} rows {rows_total} pairs {a.size} expected {e_r.size} "
                              f"{extra}{'OK' if ok else 'MISMATCH'}")
                assert ok, report[-1]

        for fused in ("1", "0"):
            os.environ["HMJ_DIST_FUSED"] = fused
            joiner = DistJoin(ctx, rank, world, rows_total)
            assert joiner.fused == (fused == "1"), getattr(joiner, "fused_error", "")
            for rep in range(3):   # repeated joins reuse the regions: the DONE/READY flags must order them
                m = joiner.join_u64(rk, sk, measure_exchange=True)
                out_r, out_s = ctx.join_result_to_host(m)
                gather_and_check("fused" if fused == "1" else "nccl", out_r, out_s,
                                 f"join {rep} fused_joins {joiner.fused_joins} fallbacks {joiner.fallback_joins} "
                                 f"sent {joiner.exchange_bytes} B received rows {joiner.local_rows} ")
            if fused == "1":
                assert joiner.fused_joins == 3 and joiner.fallback_joins == 0
                assert sum(joiner.local_rows) > 0 and joiner.exchange_measured
                # a skewed relation overflows a region: every rank falls back together, same pairs
                skew = rk.clone()
                skew[: n // 2] = skew[0]
                m = joiner.join_u64(skew, sk)
                assert joiner.fallback_joins == 1
                tot = torch.tensor([m], device="cuda")
                dist.all_reduce(tot)
                if rank == 0:
                    report.append(f"skewed relation: fell back to the NCCL path, {int(tot.item())} pairs")
                # ... and the regions still serve the next uniform join
                m = joiner.join_u64(rk, sk)
                out_r, out_s = ctx.join_result_to_host(m)
                gather_and_check("fused after a fallback", out_r, out_s)
            joiner.close()

        # string keys (C5's shape): build side every key twice, probe side Zipf(1.0); keys travel
        # with their rows over NCCL, the local join verifies them
        n_keys = 1 << 16
        gen = datagen.DeviceStrgen(ctx, n_keys, seed=31)
        per = 2 * n_keys // world
        r_ids = gen.shuffled_ids(per, seed=7, copies=2, row_base=rank * per)
        cdf = torch.from_numpy(datagen.zipf_cdf(n_keys, 1.0)).cuda()
        s_ids = gen.zipf_ids(cdf, per, seed=5, key_seed=9, row_base=rank * per)
        (roff, rchars), (soff, schars) = gen.keys(r_ids), gen.keys(s_ids)
        x_r = x_s = None
        os.environ["HMJ_DIST_FUSED"] = "1"
        # twice through the fused route (keys written into the owners' buffers by the gather kernel;
        # the second join reuses the buffers), once through the NCCL route, once with buffers too
        # small for the data (every rank falls back to NCCL together)
        for tag, hint, joins in (("fused", 2 * n_keys, 2), ("nccl", 0, 1), ("fused, buffers too small", 1024, 1)):
            joiner = DistJoin(ctx, rank, world, string_rows_hint=hint)
            for rep in range(joins):
                m, (g_r, g_s) = joiner.join_strings(roff, rchars, soff, schars, want_pairs=True)
            assert joiner.xchg_joins == (joins if tag == "fused" else 0), (tag, joiner.xchg_joins)
            gathered = [None] * world
            dist.all_gather_object(gathered, (g_r.cpu().numpy().view(np.uint32), g_s.cpu().numpy().view(np.uint32),
                                              r_ids.cpu().numpy(), s_ids.cpu().numpy()))
            if rank == 0:
                a, b = canon(np.concatenate([g[0] for g in gathered]), np.concatenate([g[1] for g in gathered]))
                if x_r is None:
                    all_r = np.concatenate([g[2] for g in gathered]).astype(np.uint64)
                    all_s = np.concatenate([g[3] for g in gathered]).astype(np.uint64)
                    x_r, x_s = canon(*expected_pairs(all_r, all_s))   # equal key id = equal key
                ok = np.array_equal(a, x_r) and np.array_equal(b, x_s)
                report.append(f"strings ({tag}): world {world} {2 * n_keys} x {2 * n_keys} rows pairs {a.size} "
                              f"expected {x_r.size} {'OK' if ok else 'MISMATCH'}")
                assert ok, report[-1]
            joiner.close()
        dist.barrier()
        ctx.close()
        dist.destroy_process_group()
        q.put((rank, "ok", report))
    except Exception:
        q.put((rank, "error", traceback.format_exc()))


def run_world(world, rows_total):
    import torch.multiprocessing as mp
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_rank_main, args=(r, world, port, rows_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = []
    try:
        for _ in range(world):
            results.append(q.get(timeout=600))
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()   # the exact processes started here
    errors = [r for r in results if r[1] != "ok"]
    assert not errors, "\n".join(f"rank {r[0]}:\n{r[2]}" for r in errors)
    lines = [ln for r in results if r[0] == 0 for ln in r[2]]
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, f"dist_parity_n{world}.txt"), "a") as f:
        f.write("\n".join(lines) + "\n")
    return lines


def _need(world):
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")


def test_dist_join_2_gpus_one_local_level():
    _need(2)
    run_world(2, 1 << 22)


def test_dist_join_2_gpus_two_local_levels():
    # 16M rows over 2 GPUs: level 1 on the GPU bit, two local levels below (C4's plan shape)
    _need(2)
    run_world(2, 1 << 24)


def test_dist_join_4_gpus():
    _need(4)
    run_world(4, 1 << 24)


def test_dist_join_8_gpus():
    # 16M rows over 8 GPUs: 3 GPU bits, then one 9-bit local level (C4's own plan, 3 + 8 + 7 bits,
    # is verified pair by pair in bench.py)
    _need(8)
    run_world(8, 1 << 24)
