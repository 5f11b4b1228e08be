#!/usr/bin/env python
"""bench.py -- the headline benchmark of the hash-merge join path (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A "step" is one complete join of two synthetic relations: hash both relations' keys, bring equal
hashes together, verify keys, write the (build_row, probe_row) pairs.  The library picks the plan
and the line says which one ran (`config.path`):
  "radix-partition + in-smem bucket join"  both relations are cut on the same top hash bits by
        one or two bucket-scatter levels and joined bucket by bucket in shared memory (uniform
        hashes: c1, c3, c4);
  "onesweep sort + merge-path join"        stable LSD passes + merge join (skewed / duplicated
        hashes: the Zipf side of c5; forced by c3-onesweep);
  "bucket-scatter sort + merge-path join"  both relations sorted by scatter levels + an
        in-shared-memory finish, then merge-joined (c3-sortmerge, or hmj_set_join_order).
Workloads (BASELINE.json configs):
  c3 (default, N=1)  256M x 256M u64 keys, 1:1                         configs[2]
  c4 (default, N>1)  1B x 1B u64 keys over N GPUs, strong scaling      configs[3]
        level 1 of the scatter writes into the owners' memory over NVLink (no NCCL all-to-all on
        this path; NCCL carries one 5-word all-reduce per join)
  c1   1M x 1M strgen strings, unique keys                             configs[0]
  c2   hash + sort of 64M strgen strings (hmj_hash_sort_strings)       configs[1]
  c5   strgen strings, build keys twice, Zipf(1.0) probe; --rows sets rows per side (default
       64M on one GPU; 512M x 512M = configs[4] with --gpus 8)
  c3-dense      c3 with keys 0..N-1 (identity hash: the top 36 hash bits are all zero; the sample
                finds the shared prefix and the bucket plan skips it)
  c3-sortmerge  c3 with both relations sorted and merge-joined (HMJ_JOIN_UNIFORM=0)
  c3-onesweep   ... and the sort done by stable LSD passes (HMJ_SORT_UNIFORM=0): the literal
                hash -> onesweep radix sort -> merge-path join pipeline
  small, mid    4M / 64M u64 rows per side (smoke / profiling sizes)
metric  = joined input tuples/sec = (|R|+|S|) / step time  (c2: keys/sec, stated in the line).
value   = inputs already resident in HBM (CUDA events).
e2e     = host buffers through the C ABI (hmj_join_u64_host / hmj_join_strings_host: H2D of the
          keys, the join, D2H of the pairs), ONE call at a time; `e2e.pipelined` = several calls
          in flight from several host threads.
verified = after the timed loop every result pair is checked on the device: both keys recomputed
          from the row ids are equal, no build row and no probe row appears twice, and (N>1) the
          key lies in the rank's hash range; with the pair total this means every row joined once.
--impl reference times the UNMODIFIED reference (oracle/_ref, HashMergeJoin with all host
threads) on a bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "joined input tuples/sec (build+probe)"
UNIT = "tuples/s"
M = 2 ** 20
WORKLOADS = {
    # name: (kind, rows per side, description)
    "c1": ("strings", 10 ** 6, "hashjoin_bench: 1M-row build x 1M-row probe, strgen random strings, unique build keys"),
    "c2": ("sort", 64 * M, "radix_hash_bench: hash+radix sort of 64M (hash,rowid) pairs, strgen strings, 1 B200"),
    "c3": ("u64", 256 * M, "hash-merge join 256M x 256M 64-bit int keys, 1:1 match rate, 1 B200"),
    "c4": ("u64", 2 ** 30, "hash-merge join 1B x 1B int keys radix-partitioned across N B200 (level-1 scatter over NVLink)"),
    "c5": ("zipf", 64 * M, "skewed many-to-many join, strgen strings, build keys twice, Zipf(1.0) probe, full-key collision check"),
    "c3-dense": ("u64", 256 * M, "hash-merge join 256M x 256M dense int keys 0..N-1 (identity hash: skewed top bits)"),
    "c3-sortmerge": ("u64", 256 * M, "hash-merge join 256M x 256M 64-bit int keys: both relations sorted (scatter levels + in-smem finish), then the merge-path join"),
    "c3-onesweep": ("u64", 256 * M, "hash-merge join 256M x 256M 64-bit int keys: onesweep-style stable LSD passes, then the merge-path join (the north-star's literal pipeline)"),
    "small": ("u64", 4 * M, "hash-merge join 4M x 4M 64-bit int keys (smoke-sized)"),
    "mid": ("u64", 64 * M, "hash-merge join 64M x 64M 64-bit int keys (profiling-sized)"),
}
CPU_SAMPLE_ROWS = 64 * M        # rows per side of the bounded CPU sample (<= 64M fit the box)
CPU_SAMPLE_ROWS_1T = 16 * M     # single-thread rows of the CPU table (time-bounded)
KERNEL_NAMES = ["onesweep_pass_kernel", "bucket_scatter_tma_kernel", "segment_sort_kernel",
                "join_tile_kernel", "radix_histogram_kernel", "hash_strings_kernel", "bucket_join_kernel",
                "bucket_scatter_tma_kernel<DIST> (level 1 over NVLink)"]
PATH_BUCKET = "radix-partition + in-smem bucket join"
PATH_SORT = "onesweep sort + merge-path join"
PATH_SCATTER_SORT = "bucket-scatter sort + merge-path join"


def path_name(tm):
    if tm["bucket_join"]:
        return PATH_BUCKET
    return PATH_SORT if tm["sort_passes"] else PATH_SCATTER_SORT


def log(*a):
    print(*a, file=sys.stderr, flush=True)


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel, alg_bytes):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed
    `ncu --set full` capture (profiles/roofline_traffic.json), scaled from the captured launch's
    pair count to this run's by the ratio of algorithmic bytes; None when no capture exists.
    (A capture of another run: it cross-checks the byte model, it cannot see this run.)"""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        t = t["kernels"][kernel]
        return t["dram_bytes_per_launch"] * alg_bytes / t["algorithmic_bytes_per_launch"]
    except Exception:
        return None


def roofline_from(tm, elapsed_ms):
    """per-kernel-class figures of the timed steps (CUDA events around every launch on the stream
    it runs on; algorithmic bytes accumulated by the library at the launch sites)"""
    peak, peak_src = peaks()
    kernels = {}
    for i, nm in enumerate(KERNEL_NAMES):
        ln, ms_sum, by = tm["kernel_launches"][i], tm["kernel_ms"][i], tm["kernel_bytes"][i]
        if ln:
            kernels[nm] = {"launches": ln, "avg_ms": ms_sum / ln, "alg_bytes_per_launch": by / ln,
                           "gbps": by / ms_sum / 1e6 if ms_sum > 0 else None,
                           "share_of_step": ms_sum / elapsed_ms if elapsed_ms else None}
    hbm = {k: v for k, v in kernels.items() if "NVLink" not in k}
    if not hbm:
        return None, kernels
    top = max(hbm, key=lambda k: hbm[k]["share_of_step"])
    k = hbm[top]
    return {"bound": "hbm", "kernel": top, "achieved": k["gbps"], "peak": peak, "unit": "GB/s",
            "frac": k["gbps"] / peak, "traffic": ncu_traffic(top, k["alg_bytes_per_launch"]),
            "peak_source": peak_src, "algorithmic_bytes_per_launch": k["alg_bytes_per_launch"],
            "avg_launch_ms": k["avg_ms"], "launches_timed": k["launches"],
            "share_of_step": k["share_of_step"], "kernels": kernels}, kernels


# ------------------------------------------------------------------------------------------
# CPU legs (the only places that load oracle/)
def _cpu_keys(kind, rows):
    """host relations of the CPU sample: same generators / shapes as the GPU arm"""
    import numpy as np
    from hashmergejoin_b200 import datagen
    from tests import oracle_binding as ob
    if kind in ("u64",):
        keys = datagen.unique_u64(rows, seed=3)
        return ob.Keys(u64=datagen.permuted(keys, 1)), ob.Keys(u64=datagen.permuted(keys, 2)), rows
    if kind in ("strings", "sort"):
        a = datagen.strgen(rows, seed=1234, shuffle_seed=1)
        b = datagen.strgen(rows, seed=1234, shuffle_seed=2)
        return ob.Keys(offsets=a[0], chars=a[1]), ob.Keys(offsets=b[0], chars=b[1]), rows
    # zipf: every key twice on the build side, Zipf(1.0) probe (true pairs = 2 * rows)
    d = rows // 2
    off, chars, _ = datagen.strgen_unshuffled(d, seed=1234)
    rng = np.random.default_rng(17)
    r = datagen.take_rows(off, chars, rng.permutation(np.repeat(np.arange(d), 2)))
    s = datagen.take_rows(off, chars, datagen.zipf_ranks(datagen.zipf_cdf(d), rows, 5).astype(np.int64))
    return ob.Keys(offsets=r[0], chars=r[1]), ob.Keys(offsets=s[0], chars=s[1]), 2 * rows


def cpu_reference_run(kind, rows, steps, warmup, threads=None, extras=False):
    """The reference's own CPU path for the workload kind on `rows` rows per side: HashMergeJoin
    ctor + full iteration (kind sort: radix_non_inplace_par) of the unmodified headers; falls back
    to the oracle port if oracle/_ref is absent.  extras: the other CPU rows BASELINE.md promises
    (single thread, sort only, unordered_map build/probe), one run each."""
    from tests import oracle_binding as ob
    r, s, true_pairs = _cpu_keys(kind, rows)
    ref = ob.load_ref()
    times, table = [], {}
    if ref is not None:
        cores = threads or ref.cores()
        for i in range(warmup + steps):
            if kind == "sort":
                _, _, t = ref.sort(r, threads=cores)
            else:
                m, _, _, t = ref.join(r, s, threads=cores, want_pairs=False)
                # (the reference's walk under-emits on duplicate runs and at the last element)
                assert m >= 1 and (kind == "zipf" or m >= rows - 1), "reference lost matches"
            if i >= warmup:
                times.append(t)
        kind_name = "reference"
        if extras:
            if kind == "sort":
                n1 = min(rows, CPU_SAMPLE_ROWS_1T // 4)
                r1 = ob.Keys(offsets=r.offsets[: n1 + 1], chars=r.chars[: int(r.offsets[n1])])
                table["radix_non_inplace_par_1_thread"] = {"rows": n1, "keys_per_s": n1 / ref.sort(r1, threads=1)[2]}
            else:
                table["unordered_map_build_probe_1_thread"] = {"rows_per_side": rows, "tuples_per_s": 2 * rows / ref.hash_join_raw(r, s)[1]}
                table["radix_non_inplace_par_%d_threads" % cores] = {"rows": rows, "keys_per_s": rows / ref.sort(r, threads=cores)[2]}
                if kind == "u64":
                    n1 = min(rows, CPU_SAMPLE_ROWS_1T)
                    r1, s1, _ = _cpu_keys(kind, n1) if n1 != rows else (r, s, 0)
                else:
                    n1, r1, s1 = rows, r, s
                if n1 <= CPU_SAMPLE_ROWS_1T:
                    table["HashMergeJoin_1_thread"] = {"rows_per_side": n1, "tuples_per_s": 2 * n1 / ref.join(r1, s1, threads=1, want_pairs=False)[3]}
                    table["radix_non_inplace_par_1_thread"] = {"rows": n1, "keys_per_s": n1 / ref.sort(r1, threads=1)[2]}
    else:
        orc = ob.load_oracle()
        cores = 1
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            if kind == "sort":
                orc.radix_sort(orc.hashes(r), r)
            else:
                orc.full_join(r, s)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        kind_name = "port"
    t = sum(times) / len(times)
    units = rows if kind == "sort" else 2 * rows
    what = {"u64": "unique u64 keys (1:1)", "strings": "strgen strings, unique keys (1:1)",
            "sort": "strgen strings", "zipf": "strgen strings, build keys twice, Zipf(1.0) probe"}[kind]
    op = "radix_non_inplace_par of %d" % rows if kind == "sort" else "HashMergeJoin of %d x %d" % (rows, rows)
    out = {"value": units / t, "unit": "keys/s" if kind == "sort" else UNIT, "cores": cores, "kind": kind_name,
           "seconds_per_step": t,
           "sample": f"{op} {what}, {'sort' if kind == 'sort' else 'ctor + full iteration'}, "
                     f"{cores} thread(s), mean of {len(times)} run(s)"}
    if table:
        out["other_cpu_paths"] = table
    return out


def pick_workload(args):
    return args.workload or ("c3" if args.gpus == 1 else "c4")


def cpu_sample_rows(kind, rows_total):
    if kind == "u64":
        return min(rows_total, CPU_SAMPLE_ROWS)
    if kind == "zipf":
        return min(rows_total, 8 * M)
    return min(rows_total, 16 * M)      # strings: ~100 B/row of host tuples, bounded time


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = pick_workload(args)
    kind, rows_total, desc = WORKLOADS[wl]
    rows_total = args.rows or rows_total
    rows = cpu_sample_rows(kind, rows_total)
    res = cpu_reference_run(kind, rows, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC if kind != "sort" else "hash+sorted keys/sec", "value": res["value"],
        "unit": res["unit"], "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["seconds_per_step"] * 1e3,
        "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": desc, "sample_rows_per_side": rows, "same_size_as_gpu_arm": rows == rows_total,
                   "note": "reference CPU path on a bounded sample of the workload (host RAM/time); "
                           "the CPU path gets slower per tuple as it grows, so the sample flatters it"},
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": res["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print_line(line)


# ------------------------------------------------------------------------------------------
def perm_idx(torch, lo, hi, mult, rows_total, device):
    return (torch.arange(lo, hi, dtype=torch.int64, device=device) * mult + 12345) & (rows_total - 1)


MULT_R, MULT_S, KEY_SEED = 0x9E3779B1, 0x85EBCA6B, 3


def keys_of_rows(torch, datagen, rows, mult, rows_total, dense):
    """the key of global row `rows` (int64 tensor) of the relation generated with `mult`"""
    idx = (rows * mult + 12345) & (rows_total - 1)
    return idx if dense else datagen.splitmix64_torch(idx + (KEY_SEED << 40))


def verify_u64(torch, dist, datagen, ctx, m, rows_total, world, rank, dense, device):
    """Device-side check of the last join's pairs (see module docstring: `verified`)."""
    out_r, out_s = ctx.join_result_tensors(m)
    seen_r = torch.zeros(rows_total, dtype=torch.int32, device=device)
    seen_s = torch.zeros(rows_total, dtype=torch.int32, device=device)
    keys_equal, in_range = True, True
    bits = int(math.log2(world))
    chunk = 1 << 26
    for lo in range(0, m, chunk):
        r = out_r[lo:lo + chunk].to(torch.int64) & 0xFFFFFFFF
        s = out_s[lo:lo + chunk].to(torch.int64) & 0xFFFFFFFF
        kr = keys_of_rows(torch, datagen, r, MULT_R, rows_total, dense)
        ks = keys_of_rows(torch, datagen, s, MULT_S, rows_total, dense)
        keys_equal = keys_equal and bool((kr == ks).all().item())
        if bits:
            top = (kr >> (64 - bits)) & ((1 << bits) - 1)
            in_range = in_range and bool((top == rank).all().item())
        one = torch.ones(r.numel(), dtype=torch.int32, device=device)
        seen_r.index_add_(0, r, one)
        seen_s.index_add_(0, s, one)
    no_dup = int(seen_r.max().item()) <= 1 and int(seen_s.max().item()) <= 1 if m else True
    total = torch.tensor([m, 0 if (keys_equal and in_range and no_dup) else 1], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(total)
    m_tot, bad = int(total[0].item()), int(total[1].item())
    res = {"pairs": m_tot, "all_keys_equal": keys_equal, "no_row_twice": no_dup,
           "keys_in_rank_hash_range": in_range if bits else None,
           "rows_covered": bad == 0 and m_tot == rows_total,
           "how": "keys recomputed on the device from every pair's global row ids; per-row counters"}
    assert bad == 0 and m_tot == rows_total, f"verification failed: {res}"
    return res


def run_gpu(args):
    import torch
    wl = pick_workload(args)
    kind = WORKLOADS[wl][0]
    if kind == "u64":
        return run_gpu_u64(args, wl)
    return run_gpu_strings(args, wl)


def init_dist(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch N>1 with torchrun)"
    torch.cuda.set_device(local_rank)
    return world, rank, local_rank, torch.device("cuda", local_rank)


def timed_loop(torch, dist, ctx, step, args, world, rank, local_rank, device):
    """W warm-ups, then K steps between barriers, CUDA events on the launching stream, max over ranks"""
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(args.warmup):
        m = step()
    sampler = ClockSampler(local_rank)
    barrier()
    ctx.set_profiling(True)
    launches0 = ctx.launch_count()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        m = step()
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    elapsed_ms = e0.elapsed_time(e1)
    launches = ctx.launch_count() - launches0
    tm = ctx.timings()
    ctx.set_profiling(False)
    if world > 1:
        t = torch.tensor([elapsed_ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
        t = torch.tensor([launches], dtype=torch.int64, device=device)
        dist.all_reduce(t)
        launches = int(t.item())
    return m, elapsed_ms, launches, tm, clocks, barrier


def run_gpu_u64(args, wl):
    import numpy as np
    import torch
    import torch.distributed as dist
    import hashmergejoin_b200 as hmj
    from hashmergejoin_b200 import datagen

    world, rank, local_rank, device = init_dist(args)
    ctx = hmj.Context(local_rank, use_torch_stream=True)
    _, rows_total, wl_desc = WORKLOADS[wl]
    rows_total = args.rows or rows_total
    assert rows_total & (rows_total - 1) == 0, "u64 workloads need a power-of-two row count (row permutation)"
    dense = wl == "c3-dense"
    if wl in ("c3-sortmerge", "c3-onesweep"):
        os.environ["HMJ_JOIN_UNIFORM"] = "0"
    if wl == "c3-onesweep":
        os.environ["HMJ_SORT_UNIFORM"] = "0"
    rows = rows_total // world  # rows per side owned by this rank
    # same key set on both sides, two different row orders (hashjoin_bench.cc:112-113): 1:1
    lo = rank * rows
    rkeys = keys_of_rows(torch, datagen, torch.arange(lo, lo + rows, dtype=torch.int64, device=device), MULT_R, rows_total, dense)
    skeys = keys_of_rows(torch, datagen, torch.arange(lo, lo + rows, dtype=torch.int64, device=device), MULT_S, rows_total, dense)
    joiner = None
    if world > 1:
        from hashmergejoin_b200.dist import DistJoin
        joiner = DistJoin(ctx, rank, world, rows_total)
    torch.cuda.synchronize()

    def step():
        if joiner is None:
            return ctx.join_u64(rkeys, skeys)
        return joiner.join_u64(rkeys, skeys)

    m, elapsed_ms, launches, tm, clocks, barrier = timed_loop(torch, dist, ctx, step, args, world, rank, local_rank, device)
    ms_per_step = elapsed_ms / args.steps
    value = 2 * rows_total / (ms_per_step * 1e-3)
    roofline, kernels = roofline_from(tm, elapsed_ms)

    # ---- the pairs of the last timed step, checked on the device at every N
    verified = verify_u64(torch, dist, datagen, ctx, m, rows_total, world, rank, dense, device)
    nvlink = None
    if world > 1:
        m2 = joiner.join_u64(rkeys, skeys, measure_exchange=True)
        x = kernels.get(KERNEL_NAMES[7])
        per_launch = joiner.exchange_bytes / 2 if joiner.exchange_measured else None
        nvlink = {"fused_joins": joiner.fused_joins, "fallback_joins": joiner.fallback_joins,
                  "bytes_sent_to_peers_rank0_per_join": joiner.exchange_bytes,
                  "bytes_measured": joiner.exchange_measured,
                  "rows_received_rank0": list(joiner.local_rows),
                  "level1_launches_timed": x["launches"] if x else 0,
                  "level1_avg_ms": x["avg_ms"] if x else None,
                  "gbps_out_per_gpu": per_launch / x["avg_ms"] / 1e6 if x and per_launch else None,
                  "frac_of_900": per_launch / x["avg_ms"] / 1e6 / 900.0 if x and per_launch else None,
                  "note": "12 B per pair written into other GPUs' memory by the level-1 scatter (cursors read back); "
                          "level 1 of S shares the SMs with R's local levels, so its time is not NVLink alone"}

    # ---- K7 beside the join (hashjoin_bench.cc:29-63 times the hash-table join next to BM_HashMergeJoin)
    k7 = None
    if world == 1 and not args.no_k7:
        try:
            R, S = hmj.SortedRelation(rkeys, None), hmj.SortedRelation(skeys, None)
            ctx.hashjoin_build_probe(R, S)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            reps = 3
            for _ in range(reps):
                hits, _ = ctx.hashjoin_build_probe(R, S)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / reps
            assert hits == rows_total or dense
            k7 = {"kernel": "ht_build_kernel + ht_probe_kernel (open addressing, atomicCAS; aggregate only, no pairs written)",
                  "ms": dt * 1e3, "tuples_per_s": 2 * rows_total / dt, "hits": hits,
                  "join_over_hash_table": (2 * rows_total / dt) and value / (2 * rows_total / dt)}
        except hmj.HmjError as e:
            k7 = {"error": str(e)[:200]}

    # ---- e2e: host buffers through the C ABI (H2D + join + D2H of the pairs)
    e2e = None
    if world == 1 and not args.no_e2e:
        e2e = e2e_u64_single_gpu(args, torch, np, hmj, ctx, local_rank, rows, rkeys, skeys)
    elif world > 1 and not args.no_e2e:
        e2e = e2e_u64_multi_gpu(args, torch, dist, ctx, joiner, local_rank, rows, rows_total, rkeys, skeys, barrier, device)

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            n_cpu = cpu_sample_rows("u64", rows_total)
            cpu = cpu_reference_run("u64", n_cpu, 2 if n_cpu <= 16 * M else 1, 1 if n_cpu <= 16 * M else 0, extras=True)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample", "other_cpu_paths") if k in cpu}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": wl_desc, "rows_per_side": rows_total, "rows_per_side_per_gpu": rows,
                       "key": "u64 dense 0..N-1 in a permuted row order" if dense else
                              "u64 (splitmix64 of a row permutation; same key set both sides)",
                       "l2": "inputs (2 x %.1f GB) and every intermediate exceed the 126 MB L2; no flush needed" % (rows * 8 / 1e9),
                       "path": path_name(tm),
                       "bucket_join": bool(tm["bucket_join"]), "scatter_levels_per_join": tm["scatter_levels"],
                       "sort_passes_per_join": tm["sort_passes"], "bucket_finishes_per_join": tm["sort_finishes"],
                       "sort_fallbacks": tm["sort_fallbacks"],
                       "parallelism": "hash-range partition x%d" % world},
            "stage_ms_last_step": {k: tm[k] for k in ("hash_ms", "sort_ms", "join_ms", "total_ms")},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "verified": verified,
        }
        if k7:
            line["hash_table_baseline"] = k7
        if nvlink:
            line["nvlink"] = nvlink
        print_line(line)
    if world > 1:
        joiner.close()
        dist.destroy_process_group()


def e2e_u64_single_gpu(args, torch, np, hmj, ctx, local_rank, rows, rkeys, skeys):
    h_r = torch.empty(rows, dtype=torch.int64).pin_memory()
    h_s = torch.empty(rows, dtype=torch.int64).pin_memory()
    h_r.copy_(rkeys)
    h_s.copy_(skeys)
    o_r = torch.empty(rows, dtype=torch.int32).pin_memory()
    o_s = torch.empty(rows, dtype=torch.int32).pin_memory()
    a = [x.numpy().view(np.uint64) for x in (h_r, h_s)] + [x.numpy().view(np.uint32) for x in (o_r, o_s)]
    e2e_steps = max(1, min(args.steps, 3))
    ctx.join_u64_host(*a)  # warm-up (workspace growth)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        m = ctx.join_u64_host(*a)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / e2e_steps
    assert m == rows
    # every pair of the host result joins equal keys, every row once (host side of `verified`)
    ok = bool((a[0][a[2]] == a[1][a[3]]).all()) and np.array_equal(np.sort(a[2]), np.arange(rows, dtype=np.uint32))
    assert ok, "e2e result pairs do not join equal keys / do not cover the rows"
    e2e = {"value": 2 * rows / dt, "unit": UNIT, "h2d_bytes_per_step": 2 * rows * 8,
           "d2h_bytes_per_step": 2 * rows * 4, "ms_per_step": dt * 1e3, "steps": e2e_steps,
           "in_flight": 1, "pairs_checked": rows,
           "api": "hmj_join_u64_host (pinned host buffers), one call at a time"}
    if args.e2e_in_flight > 1:
        try:
            # successive joins from `in_flight` host threads, one context (stream + workspace)
            # each: a join's D2H of its pairs runs beside the next join's H2D of its keys (PCIe is
            # full duplex); every join still copies its own inputs in and its own pairs out.
            lanes = args.e2e_in_flight
            ctxs = [hmj.Context(local_rank, use_torch_stream=False) for _ in range(lanes)]
            outs = [(a[2], a[3])]
            for _ in range(lanes - 1):
                p_r = torch.empty(rows, dtype=torch.int32).pin_memory()
                p_s = torch.empty(rows, dtype=torch.int32).pin_memory()
                outs.append((p_r.numpy().view(np.uint32), p_s.numpy().view(np.uint32)))
            ctx.lib.hmj_release_workspace(ctx.h)  # its workspace goes back before the lanes grow theirs
            torch.cuda.empty_cache()
            for c, (x, y) in zip(ctxs, outs):
                c.join_u64_host(a[0], a[1], x, y)  # workspace growth
            per_lane = max(4, e2e_steps)
            got = [0] * lanes
            lane_errors = []

            def lane(i):
                try:
                    time.sleep(i * dt / lanes)   # staggered start
                    for _ in range(per_lane):
                        got[i] = ctxs[i].join_u64_host(a[0], a[1], *outs[i])
                except Exception as e:  # re-raised on the main thread below
                    lane_errors.append(e)

            th = [threading.Thread(target=lane, args=(i,)) for i in range(lanes)]
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for t in th:
                t.start()
            for t in th:
                t.join()
            torch.cuda.synchronize()
            dtp = (time.perf_counter() - t0) / (per_lane * lanes)
            if lane_errors:
                raise lane_errors[0]
            assert all(g == rows for g in got)
            for x, y in outs:
                assert (a[0][x[:100000]] == a[1][y[:100000]]).all(), "pipelined e2e result pairs do not join equal keys"
            for c in ctxs:
                c.close()
            e2e["pipelined"] = {"value": 2 * rows / dtp, "ms_per_step": dtp * 1e3, "steps": per_lane * lanes,
                                "in_flight": lanes,
                                "api": "%d joins in flight from %d host threads, a context each" % (lanes, lanes)}
        except (RuntimeError, MemoryError, hmj.HmjError) as e:
            e2e["pipelined"] = {"error": str(e)[:200]}
    return e2e


def e2e_u64_multi_gpu(args, torch, dist, ctx, joiner, local_rank, rows, rows_total, rkeys, skeys, barrier, device):
    import ctypes
    numa = bind_to_gpu_numa(torch, local_rank)
    # every rank: pinned host block of both relations -> H2D -> DistJoin.join_u64 -> D2H of this
    # rank's pairs.  Wall clock between barriers, max over ranks.
    h_r = torch.empty(rows, dtype=torch.int64).pin_memory()
    h_s = torch.empty(rows, dtype=torch.int64).pin_memory()
    h_r.copy_(rkeys)
    h_s.copy_(skeys)
    cap = int(rows * 1.25) + 1024
    o_r = torch.empty(cap, dtype=torch.int32).pin_memory()
    o_s = torch.empty(cap, dtype=torch.int32).pin_memory()
    o_ptr = [ctypes.c_void_p(o_r.data_ptr()), ctypes.c_void_p(o_s.data_ptr())]

    def e2e_step():
        rkeys.copy_(h_r, non_blocking=True)
        skeys.copy_(h_s, non_blocking=True)
        m = joiner.join_u64(rkeys, skeys)
        assert m <= cap
        rc = ctx.lib.hmj_join_result_to_host(ctx.h, o_ptr[0], o_ptr[1], cap)
        assert rc == 0, "hmj_join_result_to_host failed"
        torch.cuda.synchronize()
        return m

    e2e_steps = max(1, min(args.steps, 3))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        m = e2e_step()
    barrier()
    dt = (time.perf_counter() - t0) / e2e_steps
    t = torch.tensor([dt], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    t = torch.tensor([m], dtype=torch.int64, device=device)
    dist.all_reduce(t)
    assert int(t.item()) == rows_total
    return {"value": 2 * rows_total / dt, "unit": UNIT, "h2d_bytes_per_step": 2 * rows_total * 8,
            "d2h_bytes_per_step": 2 * rows_total * 4, "ms_per_step": dt * 1e3, "steps": e2e_steps,
            "api": "DistJoin.join_u64 (per-rank pinned host blocks; bytes are whole-job totals)",
            "host_numa_cpus_rank0": numa}


# ------------------------------------------------------------------------------------------
def run_gpu_strings(args, wl):
    """c1 (strgen x strgen join), c2 (hash + sort), c5 (build keys twice, Zipf(1.0) probe); relations
    are generated in HBM by csrc/datagen.cu."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import hashmergejoin_b200 as hmj
    from hashmergejoin_b200 import datagen

    world, rank, local_rank, device = init_dist(args)
    ctx = hmj.Context(local_rank, use_torch_stream=True)
    kind, rows_total, wl_desc = WORKLOADS[wl]
    rows_total = args.rows or rows_total
    if wl == "c5" and world == 8 and not args.rows:
        rows_total = 512 * M
    rows = rows_total // world
    lo = rank * rows
    t_gen = time.perf_counter()
    if kind == "zipf":
        n_keys = rows_total // 2
        gen = datagen.DeviceStrgen(ctx, n_keys)
        r_ids = gen.shuffled_ids(rows, seed=11, copies=2, row_base=lo)
        cdf = torch.cumsum(1.0 / torch.arange(1, n_keys + 1, dtype=torch.float64, device=device), 0)
        cdf /= cdf[-1].clone()
        cdf[-1] = 1.0
        s_ids = gen.zipf_ids(cdf, rows, seed=5, key_seed=13, row_base=lo)
        del cdf
        expect_pairs = 2 * rows_total
    else:
        gen = datagen.DeviceStrgen(ctx, rows_total)
        r_ids = gen.shuffled_ids(rows, seed=1, row_base=lo)
        s_ids = gen.shuffled_ids(rows, seed=2, row_base=lo)
        expect_pairs = rows_total
    roff, rchars = gen.keys(r_ids)
    soff, schars = gen.keys(s_ids) if kind != "sort" else (None, None)
    torch.cuda.synchronize()
    gen_s = time.perf_counter() - t_gen
    key_bytes = (int(roff[-1].item()) + (int(soff[-1].item()) if soff is not None else 0))
    mean_len = key_bytes / (rows * (1 if kind == "sort" else 2))

    joiner = None
    if world > 1:
        assert kind != "sort", "c2 is a single-GPU workload"
        from hashmergejoin_b200.dist import DistJoin
        joiner = DistJoin(ctx, rank, world, string_rows_hint=rows_total)
    out_h = torch.empty(rows, dtype=torch.int64, device=device) if kind == "sort" else None
    out_v = torch.empty(rows, dtype=torch.int32, device=device) if kind == "sort" else None
    pair_keys = {}

    def step():
        if kind == "sort":
            ctx.hash_sort_strings(roff, rchars, out_hash=out_h, out_rowid=out_v)
            return rows
        if joiner is None:
            return ctx.join_strings(roff, rchars, soff, schars)
        return joiner.join_strings(roff, rchars, soff, schars)

    m, elapsed_ms, launches, tm, clocks, barrier = timed_loop(torch, dist, ctx, step, args, world, rank, local_rank, device)
    ms_per_step = elapsed_ms / args.steps
    units = rows_total if kind == "sort" else 2 * rows_total
    value = units / (ms_per_step * 1e-3)
    if world == 1:
        # the library books 16 B per row for hash_strings_kernel (offset read + hash written); the key
        # bytes it also reads are known here: every step hashes every key once
        tm["kernel_bytes"][5] += float(key_bytes) * args.steps
    roofline, kernels = roofline_from(tm, elapsed_ms)

    # ---- verification on the device
    if kind == "sort":
        h = ctx.hash_strings(roff, rchars)
        srt = bool(((out_h[1:] ^ (-2 ** 63)) >= (out_h[:-1] ^ (-2 ** 63))).all().item())   # unsigned order
        perm_ok = bool((h[out_v.to(torch.int64)] == out_h).all().item())
        rows_once = int(torch.bincount(out_v.to(torch.int64), minlength=rows).max().item()) == 1
        verified = {"sorted_ascending": srt, "hash_of_row_matches": perm_ok, "every_row_once": rows_once,
                    "how": "sorted hash column is ascending (unsigned), equals the hash of the row it names, every row id once"}
        assert srt and perm_ok and rows_once, verified
    else:
        if joiner is None:
            out_r, out_s = ctx.join_result_tensors(m)
            kr, ks = r_ids[out_r.to(torch.int64)], s_ids[out_s.to(torch.int64)]
            eq = bool((kr == ks).all().item())
            # distinct pairs: (r, s) as one 64-bit word
            w = (out_r.to(torch.int64) << 32) | (out_s.to(torch.int64) & 0xFFFFFFFF)
            distinct = int(torch.unique(w).numel()) == m if m <= 300 * M else None
            m_tot = m
        else:
            m, (g_r, g_s) = joiner.join_strings(roff, rchars, soff, schars, want_pairs=True)
            # probe side key ids cannot be recomputed from a row number without the CDF search, so
            # gather every rank's id columns instead (4 B per row)
            all_r = torch.empty(rows_total, dtype=torch.int32, device=device)
            all_s = torch.empty(rows_total, dtype=torch.int32, device=device)
            dist.all_gather_into_tensor(all_r, r_ids)
            dist.all_gather_into_tensor(all_s, s_ids)
            kr, ks = all_r[g_r.to(torch.int64) & 0xFFFFFFFF], all_s[g_s.to(torch.int64) & 0xFFFFFFFF]
            eq = bool((kr == ks).all().item())
            w = ((g_r.to(torch.int64) & 0xFFFFFFFF) << 32) | (g_s.to(torch.int64) & 0xFFFFFFFF)
            distinct = int(torch.unique(w).numel()) == m
            t = torch.tensor([m, 0 if (eq and distinct) else 1], dtype=torch.int64, device=device)
            dist.all_reduce(t)
            m_tot = int(t[0].item())
            eq = eq and int(t[1].item()) == 0
            del all_r, all_s
        verified = {"pairs": m_tot, "expected_pairs": expect_pairs, "all_keys_equal": eq, "pairs_distinct": distinct,
                    "how": "every pair's two rows carry the same generator key id (equal id <=> equal string); "
                           "pairs distinct; pair total equals the true join size"}
        assert eq and distinct is not False and m_tot == expect_pairs, verified

    # ---- e2e: host buffers through the C ABI, one call at a time (single GPU)
    e2e = None
    if world == 1 and not args.no_e2e:
        hb = [x.cpu().pin_memory() for x in ((roff, rchars) if kind == "sort" else (roff, rchars, soff, schars))]
        hn = [x.numpy().view(np.uint64) if x.dtype == torch.int64 else x.numpy() for x in hb]
        e2e_steps = max(1, min(args.steps, 3))
        if kind == "sort":
            p_h = torch.empty(rows, dtype=torch.int64).pin_memory().numpy().view(np.uint64)
            p_v = torch.empty(rows, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
            call = lambda: ctx.hash_sort_strings_host(hn[0], hn[1], p_h, p_v)
            d2h = rows * 12
        else:
            o_r = torch.empty(expect_pairs, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
            o_s = torch.empty(expect_pairs, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
            call = lambda: ctx.join_strings_host(hn[0], hn[1], hn[2], hn[3], o_r, o_s)
            d2h = expect_pairs * 8
        call()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            res = call()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        if kind != "sort":
            assert res == expect_pairs
            hr, hs = r_ids.cpu().numpy(), s_ids.cpu().numpy()
            assert (hr[o_r[:10 ** 6]] == hs[o_s[:10 ** 6]]).all(), "e2e pairs do not join equal keys"
        e2e = {"value": units / dt, "unit": "keys/s" if kind == "sort" else UNIT,
               "h2d_bytes_per_step": sum(int(x.numel() * x.element_size()) for x in hb), "d2h_bytes_per_step": d2h,
               "ms_per_step": dt * 1e3, "steps": e2e_steps, "in_flight": 1,
               "api": ("hmj_hash_sort_strings_host" if kind == "sort" else "hmj_join_strings_host") + " (pinned host buffers), one call at a time"}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            n_cpu = cpu_sample_rows(kind, rows_total)
            cpu = cpu_reference_run(kind, n_cpu, 1, 0, extras=True)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample", "other_cpu_paths") if k in cpu}
        line = {
            "metric": METRIC if kind != "sort" else "hash+sorted keys/sec", "value": value,
            "unit": UNIT if kind != "sort" else "keys/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u64 hash of u8 keys",
            "data": "synthetic",
            "config": {"workload": wl_desc, "rows_per_side": rows_total, "rows_per_side_per_gpu": rows,
                       "key": "strgen-shaped strings generated on the device (csrc/datagen.cu), mean %.1f bytes" % mean_len,
                       "distinct_keys": gen.n_keys, "device_generation_s": gen_s,
                       "l2": ("inputs (%.2f GB) exceed the 126 MB L2; no flush needed" % (key_bytes / 1e9 + rows * 8e-9))
                             if key_bytes > 200e6 else
                             "inputs (%.0f MB) FIT the 126 MB L2 and are re-read from it between steps: the step is launch-latency bound, not HBM bound" % (key_bytes / 1e6 + rows * 16e-6),
                       "path": "hash + bucket scatter + segment sort" if kind == "sort" and tm["sort_uniform"] else
                               "hash + onesweep passes" if kind == "sort" else
                               path_name(tm),
                       "bucket_join": bool(tm["bucket_join"]), "scatter_levels_per_join": tm["scatter_levels"],
                       "sort_passes_per_join": tm["sort_passes"], "sort_fallbacks": tm["sort_fallbacks"],
                       "parallelism": "hash-range partition x%d%s" % (world, " (rows + key bytes travel to the owners)" if world > 1 else "")},
            "stage_ms_last_step": {k: tm[k] for k in ("hash_ms", "sort_ms", "join_ms", "total_ms")},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "verified": verified,
        }
        if joiner is not None and joiner.trace:
            log("string exchange phases (rank 0, ms, device-synchronised):")
            for what, ms in joiner.trace_log[-30:]:
                log("   %-36s %8.2f" % (what, ms))
        if joiner is not None:
            line["nvlink"] = {"bytes_sent_to_peers_rank0_per_join": joiner.exchange_bytes,
                              "route": "key gather fused with the exchange: one kernel writes keys, row ids and lengths into the owners' buffers over NVLink"
                                       if joiner.xchg_joins else "NCCL all_to_all_single x3",
                              "fused_exchange_joins": joiner.xchg_joins,
                              "rows_received_rank0": list(joiner.local_rows),
                              "note": "row id (4 B) + length (4 B) + key bytes per row leaving the rank; the head key's owner receives more"}
        print_line(line)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def print_line(line):
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(line), flush=True)
    if _REAL_STDOUT is not None:
        os.dup2(2, 1)


def bind_to_gpu_numa(torch, local_rank):
    """Restrict this process to the CPUs local to its GPU (sysfs local_cpulist of the PCI device)
    so that host buffers pinned afterwards are placed on that NUMA node.  Returns the CPU list, or
    None when the topology is not exposed."""
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        txt = open("/sys/bus/pci/devices/%s/local_cpulist" % bdf).read().strip()
        cpus = set()
        for part in txt.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return txt
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="rows per side (whole job) instead of the workload's own size")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-in-flight", type=int, default=2,
                    help="joins in flight in the extra e2e.pipelined leg (1 = skip it); e2e.value is always one call at a time")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-k7", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: while the run lasts, file descriptor 1 points at stderr
    # (NCCL prints its "NCCL version ..." banner on stdout from native code), print_line restores it
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
