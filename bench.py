#!/usr/bin/env python
"""bench.py -- the headline benchmark of the hash -> radix-sort -> merge-join path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A "step" is one complete hash-merge join (hash both relations, sort both, merge-join, write
the (build_row, probe_row) pairs) over one pair of synthetic relations.
  N=1 : BASELINE.json configs[2] -- 256M x 256M 64-bit integer keys, 1:1 match rate.
  N>1 : BASELINE.json configs[3] -- 1B x 1B integer keys over N GPUs (strong scaling): each rank
        owns N-th of both relations, partitions on the top hash bits, one NCCL all-to-all, then
        sort-merges its own hash range.  (launched by torchrun, one rank per GPU)
metric  = joined input tuples/sec = (|R|+|S|) / time of hash+sort+merge.
value   = inputs already resident in HBM (CUDA events).   e2e = host buffers through the C ABI
(hmj_join_u64_host: H2D of the keys, the join, D2H of the pairs).
--impl reference times the UNMODIFIED reference (oracle/_ref, HashMergeJoin with all host
threads) on a bounded sample of the same workload.
"""
import argparse
import json
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
WORKLOADS = {
    # name: (rows per side, description)
    "c3": (256 * 2 ** 20, "hash-merge join 256M x 256M 64-bit int keys, 1:1 match rate, 1 B200"),
    "c4": (2 ** 30, "hash-merge join 1B x 1B int keys radix-partitioned across N B200 with NCCL all-to-all"),
    "small": (2 ** 22, "hash-merge join 4M x 4M 64-bit int keys (smoke-sized)"),
    "mid": (2 ** 26, "hash-merge join 64M x 64M 64-bit int keys (profiling-sized)"),
}
CPU_SAMPLE_ROWS = 16 * 2 ** 20  # rows per side of the bounded CPU sample


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
    pair count to this run's by the ratio of algorithmic bytes; None when no capture exists."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        t = t["kernels"][kernel]
        return t["dram_bytes_per_launch"] * alg_bytes / t["algorithmic_bytes_per_launch"]
    except Exception:
        return None


# ------------------------------------------------------------------------------------------
def cpu_reference_run(rows, steps, warmup, threads=None):
    """HashMergeJoin(R,S,threads) + full iteration of the unmodified reference on `rows` x `rows`
    unique u64 keys; falls back to the oracle port if oracle/_ref is absent."""
    import numpy as np
    from hashmergejoin_b200 import datagen
    from tests import oracle_binding as ob
    keys = datagen.unique_u64(rows, seed=3)
    rk, sk = datagen.permuted(keys, 1), datagen.permuted(keys, 2)
    ref = ob.load_ref()
    times = []
    if ref is not None:
        cores = threads or ref.cores()
        r, s = ob.Keys(u64=rk), ob.Keys(u64=sk)
        for i in range(warmup + steps):
            m, _, _, t = ref.join(r, s, threads=cores, want_pairs=False)
            assert m >= rows - 1, "reference lost matches"
            if i >= warmup:
                times.append(t)
        kind = "reference"
    else:
        orc = ob.load_oracle()
        cores = 1
        import ctypes as C
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            sm = C.c_uint64()
            m = orc.lib.oracle_join_u64(rk.ctypes.data, rows, sk.ctypes.data, rows, None, None, 0, C.byref(sm))
            assert m == rows
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        kind = "port"
    t = sum(times) / len(times)
    return {"value": 2 * rows / t, "unit": UNIT, "cores": cores, "kind": kind, "seconds_per_step": t,
            "sample": f"HashMergeJoin of {rows} x {rows} unique u64 keys (1:1), ctor + full iteration, "
                      f"{cores} thread(s), mean of {len(times)} run(s)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = args.workload or ("c3" if args.gpus == 1 else "c4")
    res = cpu_reference_run(CPU_SAMPLE_ROWS, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["seconds_per_step"] * 1e3,
        "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOADS[wl][1], "sample_rows_per_side": CPU_SAMPLE_ROWS,
                   "note": "reference CPU path on a bounded sample of the workload (host RAM/time)"},
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print_line(line)


# ------------------------------------------------------------------------------------------
def run_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import hashmergejoin_b200 as hmj
    from hashmergejoin_b200 import datagen

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch N>1 with torchrun)"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    ctx = hmj.Context(local_rank, use_torch_stream=True)

    wl = args.workload or ("c3" if world == 1 else "c4")
    rows_total, wl_desc = WORKLOADS[wl]
    rows = rows_total // world  # rows per side owned by this rank
    # same key set on both sides, two different row orders (hashjoin_bench.cc:112-113): 1:1
    rkeys = datagen.unique_u64_torch(rows_total, seed=3, device=device, perm_mult=0x9E3779B1) \
        if world == 1 else None
    skeys = datagen.unique_u64_torch(rows_total, seed=3, device=device, perm_mult=0x85EBCA6B) \
        if world == 1 else None
    joiner = None
    if world > 1:
        from hashmergejoin_b200.dist import DistJoin
        lo = rank * rows
        full_idx = lambda mult: ((torch.arange(lo, lo + rows, dtype=torch.int64, device=device) * mult + 12345)
                                 & (rows_total - 1))
        rkeys = datagen.splitmix64_torch(full_idx(0x9E3779B1) + (3 << 40))
        skeys = datagen.splitmix64_torch(full_idx(0x85EBCA6B) + (3 << 40))
        joiner = DistJoin(ctx, rank, world, rows_total)
    torch.cuda.synchronize()

    def step():
        if joiner is None:
            return ctx.join_u64(rkeys, skeys)
        return joiner.join_u64(rkeys, skeys)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        m = step()
    expect = rows_total
    m_tot = m
    if world > 1:
        t = torch.tensor([m], dtype=torch.int64, device=device)
        dist.all_reduce(t)
        m_tot = int(t.item())
    assert m_tot == expect, f"join produced {m_tot} pairs, expected {expect}"

    sampler = ClockSampler(local_rank)
    barrier()
    ctx.set_profiling(True)
    launches0 = ctx.launch_count()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
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
    ms_per_step = elapsed_ms / args.steps
    value = 2 * rows_total / (ms_per_step * 1e-3)

    # ---- roofline of the dominant kernel class, timed per launch with CUDA events on the
    # launching stream during the timed steps (hmj_set_profiling).  Algorithmic bytes per launch
    # are accumulated by the library at the launch sites.
    peak, peak_src = peaks()
    names = ["onesweep_pass_kernel", "bucket_scatter_tma_kernel", "segment_sort_kernel",
             "join_tile_kernel", "radix_histogram_kernel", "hash_strings_kernel", "bucket_join_kernel"]
    kernels = {}
    for i, nm in enumerate(names):
        ln, ms_sum, by = tm["kernel_launches"][i], tm["kernel_ms"][i], tm["kernel_bytes"][i]
        if ln:
            kernels[nm] = {"launches": ln, "avg_ms": ms_sum / ln, "alg_bytes_per_launch": by / ln,
                           "gbps": by / ms_sum / 1e6 if ms_sum > 0 else None,
                           "share_of_step": ms_sum / elapsed_ms if elapsed_ms else None}
    top = max(kernels, key=lambda k: kernels[k]["share_of_step"]) if kernels else None
    roofline = None
    if top:
        k = kernels[top]
        roofline = {"bound": "hbm", "kernel": top, "achieved": k["gbps"], "peak": peak, "unit": "GB/s",
                    "frac": k["gbps"] / peak, "traffic": ncu_traffic(top, k["alg_bytes_per_launch"]),
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": k["alg_bytes_per_launch"],
                    "avg_launch_ms": k["avg_ms"], "launches_timed": k["launches"],
                    "share_of_step": k["share_of_step"], "kernels": kernels}

    # ---- e2e: host buffers through the C ABI (H2D + join + D2H of the pairs), N=1 only here
    e2e = None
    if world == 1 and not args.no_e2e:
        h_r = torch.empty(rows, dtype=torch.int64).pin_memory()
        h_s = torch.empty(rows, dtype=torch.int64).pin_memory()
        h_r.copy_(rkeys)
        h_s.copy_(skeys)
        o_r = torch.empty(rows, dtype=torch.int32).pin_memory()
        o_s = torch.empty(rows, dtype=torch.int32).pin_memory()
        a = [x.numpy().view(np.uint64) for x in (h_r, h_s)] + [x.numpy().view(np.uint32) for x in (o_r, o_s)]
        del rkeys, skeys
        torch.cuda.empty_cache()
        e2e_steps = max(1, min(args.steps, 3))
        ctx.join_u64_host(*a)  # warm-up (workspace growth)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            m = ctx.join_u64_host(*a)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        assert m == rows
        chk = a[0][a[2][:1000]] == a[1][a[3][:1000]]
        assert chk.all(), "e2e result pairs do not join equal keys"
        e2e = {"value": 2 * rows / dt, "unit": UNIT, "h2d_bytes_per_step": 2 * rows * 8,
               "d2h_bytes_per_step": 2 * rows * 4, "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "in_flight": 1, "api": "hmj_join_u64_host (pinned host buffers)"}
        if args.e2e_in_flight > 1:
            try:
                # successive joins from `in_flight` host threads, one context (stream + workspace)
                # each: a join's D2H of its pairs runs beside the next join's H2D of its keys (PCIe is
                # full duplex); every join still copies its own inputs in and its own pairs out.
                import threading
                lanes = args.e2e_in_flight
                # a stream of its own per lane (the default Context rides torch's current stream)
                ctxs = [hmj.Context(local_rank, use_torch_stream=False) for _ in range(lanes)]
                outs = [(a[2], a[3])]
                for _ in range(lanes - 1):
                    p_r = torch.empty(rows, dtype=torch.int32).pin_memory()
                    p_s = torch.empty(rows, dtype=torch.int32).pin_memory()
                    outs.append((p_r.numpy().view(np.uint32), p_s.numpy().view(np.uint32)))
                ctx.close()  # its workspace goes back before the lanes grow theirs
                for c, (x, y) in zip(ctxs, outs):
                    c.join_u64_host(a[0], a[1], x, y)  # workspace growth
                per_lane = max(4, e2e_steps)
                got = [0] * lanes

                lane_errors = []

                def lane(i):
                    # staggered start: joins entering together would queue on the same copy engine
                    # at the same time and leave it together
                    try:
                        time.sleep(i * dt / lanes)
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
                    assert (a[0][x[:1000]] == a[1][y[:1000]]).all(), "pipelined e2e result pairs do not join equal keys"
                for c in ctxs:
                    c.close()
                e2e.update({"single_call": {"value": e2e["value"], "ms_per_step": e2e["ms_per_step"]},
                            "value": 2 * rows / dtp, "ms_per_step": dtp * 1e3, "steps": per_lane * lanes,
                            "in_flight": lanes,
                            "api": "hmj_join_u64_host (pinned host buffers), %d joins in flight from %d host "
                                   "threads with a context each; single_call = one join at a time" % (lanes, lanes)})
            except (RuntimeError, MemoryError, hmj.HmjError) as e:
                # no room for a second lane (device workspace or pinned host memory): the
                # one-at-a-time figure stands; wrong results are assertion errors and not caught
                e2e["in_flight_error"] = str(e)[:200]
    elif world > 1 and not args.no_e2e:
        # the rank's pinned blocks should live on the NUMA node its GPU hangs off (first touch)
        numa = bind_to_gpu_numa(local_rank)
        # every rank: pinned host block of both relations -> H2D -> DistJoin.join_u64 -> D2H of this
        # rank's pairs.  Wall clock between barriers, max over ranks.
        h_r = torch.empty(rows, dtype=torch.int64).pin_memory()
        h_s = torch.empty(rows, dtype=torch.int64).pin_memory()
        h_r.copy_(rkeys)
        h_s.copy_(skeys)
        cap = int(rows * 1.25) + 1024
        o_r = torch.empty(cap, dtype=torch.int32).pin_memory()
        o_s = torch.empty(cap, dtype=torch.int32).pin_memory()
        import ctypes
        o_ptr = [ctypes.c_void_p(o_r.data_ptr()), ctypes.c_void_p(o_s.data_ptr())]

        def e2e_step():
            rkeys.copy_(h_r, non_blocking=True)
            skeys.copy_(h_s, non_blocking=True)
            # the production path (level-1 scatter over NVLink, local bucket join); the rank's
            # pairs (global row ids) leave from the context's result buffers
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
        e2e = {"value": 2 * rows_total / dt, "unit": UNIT, "h2d_bytes_per_step": 2 * rows_total * 8,
               "d2h_bytes_per_step": 2 * rows_total * 4, "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "api": "DistJoin.join_u64 (per-rank pinned host blocks; bytes are whole-job totals)",
               "host_numa_cpus_rank0": numa}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            cpu = cpu_reference_run(CPU_SAMPLE_ROWS, 1, 0)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": wl_desc, "rows_per_side": rows_total, "rows_per_side_per_gpu": rows,
                       "key": "u64 (splitmix64 of a row permutation; same key set both sides)",
                       "l2": "inputs (2 x %.1f GB) and every intermediate exceed the 126 MB L2; no flush needed" % (rows * 8 / 1e9),
                       "sort_passes_per_join": tm["sort_passes"], "bucket_finishes_per_join": tm["sort_finishes"],
                       "sort_fallbacks": tm["sort_fallbacks"], "uniform_sorts_per_join": tm["sort_uniform"],
                       "parallelism": "hash-range partition x%d" % world},
            "stage_ms_last_step": {k: tm[k] for k in ("hash_ms", "sort_ms", "join_ms", "total_ms")},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }
        if world > 1:
            line["nvlink"] = {"exchange_bytes_sent_rank0_per_step": joiner.exchange_bytes,
                              "note": "12 B per pair leaving the rank, both relations"}
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


def bind_to_gpu_numa(local_rank):
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
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-in-flight", type=int, default=2, help="joins in flight in the e2e leg (1 = one at a time)")
    ap.add_argument("--no-cpu", action="store_true")
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
