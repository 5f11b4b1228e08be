#!/usr/bin/env python
"""Turns the raw artefacts of scripts/gpu_round1_measure.sh (round 1; removed) (gpurun_out/<tag>_*) into the small
tracked summaries under profiles/:
  <tag>_bench_c3.json / <tag>_bench_reference.json   the bench lines as printed
  <tag>_launches_c3.csv + <tag>_launch_shares.txt    ncu launch list and per-kernel share of step
  <tag>_ncu_full.json                                 key `ncu --set full` metrics per kernel
  <tag>_top_stalls_<kernel>.txt                       hottest SASS lines by stall samples
  roofline_traffic.json                               DRAM bytes per onesweep launch (bench.py)
usage: python scripts/digest_profiles.py <tag>   (needs `ncu` on PATH to read the .ncu-rep)"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
]


def short(name):
    n = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    return n.strip()


def launch_shares(tag):
    src = os.path.join(OUT, f"{tag}_launches_c3.csv")
    if not os.path.exists(src):
        return
    shutil.copy(src, os.path.join(PROF, f"{tag}_launches_c3.csv"))
    rows = list(csv.reader(open(src)))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr = rows[hi]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hi + 1:]:
        if len(r) <= mv:
            continue
        v = float(r[mv].replace(",", ""))
        ms = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}.get(r[mu], 1e-6) * v
        agg[short(r[kn])[:70]][0] += 1
        agg[short(r[kn])[:70]][1] += ms
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(PROF, f"{tag}_launch_shares.txt"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none, bench.py --steps 2 --warmup 1 "
                "--no-e2e --no-cpu (C3); cold-cache serialised times: compare SHARES\n")
        f.write("# includes torch kernels that generate the synthetic input (outside the timed region)\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k:72s} n={v[0]:4d} total_ms={v[1]:9.3f} avg_ms={v[1] / v[0]:8.4f} share={v[1] / tot:.3f}\n")


def ncu_full(tag):
    rep = os.path.join(OUT, f"{tag}_prof_mid.ncu-rep")
    if not os.path.exists(rep) or not shutil.which("ncu"):
        return
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    per_kernel = collections.OrderedDict()
    for ki, r in enumerate(rows[2:]):
        name = short(r[hdr.index("Kernel Name")])
        d = {"launch_index_in_report": ki}
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                d[k] = f"{r[i]} {units[i]}".strip()
        per_kernel.setdefault(name, []).append(d)
    json.dump({"capture": "ncu --set full --clock-control none --import-source on; bench.py --workload mid "
                          "(64M x 64M u64 keys)", "kernels": per_kernel},
              open(os.path.join(PROF, f"{tag}_ncu_full.json"), "w"), indent=1)
    # DRAM traffic per launch for bench.py's roofline.traffic, per kernel class: median over the
    # captured launches; algorithmic bytes of a 64M-row launch (bench.py scales by their ratio)
    def num(s):
        v, u = s.split()[0], s.split()[1]
        return float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
    n_mid = 64 * 2 ** 20
    alg = {"onesweep_pass_kernel": n_mid * 24.0, "bucket_scatter_tma_kernel": n_mid * 22.0,
           "segment_sort_kernel": n_mid * 24.0, "bucket_join_kernel": 2 * n_mid * 12.0 + n_mid * 8.0,
           "join_tile_kernel": 2 * n_mid * 12.0 + n_mid * 8.0}
    traffic = {}
    for k, v in per_kernel.items():
        base = k.split("<")[0]
        if base in alg:
            tr = sorted(num(d["dram__bytes_read.sum"]) + num(d["dram__bytes_write.sum"]) for d in v)
            traffic[base] = {"dram_bytes_per_launch": tr[len(tr) // 2],
                             "algorithmic_bytes_per_launch": alg[base], "launches_captured": len(tr)}
    if traffic:
        json.dump({"source": f"profiles/{tag}_ncu_full.json",
                   "note": "ncu --set full on bench.py --workload mid (64M x 64M); bucket_scatter: mean of "
                           "level 1 (20 B/pair) and level 2 (24 B/pair)", "kernels": traffic},
                  open(os.path.join(PROF, "roofline_traffic.json"), "w"), indent=1)
    # top stalls of the first launch of each kernel
    seen = set()
    for ki, r in enumerate(rows[2:]):
        name = short(r[hdr.index("Kernel Name")]).split("<")[0]
        if name in seen:
            continue
        seen.add(name)
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(ki), "--launch-count", "1"], capture_output=True, text=True).stdout
        tmp = os.path.join(OUT, "_src.csv")
        open(tmp, "w").write(src)
        top = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_top_stalls.py"), tmp, "40"],
                             capture_output=True, text=True).stdout
        lines = top.splitlines()
        dedup = lines[:2] + [l for i, l in enumerate(lines[2:]) if i % 2 == 0]
        open(os.path.join(PROF, f"{tag}_top_stalls_{name}.txt"), "w").write("\n".join(dedup) + "\n")


def main():
    tag = sys.argv[1]
    os.makedirs(PROF, exist_ok=True)
    for f in (f"{tag}_bench_c3.json", f"{tag}_bench_reference.json"):
        p = os.path.join(OUT, f)
        if os.path.exists(p):
            shutil.copy(p, os.path.join(PROF, f))
    launch_shares(tag)
    ncu_full(tag)
    print("\n".join(sorted(x for x in os.listdir(PROF) if x.startswith(tag) or x == "roofline_traffic.json")))


if __name__ == "__main__":
    main()
