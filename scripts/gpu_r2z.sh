#!/bin/bash
# 1 GPU, last pass of round 2 on the final code: the default bench line and the ncu launch list of
# the same command (the whole -m gpu suite ran on this code in the pass before: profiles/r2z9_gpu_tests.log)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py > gpurun_out/r2z10_bench_c3.json 2> gpurun_out/r2z10_bench_c3.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2z10_bench_c3.json"))
print("c3", d["ms_per_step"], d["value"]/1e9, "e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"]/1e9, "frac", d["roofline"]["frac"], d["verified"]["rows_covered"], d["clocks"])
PY
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2z10_launches_c3.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-k7 > gpurun_out/r2z10_ncu_launches.log 2>&1; echo "launch list rc=$?"
