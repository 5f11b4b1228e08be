#!/bin/bash
# 1 GPU: suite after the scatter-kernel rework (region starts) and the optimistic string join; c3, c5, c2 lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r2l_pytest.log
for w in c3 c5 c2; do
  python bench.py --workload $w --no-e2e --no-cpu --no-k7 > gpurun_out/r2l_bench_$w.json 2> gpurun_out/r2l_bench_$w.err; echo "$w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2l_bench_$w.json"))
print(d["ms_per_step"], d["value"]/1e9, d["config"]["path"], d["stage_ms_last_step"], d["verified"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
PY
  tail -2 gpurun_out/r2l_bench_$w.err
done
