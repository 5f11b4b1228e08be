#!/bin/bash
# 1 GPU: whole parity suite, full C3 line (chunked host-buffer join), string workloads
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r2f_pytest.log
python bench.py > gpurun_out/r2f_bench_c3.json 2> gpurun_out/r2f_bench_c3.err; echo "c3 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2f_bench_c3.json"))
print("c3", d["ms_per_step"], d["value"]/1e9, "e2e", d["e2e"]["ms_per_step"], d["e2e"].get("pipelined"), d["roofline"]["frac"])
PY
tail -3 gpurun_out/r2f_bench_c3.err
for w in c1 c2 c5; do
  python bench.py --workload $w > gpurun_out/r2f_bench_$w.json 2> gpurun_out/r2f_bench_$w.err; echo "$w rc=$?"
  tail -c 1800 gpurun_out/r2f_bench_$w.json; echo; tail -4 gpurun_out/r2f_bench_$w.err
done
