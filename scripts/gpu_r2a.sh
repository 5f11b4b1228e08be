#!/bin/bash
# round 2, first GPU pass: parity tests, scatter ranking variants, full bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
for r in 0 1 2; do
  HMJ_SCATTER_RANK=$r python bench.py --workload c3 --steps 5 --warmup 3 --no-e2e --no-cpu --no-k7 > gpurun_out/r2a_c3_rank$r.json 2> gpurun_out/r2a_c3_rank$r.err
  echo "rank=$r rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2a_c3_rank$r.json"))
    print(d["ms_per_step"], {k:(round(v["avg_ms"],3), round(v["gbps"])) for k,v in d["roofline"]["kernels"].items()}, d["verified"]["rows_covered"])
except Exception as e:
    print("no line", e)
PY
done
python bench.py > gpurun_out/r2a_bench_c3.json 2> gpurun_out/r2a_bench_c3.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r2a_bench_c3.json
tail -5 gpurun_out/r2a_bench_c3.err
