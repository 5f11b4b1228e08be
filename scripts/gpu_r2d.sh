#!/bin/bash
# table bucket join: 2^13 vs 2^14 slots (chain flag in the slot, fused start-up loads)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "operator or join or device_generated" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2d_pytest.log
for v in 13 14; do
  HMJ_BUCKET_TABLE_BITS=$v python bench.py --workload c3 --steps 5 --warmup 3 --no-e2e --no-cpu --no-k7 > gpurun_out/r2d_c3_$v.json 2> gpurun_out/r2d_c3_$v.err
  echo "tbits=$v rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2d_c3_$v.json"))
    print(d["ms_per_step"], {k:(round(v["avg_ms"],3), round(v["gbps"])) for k,v in d["roofline"]["kernels"].items()}, d["verified"]["rows_covered"])
except Exception as e:
    print("no line", e)
PY
  tail -3 gpurun_out/r2d_c3_$v.err
done
