#!/bin/bash
# round 2, second GPU pass (1 GPU): table bucket join vs counting-sort bucket join
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "operator or join or smoke or device_generated" > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2b_pytest.log
for v in table sort; do
  HMJ_BUCKET_JOIN=$v python bench.py --workload c3 --steps 5 --warmup 3 --no-e2e --no-cpu --no-k7 > gpurun_out/r2b_c3_$v.json 2> gpurun_out/r2b_c3_$v.err
  echo "join=$v rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2b_c3_$v.json"))
    print(d["ms_per_step"], {k:(round(v["avg_ms"],3), round(v["gbps"])) for k,v in d["roofline"]["kernels"].items()}, d["verified"]["rows_covered"])
except Exception as e:
    print("no line", e)
PY
  tail -3 gpurun_out/r2b_c3_$v.err
done
