#!/bin/bash
# 2 GPUs: level 1 over NVLink with bulk async stores to the peers vs plain stores; parity first
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/dist_parity_n2.txt
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k 2_gpus > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2u_pytest.log; grep -c OK gpurun_out/dist_parity_n2.txt
for bk in 1 0; do
  HMJ_DIST_BULK=$bk timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2959$bk bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/r2u_c4_n2_b$bk.json 2> gpurun_out/r2u_c4_n2_b$bk.err; echo "bulk=$bk rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2u_c4_n2_b$bk.json"))
print(d["ms_per_step"], d["value"]/1e9, d["verified"]["rows_covered"], {k[:32]:(round(v["avg_ms"],3)) for k,v in d["roofline"]["kernels"].items()}, round(d["nvlink"]["gbps_out_per_gpu"]))
PY
  tail -2 gpurun_out/r2u_c4_n2_b$bk.err
done
