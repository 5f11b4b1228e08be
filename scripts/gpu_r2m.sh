#!/bin/bash
# 2 GPUs: parity of the part-wise multi-GPU join, then C4 with 1 / 2 / 4 parts per relation
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/dist_parity_n2.txt
timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k 2_gpus > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r2m_pytest.log; grep -c OK gpurun_out/dist_parity_n2.txt
HMJ_DIST_PARTS=3 timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k one_local_level > gpurun_out/r2m_pytest3.log 2>&1; echo "pytest parts=3 rc=$?"
for p in 1 2 4; do
  HMJ_DIST_PARTS=$p timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$p bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/r2m_c4_n2_p$p.json 2> gpurun_out/r2m_c4_n2_p$p.err; echo "parts=$p rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2m_c4_n2_p$p.json"))
print(d["ms_per_step"], d["value"]/1e9, d["stage_ms_last_step"], d["verified"]["rows_covered"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
PY
done
