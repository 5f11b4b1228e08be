#!/bin/bash
# 4 GPUs: parity test and C4 through bench.py (the N the driver's scaling run uses and this round had not run yet)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/dist_parity_n4.txt
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k 4_gpus > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2r_pytest.log; grep -c OK gpurun_out/dist_parity_n4.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2r_c4_n4.json 2> gpurun_out/r2r_c4_n4.err; echo "c4 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2r_c4_n4.json"))
print(d["ms_per_step"], d["value"]/1e9, d["stage_ms_last_step"], d["verified"]["rows_covered"], d["e2e"]["ms_per_step"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
print(d["nvlink"]["gbps_out_per_gpu"])
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29582 bench.py --gpus 4 --impl reference --steps 2 --warmup 1 | cut -c1-300
