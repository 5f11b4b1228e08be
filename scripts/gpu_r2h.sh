#!/bin/bash
# 8 GPUs: multi-GPU parity test, C4 (1B x 1B) and C5 (512M x 512M Zipf strings) through bench.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k 8_gpus > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r2h_pytest.log
cat gpurun_out/dist_parity_n8.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2h_c4_n8.json 2> gpurun_out/r2h_c4_n8.err; echo "c4 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2h_c4_n8.json"))
print(d["ms_per_step"], d["value"]/1e9, d["stage_ms_last_step"], d["verified"], d["e2e"]["ms_per_step"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
print(d["nvlink"])
PY
tail -3 gpurun_out/r2h_c4_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --workload c5 --steps 3 --warmup 2 > gpurun_out/r2h_c5_n8.json 2> gpurun_out/r2h_c5_n8.err; echo "c5 rc=$?"
tail -c 2500 gpurun_out/r2h_c5_n8.json; tail -5 gpurun_out/r2h_c5_n8.err
