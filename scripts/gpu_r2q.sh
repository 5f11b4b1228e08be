#!/bin/bash
# 8 GPUs: parity (fused string exchange included), C5 at 512M x 512M untraced and traced
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/dist_parity_n8.txt
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k 8_gpus > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2q_pytest.log; grep -c OK gpurun_out/dist_parity_n8.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --workload c5 --steps 3 --warmup 2 > gpurun_out/r2q_c5_n8.json 2> gpurun_out/r2q_c5_n8.err; echo "c5 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2q_c5_n8.json')); print(d['ms_per_step'], d['value']/1e9, d['stage_ms_last_step'], d['verified']['pairs'], d['nvlink'])"
tail -3 gpurun_out/r2q_c5_n8.err
HMJ_DIST_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 8 --workload c5 --steps 2 --warmup 1 > gpurun_out/r2q_c5_n8_trace.json 2> gpurun_out/r2q_c5_n8_trace.err; echo "c5 trace rc=$?"
grep -A20 "string exchange phases" gpurun_out/r2q_c5_n8_trace.err | tail -21
