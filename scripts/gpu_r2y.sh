#!/bin/bash
# 8 GPUs: C5 at 512M x 512M with the final string send kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 8 --workload c5 --steps 3 --warmup 2 > gpurun_out/r2y_c5_n8.json 2> gpurun_out/r2y_c5_n8.err; echo "c5 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2y_c5_n8.json')); print(d['ms_per_step'], d['value']/1e9, d['stage_ms_last_step'], d['verified']['pairs'], d['verified']['all_keys_equal'], d['nvlink']['fused_exchange_joins'])"
tail -2 gpurun_out/r2y_c5_n8.err
