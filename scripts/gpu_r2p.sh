#!/bin/bash
# 2 GPUs: parity with the fused string exchange, then C5 (128M x 128M) traced
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/dist_parity_n2.txt
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k one_local_level > gpurun_out/r2p_pytest_dist.log 2>&1; echo "dist pytest rc=$?"; tail -12 gpurun_out/r2p_pytest_dist.log; grep strings gpurun_out/dist_parity_n2.txt
HMJ_DIST_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29563 bench.py --gpus 2 --workload c5 --rows 134217728 --steps 2 --warmup 1 > gpurun_out/r2p_c5_n2.json 2> gpurun_out/r2p_c5_n2.err; echo "c5 n2 rc=$?"
grep -A18 "string exchange phases" gpurun_out/r2p_c5_n2.err | tail -19; tail -3 gpurun_out/r2p_c5_n2.err
python -c "
import json; d=json.load(open('gpurun_out/r2p_c5_n2.json')); print(d['ms_per_step'], d['verified']['pairs'], d['nvlink'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29564 bench.py --gpus 2 --workload c5 --rows 134217728 --steps 3 --warmup 2 > gpurun_out/r2p_c5_n2_plain.json 2> /dev/null; python -c "
import json; d=json.load(open('gpurun_out/r2p_c5_n2_plain.json')); print('untraced', d['ms_per_step'], d['verified']['pairs'])"
