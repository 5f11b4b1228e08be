#!/bin/bash
# 2 GPUs: multi-GPU parity tests, then C4 (1B x 1B) through bench.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/r2e_pytest.log
cat gpurun_out/dist_parity_n2.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2e_c4_n2.json 2> gpurun_out/r2e_c4_n2.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/r2e_c4_n2.json; tail -5 gpurun_out/r2e_c4_n2.err
