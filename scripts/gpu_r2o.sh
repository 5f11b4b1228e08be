#!/bin/bash
# 2 GPUs: string exchange after the row-span gather (traced), parity; pipelined hash kernel vs plain on c2 / c5
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "hash or gather or strings" > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2o_pytest.log
HMJ_HASH_TMA=1 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "hash or strings" > gpurun_out/r2o_pytest_tma.log 2>&1; echo "pytest tma rc=$?"; tail -3 gpurun_out/r2o_pytest_tma.log
for t in 0 1; do
  HMJ_HASH_TMA=$t python bench.py --workload c2 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c2 tma=$t', d['ms_per_step'], d['verified']['sorted_ascending'], {k:(round(v['avg_ms'],3), round(v['gbps'])) for k,v in d['roofline']['kernels'].items()})"
done
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k one_local_level > gpurun_out/r2o_pytest_dist.log 2>&1; echo "dist pytest rc=$?"; tail -3 gpurun_out/r2o_pytest_dist.log
HMJ_DIST_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --workload c5 --rows 134217728 --steps 2 --warmup 1 > gpurun_out/r2o_c5_n2.json 2> gpurun_out/r2o_c5_n2.err; echo "c5 n2 rc=$?"
grep -A16 "string exchange phases" gpurun_out/r2o_c5_n2.err | tail -17
python -c "
import json; d=json.load(open('gpurun_out/r2o_c5_n2.json')); print(d['ms_per_step'], d['verified']['pairs'])"
