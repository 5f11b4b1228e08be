#!/bin/bash
# 8 GPUs: parity, C4 with the relations whole / in 2 parts, C5 at 512M x 512M with the exchange's phases traced
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/dist_parity_n8.txt
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k 8_gpus > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2n_pytest.log; grep -c OK gpurun_out/dist_parity_n8.txt
for p in 1 2; do
  HMJ_DIST_PARTS=$p timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2954$p bench.py --gpus 8 --steps 5 --warmup 3 $( [ $p = 2 ] && echo --no-e2e ) > gpurun_out/r2n_c4_n8_p$p.json 2> gpurun_out/r2n_c4_n8_p$p.err; echo "parts=$p rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2n_c4_n8_p$p.json"))
print(d["ms_per_step"], d["value"]/1e9, d["stage_ms_last_step"], d["verified"]["rows_covered"], (d.get("e2e") or {}).get("ms_per_step"))
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
print(d["nvlink"]["gbps_out_per_gpu"])
PY
done
HMJ_DIST_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus 8 --workload c5 --steps 2 --warmup 1 > gpurun_out/r2n_c5_n8_trace.json 2> gpurun_out/r2n_c5_n8_trace.err; echo "c5 rc=$?"
grep -A32 "string exchange phases" gpurun_out/r2n_c5_n8_trace.err | tail -34
python -c "
import json; d=json.load(open('gpurun_out/r2n_c5_n8_trace.json')); print(d['ms_per_step'], d['stage_ms_last_step'], d['verified']['pairs'])"
