#!/bin/bash
# 2 GPUs: C4 with the level-1 width rule; S's level 1 with 1 vs 2 CTAs per SM
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for c in 1 2; do
  HMJ_DIST_S_CTAS=$c timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$c bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/r2g_c4_n2_s$c.json 2> gpurun_out/r2g_c4_n2_s$c.err; echo "s_ctas=$c rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2g_c4_n2_s$c.json"))
print(d["ms_per_step"], d["value"]/1e9, d["stage_ms_last_step"], d["verified"]["rows_covered"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
print(d["nvlink"]["gbps_out_per_gpu"])
PY
done
