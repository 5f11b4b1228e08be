#!/bin/bash
# 8 GPUs: C4 with bulk stores to the peers vs plain stores (final code)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for bk in 1 0; do
  HMJ_DIST_BULK=$bk timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2960$bk bench.py --gpus 8 --steps 5 --warmup 3 $( [ $bk = 0 ] && echo --no-e2e ) > gpurun_out/r2v_c4_n8_b$bk.json 2> gpurun_out/r2v_c4_n8_b$bk.err; echo "bulk=$bk rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2v_c4_n8_b$bk.json"))
print(d["ms_per_step"], d["value"]/1e9, d["verified"]["rows_covered"], {k[:32]:(round(v["avg_ms"],3)) for k,v in d["roofline"]["kernels"].items()}, round(d["nvlink"]["gbps_out_per_gpu"]), (d.get("e2e") or {}).get("ms_per_step"))
PY
done
