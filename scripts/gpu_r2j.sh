#!/bin/bash
# 1 GPU: suite again, facade cost, c3-onesweep, c5 after the sample-driven plan, ncu of the string kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r2j_pytest.log
rm -f gpurun_out/r2j_facade_bench.jsonl
for args in "str 1000000 1" "str 1000000 16" "u64 16777216 1" "u64 16777216 16"; do
  ./build/facade_bench $args | tee -a gpurun_out/r2j_facade_bench.jsonl
done
for w in c3-onesweep c5; do
  python bench.py --workload $w --no-e2e --no-cpu --no-k7 > gpurun_out/r2j_bench_$w.json 2> gpurun_out/r2j_bench_$w.err; echo "$w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2j_bench_$w.json"))
print(d["ms_per_step"], d["value"]/1e9, d["config"]["path"], d["stage_ms_last_step"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
PY
  tail -2 gpurun_out/r2j_bench_$w.err
done
python scripts/one_c5.py 16777216 || exit 1
bash scripts/ncu_kernel.sh r2j_hash_strings "hash_strings_kernel" 2 -- python scripts/one_c5.py 16777216
bash scripts/ncu_kernel.sh r2j_join_tile_zipf "join_tile_kernel" 1 -- python scripts/one_c5.py 16777216
