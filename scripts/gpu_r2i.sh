#!/bin/bash
# 1 GPU: whole suite with the sampling plan / chunked host join / facade changes; dense and sort-merge
# C3 lines; the facade's own cost; the level-width sweep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/r2i_pytest.log
for w in c3-dense c3-sortmerge; do
  python bench.py --workload $w --no-e2e --no-cpu --no-k7 > gpurun_out/r2i_bench_$w.json 2> gpurun_out/r2i_bench_$w.err; echo "$w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2i_bench_$w.json"))
print(d["ms_per_step"], d["value"]/1e9, d["config"]["path"], d["stage_ms_last_step"], d["verified"]["rows_covered"])
for k,v in d["roofline"]["kernels"].items(): print("   ", k, v["launches"], "avg_ms %.3f"%v["avg_ms"], "GB/s %.0f"%v["gbps"])
PY
  tail -2 gpurun_out/r2i_bench_$w.err
done
for args in "str 1000000 1" "str 1000000 16" "u64 16777216 1" "u64 16777216 16"; do
  ./build/facade_bench $args | tee -a gpurun_out/r2i_facade_bench.jsonl
done
python scripts/sweep_bits.py > gpurun_out/r2i_sweep_bits.jsonl 2> gpurun_out/r2i_sweep_bits.err; echo "sweep rc=$?"; cat gpurun_out/r2i_sweep_bits.jsonl | cut -c1-400
python bench.py --workload c2 --no-cpu > gpurun_out/r2i_bench_c2.json 2> gpurun_out/r2i_bench_c2.err; echo "c2 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2i_bench_c2.json')); print(d['ms_per_step'], d['e2e'])"
