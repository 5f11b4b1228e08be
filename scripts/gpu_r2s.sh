#!/bin/bash
# 1 GPU, end of round 2: whole suite, the default bench line, the reference arm, the ncu launch list of
# the same command and `ncu --set full` captures of the four kernels that carry the configurations
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2s_pytest.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2s_bench_reference.json 2> gpurun_out/r2s_bench_reference.err; echo "reference rc=$?"
python bench.py > gpurun_out/r2s_bench_c3.json 2> gpurun_out/r2s_bench_c3.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2s_bench_c3.json")); r=json.load(open("gpurun_out/r2s_bench_reference.json"))
print("c3", d["ms_per_step"], d["value"]/1e9, "e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"]/1e9, "ref", r["value"]/1e9, "ratio e2e", d["e2e"]["value"]/r["value"], "frac", d["roofline"]["frac"], d["verified"]["rows_covered"])
PY
for w in c1 c2 c5; do python bench.py --workload $w > gpurun_out/r2s_bench_$w.json 2> gpurun_out/r2s_bench_$w.err; echo "$w rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2s_bench_$w.json')); print('$w', d['ms_per_step'], d['value']/1e9, d['e2e']['ms_per_step'], d['roofline']['kernel'], round(d['roofline']['frac'],3))"; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2s_launches_c3.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-k7 > gpurun_out/r2s_ncu_launches.log 2>&1; echo "launch list rc=$?"
bash scripts/ncu_kernel.sh r2s_join_table "bucket_join_table_kernel" 1 -- python scripts/one_join.py 67108864
bash scripts/ncu_kernel.sh r2s_scatter "bucket_scatter_tma_kernel" 5 -- python scripts/one_join.py 67108864
bash scripts/ncu_kernel.sh r2s_hash_strings "hash_strings_tma_kernel" 2 -- python scripts/one_c5.py 16777216
bash scripts/ncu_kernel.sh r2s_join_tile_zipf "join_tile_kernel" 1 -- python scripts/one_c5.py 16777216
