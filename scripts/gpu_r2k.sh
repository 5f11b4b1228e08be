#!/bin/bash
# 1 GPU: facade tests + what the facade costs a caller (copied vs borrowed inputs), K7 after the word-wise compare
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "facade or hashjoin_baseline" > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2k_pytest.log
rm -f gpurun_out/r2k_facade_bench.jsonl
for b in facade_bench facade_bench_borrow; do
  for args in "str 1000000 1" "str 1000000 16" "u64 16777216 1" "u64 16777216 16"; do
    ./build/$b $args | tee -a gpurun_out/r2k_facade_bench.jsonl
  done
done
