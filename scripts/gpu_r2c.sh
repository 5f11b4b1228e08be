#!/bin/bash
# ncu --set full captures of the two kernels of the C3 step (64M x 64M rows; after the plain run exits 0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/one_join.py 67108864 || exit 1
bash scripts/ncu_kernel.sh r2c_join_table "bucket_join_table_kernel" 1 -- python scripts/one_join.py 67108864
bash scripts/ncu_kernel.sh r2c_scatter "bucket_scatter_tma_kernel" 5 -- python scripts/one_join.py 67108864
ls -la gpurun_out/*.ncu-rep
