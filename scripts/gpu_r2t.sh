#!/bin/bash
# ncu --set full of the kernels not yet looked at instruction by instruction: segment sort (C2), stable pass and
# merge-join tile (C5), and the table join after the staging swizzle
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/one_sort.py 67108864 || exit 1
bash scripts/ncu_kernel.sh r2t_segment_sort "segment_sort_kernel" 1 -- python scripts/one_sort.py 67108864
bash scripts/ncu_kernel.sh r2t_onesweep "onesweep_pass_kernel" 10 -- python scripts/one_c5.py 16777216
bash scripts/ncu_kernel.sh r2t_join_table "bucket_join_table_kernel" 1 -- python scripts/one_join.py 67108864
