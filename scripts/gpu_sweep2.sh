timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for d in uniform seq dup16 zipf; do python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 0,1 --dist $d; done
python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 0,1 --n 268435456
python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 0,1 --n 1000000
python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 0,1 --n 16777216
