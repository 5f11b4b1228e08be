timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 1 --n 268435456
python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 1 --n 67108864
python scripts/sort_sweep.py --var HMJ_SORT_HYBRID --cfgs 1 --n 1000000
timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 > gpurun_out/bench_t4.json; python -c "
import json; d=json.loads(open('gpurun_out/bench_t4.json').read()); print(d['value'], d['ms_per_step'], d['stage_ms_last_step'], d['roofline']['frac'], d['e2e']['value'])"
