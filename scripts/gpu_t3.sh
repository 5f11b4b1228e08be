timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 > gpurun_out/bench_t3.json; cat gpurun_out/bench_t3.json | cut -c1-1500
