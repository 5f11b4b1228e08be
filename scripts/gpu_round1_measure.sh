set -x
python bench.py > gpurun_out/r1_bench_c3.log 2> gpurun_out/r1_bench_c3.err; echo "bench rc=$?"; tail -3 gpurun_out/r1_bench_c3.log; tail -5 gpurun_out/r1_bench_c3.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_bench_ref.log 2>&1; echo "ref rc=$?"; tail -2 gpurun_out/r1_bench_ref.log
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/plain_c3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c3.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_c3.log 2>&1; echo "ncu launches rc=$?"
python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/plain_mid.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:onesweep -s 18 -c 3 -o gpurun_out/prof_onesweep_r1 python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_mid.log 2>&1; echo "ncu full rc=$?"; tail -5 gpurun_out/ncu_mid.log
ls -la gpurun_out
