# Round-end measurement set: full C3 bench line, reference arm, ncu launch list of the same
# command, one `ncu --set full` capture of the dominant kernels (on the 64M-row workload).
set -x
TAG=${1:-r1}
python bench.py > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err; echo "bench rc=$?"; tail -c 600 gpurun_out/${TAG}_bench_c3.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>&1; echo "ref rc=$?"
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/plain_c3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_c3.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_c3.log 2>&1; echo "ncu launches rc=$?"
python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/plain_mid.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'bucket_scatter|bucket_join|segment_sort|onesweep|join_tile' -s 5 -c 10 -o gpurun_out/${TAG}_prof_mid python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_mid.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_mid.log | cut -c1-200
