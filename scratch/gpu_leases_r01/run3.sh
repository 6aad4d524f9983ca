set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r3_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3_gpu_tests.log
tail -n 6 gpurun_out/r3_gpu_tests.log
( time python bench.py --workload cfg5 --steps 3 --warmup 3 ) > gpurun_out/r3_bench_cfg5.json 2> gpurun_out/r3_bench_cfg5.err
tail -n 4 gpurun_out/r3_bench_cfg5.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r3_launches_cfg5.csv python bench.py --workload cfg5 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r3_ncu_cfg5.log 2>&1
exit 0
