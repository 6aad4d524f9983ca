set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_gpu_tests.log
tail -n 30 gpurun_out/r2_gpu_tests.log
( time python bench.py ) > gpurun_out/r2_bench_cfg2.json 2> gpurun_out/r2_bench_cfg2.err
tail -n 5 gpurun_out/r2_bench_cfg2.err
exit 0
