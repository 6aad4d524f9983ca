set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r4_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r4_gpu_tests.log
tail -n 15 gpurun_out/r4_gpu_tests.log
( time python bench.py ) > gpurun_out/r4_bench_cfg2.json 2> gpurun_out/r4_bench_cfg2.err
tail -n 4 gpurun_out/r4_bench_cfg2.err
exit 0
