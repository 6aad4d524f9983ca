set -x
mkdir -p gpurun_out
( time python bench.py ) > gpurun_out/r5_bench_cfg2.json 2> gpurun_out/r5_bench_cfg2.err
tail -n 4 gpurun_out/r5_bench_cfg2.err
exit 0
