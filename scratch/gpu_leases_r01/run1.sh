set -x
mkdir -p gpurun_out
nproc; free -g | head -2
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r1_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r1_smoke.log
( time python bench.py ) > gpurun_out/r1_bench_cfg2.json 2> gpurun_out/r1_bench_cfg2.err
( time python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/r1_bench_cfg2_ref.json 2> gpurun_out/r1_bench_cfg2_ref.err
( time python bench.py --workload cfg5 --steps 3 --warmup 3 ) > gpurun_out/r1_bench_cfg5.json 2> gpurun_out/r1_bench_cfg5.err
( time python bench.py --workload cfg3 --steps 3 --warmup 3 --reads 5000000 ) > gpurun_out/r1_bench_cfg3.json 2> gpurun_out/r1_bench_cfg3.err
tail -3 gpurun_out/r1_gpu_tests.log gpurun_out/r1_smoke.log gpurun_out/*.err
