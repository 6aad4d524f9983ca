#!/bin/bash
# round-1 late run: full GPU suite (guard-canary tests included) and the cfg2 bench line
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_guards.py -m gpu -q > gpurun_out/r10_guards.log 2>&1; tail -n 30 gpurun_out/r10_guards.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r10_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r10_tests.log
tail -n 3 gpurun_out/r10_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r10_bench_cfg2.json 2> gpurun_out/r10_bench_cfg2.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r10_bench_cfg2.json
tail -n 5 gpurun_out/r10_bench_cfg2.err
