#!/bin/bash
# round-1 late run: chunked drop-in batches (bench extra), new parity test, full GPU suite
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r7_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r7_tests.log
tail -n 3 gpurun_out/r7_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r7_bench_cfg2.json 2> gpurun_out/r7_bench_cfg2.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r7_bench_cfg2.json
tail -n 5 gpurun_out/r7_bench_cfg2.err
