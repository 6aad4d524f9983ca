set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/f1_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f1_gpu_tests.log
tail -n 6 gpurun_out/f1_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/f1_smoke.log
( time python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/f1_cfg2_ref.json 2> gpurun_out/f1_cfg2_ref.err
( time python bench.py ) > gpurun_out/f1_cfg2.json 2> gpurun_out/f1_cfg2.err
( time python bench.py --workload cfg3 --steps 3 --warmup 3 --reads 5000000 ) > gpurun_out/f1_cfg3.json 2> gpurun_out/f1_cfg3.err
( time python bench.py --workload cfg5 --steps 3 --warmup 3 ) > gpurun_out/f1_cfg5.json 2> gpurun_out/f1_cfg5.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/f1_launches_cfg2.csv python bench.py --reads 2000000 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-early-exit --no-ingest > gpurun_out/f1_ncu.log 2>&1
tail -n 3 gpurun_out/f1_smoke.log gpurun_out/*.err | cut -c1-200
exit 0
