set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "bucket or build or smoke" > gpurun_out/ab_tests.log 2>&1; tail -n 3 gpurun_out/ab_tests.log
python tools/build_prof.py --gib 16 --glen 4000000 --bucket-mib -1 > gpurun_out/ab_prof.log 2>&1; cat gpurun_out/ab_prof.log
exit 0
