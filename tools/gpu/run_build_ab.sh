set -x
mkdir -p gpurun_out
for m in 1 0; do
  export DYIBF_SPLIT_MATCH=$m
  python -m pytest tests/test_gpu_parity.py -x -q -k "bucket or build or smoke" > gpurun_out/ab_tests_$m.log 2>&1; tail -n 3 gpurun_out/ab_tests_$m.log
  python tools/build_prof.py --gib 16 --glen 4000000 --bucket-mib -1 > gpurun_out/ab_prof_$m.log 2>&1; cat gpurun_out/ab_prof_$m.log
done
exit 0
