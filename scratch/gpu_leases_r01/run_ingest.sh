set -x
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_ingest.py -x -q ) > gpurun_out/ingest_tests.log 2>&1; echo "rc=$?" >> gpurun_out/ingest_tests.log
tail -n 40 gpurun_out/ingest_tests.log
exit 0
