set -x
mkdir -p gpurun_out
( python -m pytest tests/test_gpu_ingest.py -x -q ) > gpurun_out/ingest_tests.log 2>&1; echo "rc=$?" >> gpurun_out/ingest_tests.log
tail -n 25 gpurun_out/ingest_tests.log
python tools/ingest_bench.py > gpurun_out/ingest_bench.log 2>&1
python tools/ingest_bench.py --fasta >> gpurun_out/ingest_bench.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ingest_launches.csv python tools/ingest_bench.py --calls 1 > gpurun_out/ingest_ncu.log 2>&1
cat gpurun_out/ingest_bench.log
exit 0
