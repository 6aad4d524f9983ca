set -x
mkdir -p gpurun_out
A="--bins 64 --filter-gib 1 --read-len 100 --genome-len 8000 --reads 10000000 --no-ingest"
( time python bench.py $A ) > gpurun_out/cfg1_bench.json 2> gpurun_out/cfg1_bench.err
tail -n 5 gpurun_out/cfg1_bench.err | cut -c1-300
timeout 600 ncu --set full --import-source on --clock-control none -k regex:query_kernel -s 3 -c 1 -o gpurun_out/cfg1_query -f python bench.py $A --reads 2000000 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-early-exit > gpurun_out/cfg1_ncu.log 2>&1
tail -n 3 gpurun_out/cfg1_ncu.log | cut -c1-200
exit 0
