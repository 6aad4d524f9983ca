# ncu: launch list of the final bench command, and one --set full capture of the small kernels of the step (pack, compaction)
set -x
mkdir -p gpurun_out
CMD="python bench.py --reads 2000000 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-early-exit --no-ingest --no-dropin"
timeout 300 $CMD > gpurun_out/nc_plain.json 2> gpurun_out/nc_plain.err; echo "plain rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/nc_launches.csv $CMD > gpurun_out/nc_l.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pack_kernel|compact' --launch-skip 8 -c 4 -o gpurun_out/nc_small -f $CMD > gpurun_out/nc_s.log 2>&1; echo "set full rc=$?"
ncu -i gpurun_out/nc_small.ncu-rep --page raw --csv > gpurun_out/nc_small_raw.csv 2> /dev/null
ls -la gpurun_out/ | tail -n 8
tail -n 3 gpurun_out/nc_s.log | cut -c1-300
exit 0
