set -x
mkdir -p gpurun_out
python tools/build_prof.py > gpurun_out/build_prof.log 2>&1; cat gpurun_out/build_prof.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'split|apply' -c 6 -o gpurun_out/build_v5 -f python tools/build_prof.py > gpurun_out/build_prof_ncu.log 2>&1
tail -n 5 gpurun_out/build_prof_ncu.log
ls -la gpurun_out/*.ncu-rep
exit 0
