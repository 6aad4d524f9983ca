# 8-GPU runs (BASELINE.json configs[1], [3], [4]); one rank per GPU over NCCL
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
( time timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 ) > gpurun_out/n8_cfg2.json 2> gpurun_out/n8_cfg2.err
( time timeout 600 $TR bench.py --gpus 8 --workload cfg4 --steps 5 --warmup 3 --reads 10000000 ) > gpurun_out/n8_cfg4.json 2> gpurun_out/n8_cfg4.err
( time timeout 900 $TR bench.py --gpus 8 --workload cfg5 --steps 3 --warmup 3 --no-cpu-baseline ) > gpurun_out/n8_cfg5.json 2> gpurun_out/n8_cfg5.err
for f in gpurun_out/n8_*.err; do echo "== $f"; tail -n 5 $f | cut -c1-300; done
exit 0
