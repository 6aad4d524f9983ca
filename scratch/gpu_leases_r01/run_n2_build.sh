set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -q -k "bin_sharded_build" > gpurun_out/n2_shardtest.log 2>&1; tail -n 3 gpurun_out/n2_shardtest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513"
( time timeout 900 $TR bench.py --gpus 2 --workload cfg5 --steps 3 --warmup 3 --no-cpu-baseline ) > gpurun_out/n2_cfg5.json 2> gpurun_out/n2_cfg5.err
tail -n 12 gpurun_out/n2_cfg5.err
exit 0
