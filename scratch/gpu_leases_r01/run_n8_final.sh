# 8-GPU cfg2 with the final bench.py (e2e through dyibf_classify), as the driver launches it
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
( time timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 ) > gpurun_out/n8f_cfg2.json 2> gpurun_out/n8f_cfg2.err
tail -n 6 gpurun_out/n8f_cfg2.err | cut -c1-300
python - <<'PY'
import json
d = json.loads(open("gpurun_out/n8f_cfg2.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], json.dumps(d["e2e"]), json.dumps(d.get("dropin_batches")), d["clocks"])
PY
exit 0
