# 2-GPU check of bench.py as the driver launches it (read-sharded cfg2 and bin-sharded cfg4), after the e2e change
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
( time timeout 600 $TR bench.py --gpus 2 --steps 5 --warmup 3 ) > gpurun_out/n2f_cfg2.json 2> gpurun_out/n2f_cfg2.err
( time timeout 600 $TR bench.py --gpus 2 --workload cfg4 --steps 3 --warmup 3 --reads 4000000 ) > gpurun_out/n2f_cfg4.json 2> gpurun_out/n2f_cfg4.err
for f in gpurun_out/n2f_*.err; do echo "== $f"; tail -n 5 $f | cut -c1-300; done
python - <<'PY'
import json
for fn in ("n2f_cfg2", "n2f_cfg4"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % fn).read().strip().splitlines()[-1])
        print(fn, d["value"], d["ms_per_step"], json.dumps(d["e2e"]), json.dumps(d.get("dropin_batches")))
    except Exception as e:
        print(fn, "FAILED", e)
PY
exit 0
