# refresh of the cfg3 and cfg5 lines with the final library of the round (1 GPU)
mkdir -p gpurun_out
( time timeout 600 python bench.py --workload cfg3 --steps 5 --warmup 3 ) > gpurun_out/rf_cfg3.json 2> gpurun_out/rf_cfg3.err; echo "cfg3 rc=$?"
( time timeout 600 python bench.py --workload cfg5 --steps 5 --warmup 3 ) > gpurun_out/rf_cfg5.json 2> gpurun_out/rf_cfg5.err; echo "cfg5 rc=$?"
for f in gpurun_out/rf_*.err; do echo "== $f"; tail -n 4 $f | cut -c1-300; done
python - <<'PY'
import json
for fn in ("rf_cfg3", "rf_cfg5"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % fn).read().strip().splitlines()[-1])
        print(fn, d["value"], d["unit"], d["ms_per_step"], json.dumps(d["e2e"])[:300], json.dumps(d["roofline"])[:400])
    except Exception as e:
        print(fn, "FAILED", e)
PY
exit 0
