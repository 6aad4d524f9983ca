#!/bin/bash
# bench.py with e2e through dyibf_classify: cfg2 headline, cfg3 short, smoke
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r9_bench_cfg2.json 2> gpurun_out/r9_bench_cfg2.err; echo "bench cfg2 rc=$?"
tail -n 5 gpurun_out/r9_bench_cfg2.err
timeout 600 python bench.py --workload cfg3 --reads 2000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r9_bench_cfg3_short.json 2> gpurun_out/r9_bench_cfg3.err; echo "bench cfg3 rc=$?"
tail -n 5 gpurun_out/r9_bench_cfg3.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r9_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/r9_smoke.log
python - <<'PY'
import json
for fn in ("r9_bench_cfg2", "r9_bench_cfg3_short"):
    d = json.loads(open("gpurun_out/%s.json" % fn).read().strip().splitlines()[-1])
    print(fn, d["value"], d["ms_per_step"], json.dumps(d["e2e"]))
PY
