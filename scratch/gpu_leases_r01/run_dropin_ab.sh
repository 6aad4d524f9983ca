#!/bin/bash
# A/B of the small-batch cut rule of the host-buffer query pipeline (100 000-read dyibf_classify calls)
mkdir -p gpurun_out
for cfg in "4 16384" "2 16384" "6 8192" "8 8192" "12 4096" "16 4096"; do
  set -- $cfg
  DYIBF_QUERY_CUTS=$1 DYIBF_QUERY_MIN_ROWS=$2 timeout 300 python bench.py --reads 2000000 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-early-exit --no-ingest \
     > gpurun_out/dropin_ab_$1.json 2> gpurun_out/dropin_ab_$1.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/dropin_ab_$1.json").read().strip().splitlines()[-1])["dropin_batches"]
print("cuts $1 min $2:", round(d["value"]/1e6,2), "M reads/s", round(d["ms_per_call"],3), "ms/call; one chunk", round(d["one_chunk_per_call"]/1e6,2), "pinned", round(d["pinned_input"]/1e6,2), d["kernel_ms_per_call"])
PY
done
