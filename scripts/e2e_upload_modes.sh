#!/bin/bash
# Dev tool: pipelined e2e of bench.py under different in-flight depths of the SM upload kernel (bff_upload).
mkdir -p gpurun_out
for cfg in "32 4" "32 1" "16 1" "8 1" "4 1" "8 2" "16 2" "4 4"; do
  set -- $cfg
  BFF_UPLOAD_BLOCKS=$1 BFF_UPLOAD_UNROLL=$2 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --upload sm > gpurun_out/up.json 2> gpurun_out/up.err
  python - "$cfg" <<'PY'
import json,sys
m=sys.argv[1]
try:
    d=json.load(open("gpurun_out/up.json"))
    print("blocks,unroll", m, "| step %.3f"%d["ms_per_step"], "e2e %.3f"%d["e2e"]["ms_per_step"], "serial %.3f"%d["e2e"]["serial_ms_per_step"], d["e2e"]["upload_gbs_probe"], flush=True)
except Exception as e:
    print(m, "failed", e)
PY
done
