#!/bin/bash
for cfg in "1 16" "2 7" "2 4"; do
  set -- $cfg; iss=$1; ht=$2
  for w in "C1 300" "C1 500" "C1 1000" "C2 400" "C2 700" "C2 1000"; do
    set -- $w
    OTREF_ISSUERS=$iss OTREF_HOST_THREADS=$ht timeout 600 python bench.py --path pipeline --workload $1 --regions $2 --steps 6 --no-cpu-baseline > gpurun_out/p.json 2> gpurun_out/pipe.err
    python - <<PY
import json
try:
    b=json.loads(open("gpurun_out/p.json").read().strip().splitlines()[-1])
    print("issuers $iss host threads $ht $1 $2: %.0f regions/s  %.1f ms" % (b["value"], b["ms_per_step"]))
except Exception as e: print("issuers $iss $1 $2 failed", e)
PY
  done
done
