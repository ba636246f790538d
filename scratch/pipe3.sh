#!/bin/bash
for iss in 1 2; do
  for w in "C2 300" "C2 500" "C2 700" "C1 300" "C1 500"; do
    set -- $w
    OTREF_ISSUERS=$iss timeout 600 python bench.py --path pipeline --workload $1 --regions $2 --steps 6 --no-cpu-baseline > gpurun_out/p.json 2> gpurun_out/pipe.err
    python - <<PY
import json
try:
    b=json.loads(open("gpurun_out/p.json").read().strip().splitlines()[-1])
    print("issuers $iss $1 $2: %.0f regions/s  %.1f ms" % (b["value"], b["ms_per_step"]))
except Exception as e: print("issuers $iss $1 $2 failed", e)
PY
  done
done
