#!/bin/bash
for rep in 1 2; do
for iss in 1 2; do
  for w in "C2 1000" "C1 1000" "C2 200"; do
    set -- $w
    OTREF_ISSUERS=$iss timeout 600 python bench.py --path pipeline --workload $1 --regions $2 --steps 6 --no-cpu-baseline > gpurun_out/p.json 2> gpurun_out/pipe.err
    python - <<PY
import json
try:
    b=json.loads(open("gpurun_out/p.json").read().strip().splitlines()[-1])
    print("rep $rep issuers $iss $1 $2: %.0f regions/s  %.1f ms" % (b["value"], b["ms_per_step"]))
except Exception as e: print("issuers $iss $1 $2 failed", e)
PY
  done
done
done
