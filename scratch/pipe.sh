#!/bin/bash
# pipeline bench, one and two device-stage issuers, C2 and C1 (200 and 1000 regions)
timeout 300 python -m pytest tests/test_gpu_affine.py tests/test_gpu_kde.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for iss in 1 2; do
  for w in "C2 200" "C2 1000" "C1 200" "C1 1000"; do
    set -- $w
    OTREF_ISSUERS=$iss timeout 600 python bench.py --path pipeline --workload $1 --regions $2 > gpurun_out/pipe_${1}_${2}_i$iss.json 2> gpurun_out/pipe.err
    python - <<PY
import json
try:
    b=json.loads(open("gpurun_out/pipe_${1}_${2}_i$iss.json").read().strip().splitlines()[-1])
    print("issuers $iss $1 $2: %.0f regions/s  %.1f ms  identical=%s  ref %.1f" % (b["value"], b["ms_per_step"], b.get("identical_to_reference"), b.get("cpu_baseline",{}).get("value",0)))
except Exception as e: print("issuers $iss $1 $2 failed", e)
PY
  done
done
