#!/bin/bash
# A/B libs on the consensus / compare / pipeline benches.  ab_aff.sh <lib|-> ...
for lib in "$@"; do
  if [ "$lib" != "-" ]; then export OTTER_B200_LIB=$PWD/$lib; else unset OTTER_B200_LIB; fi
  for p in consensus compare; do
    timeout 300 python bench.py --path $p --no-cpu-baseline > gpurun_out/ab.json 2>gpurun_out/ab.err
    python - <<PY
import json
try:
    b=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
    print("$lib", "$p", "%.0f" % b["value"], b["unit"], "ms/step %.2f" % b["ms_per_step"], "retried", b.get("retried"))
except Exception as e: print("$lib $p failed", e)
PY
  done
done
