#!/bin/bash
# A/B two libs: bench C2-200 + C1 align ms and instruction counts.  ab.sh <libA|-> <libB>
for lib in "$@"; do
  if [ "$lib" != "-" ]; then export OTTER_B200_LIB=$PWD/$lib; else unset OTTER_B200_LIB; fi
  for w in C2 C1; do
    python bench.py --workload $w --steps 3 --warmup 3 --regions 200 --no-cpu-baseline > gpurun_out/ab.json 2>/dev/null
    python - <<PY
import json
b=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
print("$lib", "$w", "align ms %.3f" % b["stage_ms_per_step"]["align"])
PY
  done
done
