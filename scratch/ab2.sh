#!/bin/bash
# A/B libs: align-stage ms on C1 (1,000 regions), C2 (300), C3 (60).  ab2.sh <lib|-> ...
for lib in "$@"; do
  if [ "$lib" != "-" ]; then export OTTER_B200_LIB=$PWD/$lib; else unset OTTER_B200_LIB; fi
  for w in "C1 1000" "C2 300" "C3 60"; do
    set -- $w
    python bench.py --workload $1 --steps 4 --warmup 3 --regions $2 --no-cpu-baseline > gpurun_out/ab.json 2>/dev/null
    python - <<PY
import json
b=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
print("$lib", "$1", "align ms %.3f" % b["stage_ms_per_step"]["align"], "wide %.3f" % b["stage_ms_per_step"]["wide"], "det", b["deterministic"])
PY
  done
done
