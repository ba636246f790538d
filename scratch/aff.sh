#!/bin/bash
# gap-affine / edit-CIGAR tiers on the GPU box: parity tests, then the three benches that use them
tag=${1:-aff}
timeout 600 python -m pytest tests/test_gpu_affine.py tests/test_gpu_editcigar.py tests/test_gpu_kde.py -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; tail -3 gpurun_out/${tag}_tests.log
for p in consensus compare; do
  timeout 300 python bench.py --path $p --no-cpu-baseline > gpurun_out/${tag}_$p.json 2> gpurun_out/${tag}_$p.err
  python - <<PY
import json
try:
    b=json.loads(open("gpurun_out/${tag}_$p.json").read().strip().splitlines()[-1])
    print("$p", b["value"], b["unit"], "ms/step", b["ms_per_step"], {k:b[k] for k in b if k in ("stage_ms_per_step","pairs_wide","kernel_ms","e2e")})
except Exception as e: print("$p failed", e)
PY
done
