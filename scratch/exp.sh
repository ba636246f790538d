#!/bin/bash
# one kernel experiment on the GPU box: parity, timing (200 regions of C2), instruction count
#   exp.sh <tag> [lib.so] [notest]
tag=${1:-exp}
if [ -n "$2" ] && [ "$2" != "-" ]; then export OTTER_B200_LIB=$PWD/$2; fi
if [ "$3" != "notest" ]; then
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${tag}_tests.log
tail -2 gpurun_out/${tag}_tests.log
fi
python bench.py --steps 3 --warmup 3 --regions 200 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err && \
ncu --metrics smsp__inst_executed.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_lsu.sum,smsp__inst_executed_pipe_xu.sum,gpu__time_duration.sum,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_align -s 6 -c 1 --csv --log-file gpurun_out/${tag}_inst.csv python bench.py --steps 1 --warmup 3 --regions 200 --no-cpu-baseline > gpurun_out/${tag}_ncu.log 2>&1
python - <<PY
import json,csv
b=json.load(open("gpurun_out/${tag}_bench.json"))
print("${tag}", "align ms %.2f"%b["stage_ms_per_step"]["align"], "wide", b["pairs_wide"], "det", b["deterministic"])
try:
    for r in csv.reader(open("gpurun_out/${tag}_inst.csv")):
        if len(r)>5 and r[0].isdigit(): print("  ", r[-3], r[-1])
except Exception as e: print("ncu", e)
PY
