#!/bin/bash
# warp instructions + time of k_align<1024> on C1 for each lib.  inst.sh <lib|-> ...
for lib in "$@"; do
  if [ "$lib" != "-" ]; then export OTTER_B200_LIB=$PWD/$lib; else unset OTTER_B200_LIB; fi
  ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_align -s 4 -c 1 --csv --log-file gpurun_out/inst.csv python bench.py --workload ${W:-C1} --steps 1 --warmup 3 --no-cpu-baseline ${R:+--regions $R} > /dev/null 2>&1
  echo "$lib: $(grep -v '^==' gpurun_out/inst.csv | awk -F'","' 'NR>1{printf "%s=%s  ", $(NF-2), $NF}')"
done
