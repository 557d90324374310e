#!/bin/bash
# dram traffic of the dominant kernels at the bench size (1 GiB): one ncu --set full capture each.
mkdir -p gpurun_out
for m in exact fast; do
  R="python scripts/run_one.py --mib 1024 --reps 2 --mode $m"
  k=parse_pipe; [ $m = fast ] && k=parse_fast
  $R > gpurun_out/plain_1g_$m.log 2>&1 &&
  ncu --set full --clock-control none -k regex:$k -s 1 -c 1 -f -o gpurun_out/prof_1g_$m $R > gpurun_out/ncu_1g_$m.log 2>&1
  tail -1 gpurun_out/plain_1g_$m.log
done
