#!/bin/bash
# ncu evidence: launch list of the bench command + one full capture of the parse kernel.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3"
R="python scripts/run_one.py --mib 55.5 --reps 2"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_bench.log 2>&1
$R > gpurun_out/plain_run.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:parse_pipe -s 1 -c 1 -f -o gpurun_out/prof_parse $R > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/plain_run.log; tail -2 gpurun_out/ncu_run.log; wc -l gpurun_out/launches.csv
