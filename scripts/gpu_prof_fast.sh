#!/bin/bash
# ncu evidence for the fast mode: launch list of one 256 MiB call + full capture of parse_fast_kernel.
mkdir -p gpurun_out
R="python scripts/run_one.py --mib 256 --reps 2 --mode fast"
$R > gpurun_out/plain_fast.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_fast.csv $R > gpurun_out/ncu_fast_l.log 2>&1
S="python scripts/run_one.py --mib 37 --reps 2 --mode fast"
$S > gpurun_out/plain_fast2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:parse_fast -s 1 -c 1 -f -o gpurun_out/prof_fast $S > gpurun_out/ncu_fast_f.log 2>&1
tail -2 gpurun_out/plain_fast.log; grep -E "parse_fast|size_sparse|emit_sparse|scan_sizes" gpurun_out/launches_fast.csv | awk -F'","' '{print $5, $(NF-1), $NF}' | tail -8
