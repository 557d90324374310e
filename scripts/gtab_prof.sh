#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/run_one.py --mib 512 --reps 2 --mode exact_gtab"
SNAPPY_B200_GTAB_W=28 $CMD > gpurun_out/plain.log 2>&1 &&
SNAPPY_B200_GTAB_W=28 ncu --set full --clock-control none --import-source on -k regex:parse_gtab -s 1 -c 1 -o gpurun_out/r02_gtab $CMD > gpurun_out/ncu.log 2>&1
cat gpurun_out/plain.log | tail -2; ls -la gpurun_out/*.ncu-rep
