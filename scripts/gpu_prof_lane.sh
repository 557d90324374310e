#!/bin/bash
# ncu --set full of the lane parser (prefetch variants 1 and 0) + the L2 fetch-granularity knob.
mkdir -p gpurun_out
for g in 32 64 128; do echo "== L2 fetch $g"; SNAPPY_B200_L2_FETCH=$g timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1; done 2>&1 | tee gpurun_out/lane_l2fetch.log
python scripts/run_one.py --mib 256 --reps 2 --mode exact_lane > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:parse_lane -s 1 -c 1 -o gpurun_out/r02_lane_pf1 \
    python scripts/run_one.py --mib 256 --reps 2 --mode exact_lane > gpurun_out/ncu1.log 2>&1
SNAPPY_B200_LANE_PF=0 python scripts/run_one.py --mib 64 --reps 2 --mode exact_lane > gpurun_out/plain0.log 2>&1 &&
SNAPPY_B200_LANE_PF=0 ncu --set full --clock-control none --import-source on -k regex:parse_lane -s 1 -c 1 -o gpurun_out/r02_lane_pf0 \
    python scripts/run_one.py --mib 64 --reps 2 --mode exact_lane > gpurun_out/ncu0.log 2>&1
tail -2 gpurun_out/plain.log gpurun_out/plain0.log gpurun_out/ncu1.log gpurun_out/ncu0.log
ls -la gpurun_out
