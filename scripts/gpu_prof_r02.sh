#!/bin/bash
# ncu evidence for round 2: launch list of the bench step, full capture of the lane parser at 1 GiB,
# DRAM bytes of every kernel of one step (exact and fast), the pipelined parser for comparison.
mkdir -p gpurun_out
CMD="python scripts/run_one.py --mib 1024 --reps 2 --mode exact"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv \
    --log-file gpurun_out/r02_step_launches_exact.csv $CMD > gpurun_out/ncu_l.log 2>&1
CMDF="python scripts/run_one.py --mib 1024 --reps 2 --mode fast"
$CMDF >> gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv \
    --log-file gpurun_out/r02_step_launches_fast.csv $CMDF >> gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:parse_lane -s 1 -c 1 -o gpurun_out/r02_lane_1g $CMD > gpurun_out/ncu_f.log 2>&1
cat gpurun_out/plain.log; tail -3 gpurun_out/ncu_l.log gpurun_out/ncu_f.log; ls -la gpurun_out
