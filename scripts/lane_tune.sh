#!/bin/bash
# GPU visit: lane-parser parity + a sweep of its two knobs on the 1 GiB headline (device-resident).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv | tail -1
timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -6
for pf in 1 0 3 5 11; do for fpw in 8 4 16 32; do
  echo "== pf=$pf fpw=$fpw"; SNAPPY_B200_LANE_PF=$pf SNAPPY_B200_LANE_FPW=$fpw timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1
done; done 2>&1 | tee gpurun_out/lane_tune.log
echo "== pipe"; timeout 200 python scripts/run_one.py --mib 1024 --reps 2 --mode exact_pipe | tail -1
for mib in 64 256 512 2048; do echo "== lane $mib MiB"; timeout 200 python scripts/run_one.py --mib $mib --reps 3 --mode exact_lane | tail -1; done 2>&1 | tee -a gpurun_out/lane_tune.log
for d in random alice zeros; do echo "== lane $d"; timeout 200 python scripts/run_one.py --data $d --mib 1024 --reps 3 --mode exact_lane | tail -1; done 2>&1 | tee -a gpurun_out/lane_tune.log
