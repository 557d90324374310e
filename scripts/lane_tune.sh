#!/bin/bash
# GPU visit: lane-parser parity + a sweep of its knobs on the 1 GiB headline (device-resident).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv | tail -1
timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -3
for v in 10 14; do SNAPPY_B200_LANE_VAR=$v timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -1; done
for v in 2 6 10 14 18 1 0 4 8 12 9; do
  echo "== var=$v fpw=8"; SNAPPY_B200_LANE_VAR=$v timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1
done 2>&1 | tee gpurun_out/lane_tune.log
for v in 2 10 14; do for fpw in 4 16 32; do
  echo "== var=$v fpw=$fpw"; SNAPPY_B200_LANE_VAR=$v SNAPPY_B200_LANE_FPW=$fpw timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1
done; done 2>&1 | tee -a gpurun_out/lane_tune.log
for v in 2 10; do for mib in 64 256 2048; do echo "== var=$v lane $mib MiB"; SNAPPY_B200_LANE_VAR=$v timeout 200 python scripts/run_one.py --mib $mib --reps 3 --mode exact_lane | tail -1; done; done 2>&1 | tee -a gpurun_out/lane_tune.log
