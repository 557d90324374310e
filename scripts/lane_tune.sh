#!/bin/bash
mkdir -p gpurun_out
B=scripts/probes/randlat2.bin
echo "--- exch+full-sector store b / exch + 16B store / exch+exch / 64-bit exch"; for j in 4096 16384 65536; do $B 7 0 0 $j; $B 3 0 1 $j; $B 9 0 0 $j; $B 8 0 0 $j; done
timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -2
for v in 34 40 42; do SNAPPY_B200_LANE_VAR=$v timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -1; done
for v in 32 33 34 40 41 42 1; do for fpw in 8 4; do
  echo "== var=$v fpw=$fpw"; SNAPPY_B200_LANE_VAR=$v SNAPPY_B200_LANE_FPW=$fpw timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1
done; done 2>&1 | tee gpurun_out/lane_tune.log
for v in 32 34 42; do for mib in 64 256 2048; do echo "== var=$v lane $mib MiB"; SNAPPY_B200_LANE_VAR=$v timeout 200 python scripts/run_one.py --mib $mib --reps 3 --mode exact_lane | tail -1; done; done 2>&1 | tee -a gpurun_out/lane_tune.log
