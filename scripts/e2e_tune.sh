#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_lane_gpu.py tests/test_gpu_parity.py -x -q 2>&1 | tail -2
for pct in 0 20 33 45 60; do echo "== tail_pct=$pct"; SNAPPY_B200_E2E_LANE_PCT=$pct timeout 200 python scripts/e2e_one.py --mib 1024 --reps 4 | tail -1; done 2>&1 | tee gpurun_out/e2e_tune.log
for cfg in "33 1 4" "33 2 4" "33 4 2" "45 1 4" "25 1 4"; do set -- $cfg; echo "== tail_pct=$1 fpw=$2 wpc=$3"; SNAPPY_B200_E2E_LANE_PCT=$1 SNAPPY_B200_LANE_FPW=$2 SNAPPY_B200_LANE_WPC=$3 timeout 200 python scripts/e2e_one.py --mib 1024 --reps 4 | tail -1; done 2>&1 | tee -a gpurun_out/e2e_tune.log
for c in 8 32; do echo "== tail 33 chunk $c"; SNAPPY_B200_CHUNK_MIB=$c timeout 200 python scripts/e2e_one.py --mib 1024 --reps 4 | tail -1; done 2>&1 | tee -a gpurun_out/e2e_tune.log
echo "== fast"; timeout 200 python scripts/e2e_one.py --mib 1024 --reps 4 --mode fast | tail -1
echo "== 2 GiB"; timeout 200 python scripts/e2e_one.py --mib 2048 --reps 3 | tail -1
echo "== 512 MiB"; timeout 200 python scripts/e2e_one.py --mib 512 --reps 4 | tail -1
