#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -2
echo "### hybrid: pipe share x shape, 1 GiB"
for cfg in "4 2" "2 4"; do set -- $cfg; for pct in 0 40 50 55 60 65 70; do
  echo "== fpw=$1 wpc=$2 pipe=$pct%"; SNAPPY_B200_LANE_FPW=$1 SNAPPY_B200_LANE_WPC=$2 SNAPPY_B200_HYBRID_PIPE_PCT=$pct timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact | tail -1
done; done 2>&1 | tee gpurun_out/lane_tune.log
echo "### hybrid 2 GiB / 512 MiB"
for mib in 2048 512; do for pct in 40 55 70; do echo "== $mib MiB pipe=$pct%"; SNAPPY_B200_HYBRID_PIPE_PCT=$pct timeout 200 python scripts/run_one.py --mib $mib --reps 3 --mode exact | tail -1; done; done 2>&1 | tee -a gpurun_out/lane_tune.log
SNAPPY_B200_HYBRID_PIPE_PCT=55 timeout 200 python scripts/run_one.py --mib 1024 --reps 2 --mode exact --check | tail -3
