#!/bin/bash
for cfg in "3 2 40" "3 1 40" "3 4 40" "3 2 32" "5 2 40" "6 2 40"; do set -- $cfg
  echo "== fpw=$1 wpc=$2 var=$3"; SNAPPY_B200_LANE_FPW=$1 SNAPPY_B200_LANE_WPC=$2 SNAPPY_B200_LANE_VAR=$3 timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1
done
