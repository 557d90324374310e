#!/bin/bash
for v in 40 104 168 232; do echo "== var $v"; SNAPPY_B200_LANE_VAR=$v timeout 200 python scripts/run_one.py --data random --mib 256 --reps 4 --mode exact_lane | tail -1; done
B=scripts/probes/randlat2.bin
$B 2 0 0 4096 3 266 524288; $B 2 0 0 4096 3 2000 524288; $B 2 0 0 4096 8 266
