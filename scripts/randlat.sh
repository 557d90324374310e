#!/bin/bash
B=scripts/probes/randlat.bin
for mode in 0 1 2 3 4; do for jobs in 1024 4096 16384 32768; do $B $mode $jobs 8 2000; done; done
echo "--- fpw"; for fpw in 1 4 32; do for mode in 1 2; do $B $mode 16384 $fpw 2000; done; done
echo "--- region"; for region in 8192 32768 65536 1048576; do for mode in 1 2; do $B $mode 16384 8 2000 $region; done; done
