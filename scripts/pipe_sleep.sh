#!/bin/bash
for s in 0 400 1000; do echo "== SB_PIPE_SLEEP=$s"; SNAPPY_B200_LIB=$PWD/compression-accelerator_b200/libsnappy_b200_sleep$s.so timeout 200 python scripts/run_one.py --mib 512 --reps 3 --mode exact_pipe | tail -1; done
echo "== default (100)"; timeout 200 python scripts/run_one.py --mib 512 --reps 3 --mode exact_pipe | tail -1
