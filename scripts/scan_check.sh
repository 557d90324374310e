#!/bin/bash
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_lane_gpu.py tests/test_fast_mode.py -x -q 2>&1 | tail -3
timeout 200 python scripts/run_one.py --data random --batch 4096 --mib 1024 --reps 3 | tail -1
timeout 200 python scripts/run_one.py --data random --mib 256 --reps 3 --mode exact_pipe | tail -1
timeout 200 python scripts/run_one.py --batch 4096 --mib 1024 --reps 3 | tail -1
timeout 200 python scripts/run_one.py --mib 512 --reps 3 --mode exact_pipe | tail -1
timeout 200 python scripts/run_one.py --data alice --mib 512 --reps 3 --mode exact_pipe | tail -1
