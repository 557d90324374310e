#!/bin/bash
timeout 900 python -m pytest tests/test_lane_gpu.py -x -q 2>&1 | tail -2
timeout 200 python scripts/run_one.py --mib 1024 --reps 4 --mode exact_lane | tail -2
timeout 200 python scripts/run_one.py --data random --mib 1024 --reps 3 --mode exact_lane | tail -1
timeout 200 python scripts/run_one.py --mib 2048 --reps 3 --mode exact_lane | tail -1
