#!/bin/bash
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
timeout 200 python scripts/run_one.py --batch 4096 --mib 1024 --reps 3 | tail -1
timeout 200 python scripts/run_one.py --data random --batch 4096 --mib 1024 --reps 3 | tail -1
timeout 200 python scripts/run_one.py --data alice --batch 2048 --mib 1024 --reps 3 | tail -1
