#!/bin/bash
timeout 900 python -m pytest tests/test_fast_mode.py tests/test_uncompress.py -x -q -m gpu 2>&1 | tail -2
timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode fast --check | tail -3
