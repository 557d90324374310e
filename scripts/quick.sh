#!/bin/bash
# quick GPU check: parity tests + a few throughput points
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for m in exact exact_pipe_smem exact_window; do timeout 120 python scripts/run_one.py --data mtg --mib 18.5 --reps 2 --mode $m --check | tail -3; done
for d in alice random zeros; do timeout 120 python scripts/run_one.py --data $d --mib 18.5 --reps 2 | tail -1; done
timeout 120 python scripts/run_one.py --batch 4096 --mib 256 --reps 2 | tail -1
timeout 200 python scripts/run_one.py --mib 1024 --reps 2 | tail -1
