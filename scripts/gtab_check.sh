#!/bin/bash
for b in 1024 2048 4096 8192; do for mode in exact exact_page; do echo "== $b-byte text pages, $mode"; timeout 200 python scripts/run_one.py --batch $b --mib 1024 --reps 3 --mode $mode --decode | tail -2; done; done
