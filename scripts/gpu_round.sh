#!/bin/bash
# One GPU visit: parity tests, the bench line, the parser's per-window counters, a few throughput points.
mkdir -p gpurun_out
python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 600 python bench.py --steps 5 --warmup 3 2> gpurun_out/bench.err | grep "^{" > gpurun_out/bench.json; tail -c 3000 gpurun_out/bench.json
make -s -C compression-accelerator_b200/csrc stats > /dev/null 2>&1 &&
SNAPPY_B200_LIB=$PWD/compression-accelerator_b200/libsnappy_b200_stats.so timeout 300 python scripts/run_one.py --mib 1024 --reps 2 2>&1 | tail -5
for d in alice random zeros; do timeout 120 python scripts/run_one.py --data $d --mib 256 --reps 2 | tail -1; done
timeout 120 python scripts/run_one.py --batch 4096 --mib 1024 --reps 2 --decode | tail -2
