#!/bin/bash
# Round-1 v6 evidence: parity suite, the bench line, its launch list.
mkdir -p gpurun_out
python __graft_entry__.py smoke 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py --steps 5 --warmup 3 2> gpurun_out/bench_v6.err | grep "^{" > gpurun_out/bench_v6.json
head -c 400 gpurun_out/bench_v6.json; echo; tail -2 gpurun_out/bench_v6.err
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | grep "^{" > gpurun_out/bench_v6_ref.json; head -c 300 gpurun_out/bench_v6_ref.json; echo
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
wc -l gpurun_out/launches.csv
