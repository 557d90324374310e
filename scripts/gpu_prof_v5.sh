#!/bin/bash
# Round-1 v5 evidence: the bench line with the configuration rows, its launch list, and a full
# capture of the block decode kernel (one 1 GiB raw stream, decode_par.cuh) from the same command.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err
tail -c 2500 gpurun_out/bench_v5.json; tail -3 gpurun_out/bench_v5.err
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_bench.log 2>&1
wc -l gpurun_out/launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:^decode_kernel -c 1 -f -o gpurun_out/prof_decode_blocks $B > gpurun_out/ncu_decode_blocks.log 2>&1
ls -la gpurun_out/*.ncu-rep
