#!/bin/bash
# ncu --set full of (a) the page parser with its own emission on 1,000-byte text buffers and (b) the global-table
# parser as it is now (sector-granular candidate loads); each only after the same command has exited 0 without ncu.
mkdir -p gpurun_out
python scripts/run_one.py --mib 256 --batch 1000 --reps 2 > gpurun_out/plain_page.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:parse_kernel -s 1 -c 1 -o gpurun_out/r02_page_fused \
    python scripts/run_one.py --mib 256 --batch 1000 --reps 2 > gpurun_out/ncu_page.log 2>&1
python scripts/run_one.py --mib 512 --mode exact_gtab --reps 2 > gpurun_out/plain_gtab.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:parse_gtab -s 1 -c 1 -o gpurun_out/r02_gtab_v2 \
    python scripts/run_one.py --mib 512 --mode exact_gtab --reps 2 > gpurun_out/ncu_gtab.log 2>&1
tail -2 gpurun_out/plain_page.log gpurun_out/plain_gtab.log gpurun_out/ncu_page.log gpurun_out/ncu_gtab.log
ls -la gpurun_out/*.ncu-rep
