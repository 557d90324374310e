#!/bin/bash
# Round-1 v4 evidence: launch list of the bench command, full captures (with source) of the dominant
# kernel of each mode at the bench size, and of the decompressor / CRC kernels.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_bench.log 2>&1
for m in exact fast; do
  R="python scripts/run_one.py --mib 1024 --reps 2 --mode $m"
  k=parse_pipe; [ $m = fast ] && k=parse_fast
  $R > gpurun_out/plain_1g_$m.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o gpurun_out/prof_1g_$m $R > gpurun_out/ncu_1g_$m.log 2>&1
  tail -1 gpurun_out/plain_1g_$m.log
done
R="python scripts/run_one.py --batch 4096 --mib 1024 --reps 2 --decode"
$R > gpurun_out/plain_decode.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:decode_staged -s 1 -c 1 -f -o gpurun_out/prof_decode_pages $R > gpurun_out/ncu_decode.log 2>&1
R="python scripts/frame_one.py --mib 1024 --reps 2"
$R > gpurun_out/plain_frame.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:crc32c -s 1 -c 1 -f -o gpurun_out/prof_crc32c $R > gpurun_out/ncu_frame.log 2>&1
tail -3 gpurun_out/plain_frame.log; tail -1 gpurun_out/plain_decode.log; wc -l gpurun_out/launches.csv; ls -la gpurun_out/*.ncu-rep
