#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_all_engines.py -x -q 2>&1 | tail -4
for v in base r64; do
  if [ $v = base ]; then unset SNAPPY_B200_LIB; else export SNAPPY_B200_LIB=$PWD/build_ab/libsnappy_b200_$v.so; fi
  echo "== lib $v"
  for b in 300 600 1000 2000 4096; do
    for d in mtg random; do timeout 100 python scripts/run_one.py --mib 1024 --batch $b --reps 3 --data $d | tail -1; done
  done
done
