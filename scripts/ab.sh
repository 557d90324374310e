#!/bin/bash
# A/B several builds of the library: AB_MODE=exact|fast AB_MIB=1024 scripts/ab.sh var_a.so var_b.so ...
for so in "$@"; do
  echo "== $so"
  SNAPPY_B200_LIB=$PWD/compression-accelerator_b200/$so timeout 300 python scripts/run_one.py --mib ${AB_MIB:-1024} --reps 2 --mode ${AB_MODE:-exact} ${AB_CHECK:---check} 2>&1 | tail -${AB_TAIL:-3}
done
