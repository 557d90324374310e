#!/bin/bash
for cfg in "32 8" "32 12" "32 16" "16 16" "24 12" "48 8" "8 16"; do set -- $cfg
  echo "== chunk=$1 MiB streams=$2"; SNAPPY_B200_CHUNK_MIB=$1 SNAPPY_B200_PARSE_STREAMS=$2 timeout 300 python scripts/e2e_one.py --mib 1024 --reps 4 | tail -1; done
echo "== prio off"; SNAPPY_B200_STREAM_PRIO=0 timeout 300 python scripts/e2e_one.py --mib 1024 --reps 4 | tail -1
