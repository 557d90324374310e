#!/bin/bash
B=scripts/probes/randlat2.bin
echo "--- loads only, more jobs"; for j in 16384 65536 131072; do $B 4 0 0 $j; done
echo "--- exch+full-sector store b / exch + 16B store / exch+exch"; for j in 4096 16384 65536; do $B 7 0 0 $j; $B 3 0 1 $j; $B 9 0 0 $j; done
echo "--- 64-bit exch"; for j in 4096 16384 65536; do $B 8 0 0 $j; done
echo "--- exch, L2-resident regions (4 KB, 8 KB, 32 KB x jobs)"; for r in 4096 8192 32768; do for j in 2048 8192 16384; do $B 2 0 0 $j 8 2000 $r; done; done
echo "--- load a+store a delayed, L2-resident"; for r in 4096 32768; do for j in 2048 16384; do $B 1 0 0 $j 8 2000 $r; done; done
echo "--- loads only L2-resident"; for r in 4096 32768; do for j in 2048 16384; do $B 4 0 0 $j 8 2000 $r; done; done
echo "--- delayed store more jobs"; for j in 4096 65536; do $B 1 0 0 $j; done
