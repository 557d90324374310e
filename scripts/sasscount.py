#!/usr/bin/env python
"""Static SASS count per source region of a kernel: scripts/sasscount.py <lib.so> <kernel-substring> <file> <lo> <hi>"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import srcprof
srcprof.SO = os.path.abspath(sys.argv[1])
sl = srcprof.sass_lines(sys.argv[2])
f, lo, hi = sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
n = sum(1 for (loc, _) in sl if loc[0] == f and lo <= loc[1] <= hi)
print("%s: %d SASS total, %d in %s:%d-%d" % (os.path.basename(sys.argv[1]), len(sl), n, f, lo, hi))
if len(sys.argv) > 6:
    for (loc, ins) in sl:
        if loc[0] == f and lo <= loc[1] <= hi or sys.argv[6] == "all":
            print("  %s:%d  %s" % (loc[0][:16], loc[1], ins))
