#!/usr/bin/env python
"""Repro driver: fast mode on purely periodic input (found by scripts/fuzz_decode.py, seed 5 run 721)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ONE = r'''
import sys, os
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
import numpy as np
import oracle_lib
import __graft_entry__ as g
pkg = g.load_package(); ctx = pkg.Context(0)
per, n, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
base = bytes(np.random.default_rng(seed).integers(0, 256, size=per, dtype=np.uint8))
data = (base * (n // per + 2))[:n]
got = ctx.compress(data, mode="fast")
want = oracle_lib.port().fast_compress(data)
print("ok" if got == want else "DIFFERS", len(got), len(want))
''' % (ROOT, ROOT)
for per in (1, 60, 64, 100, 128, 129, 256, 300):
    for n in (65536, 100000, 190589):
        r = subprocess.run([sys.executable, "-c", ONE, str(per), str(n), "1"], capture_output=True, text=True, timeout=300)
        tail = (r.stdout.strip().splitlines() or [""])[-1] if r.returncode == 0 else "FAULT: " + r.stderr.strip().splitlines()[-1][-120:]
        print("period %4d  n %7d : %s" % (per, n, tail), flush=True)
