#!/usr/bin/env python
"""Write the size sweep of benchmark/hw_results.csv for this library: scripts/size_sweep_csv.py [out.csv]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g
import corpus

pkg = g.load_package()
from compression_accelerator_b200 import sweep_csv
dev = pkg.RoccDevice()
kind = {"random": "random", "real": "real", "repeat": "repeating"}
rows = sweep_csv.sweep_rows(sweep_csv.device_runner(dev), lambda t, n: corpus.benchmark_file(kind[t], n))
dev.close()
text = sweep_csv.to_csv(rows)
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "b200_results.csv")
os.makedirs(os.path.dirname(out), exist_ok=True)
open(out, "w").write(text)
print(text)
