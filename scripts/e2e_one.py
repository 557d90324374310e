#!/usr/bin/env python
"""End-to-end (host pointers, pinned) throughput of snappy_b200_compress: scripts/e2e_one.py [--mib N] [--mode exact|fast] [--reps K]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import __graft_entry__ as g
import corpus

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=1024)
ap.add_argument("--mode", default="exact")
ap.add_argument("--reps", type=int, default=4)
a = ap.parse_args()
pkg = g.load_package()
ctx = pkg.Context(0)
n = a.mib << 20
host = torch.from_numpy(corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n).copy()).pin_memory()
cap = pkg.max_compressed_length(n)
out = torch.empty(cap, dtype=torch.uint8).pin_memory()
for _ in range(2):
    got = ctx.compress_into(host.data_ptr(), n, out.data_ptr(), cap, mode=a.mode)
t0 = time.perf_counter()
for _ in range(a.reps):
    got = ctx.compress_into(host.data_ptr(), n, out.data_ptr(), cap, mode=a.mode)
dt = (time.perf_counter() - t0) / a.reps
print("e2e %s %d MiB chunk=%s: %.2f ms  %.2f GB/s  (out %d)" % (a.mode, a.mib, os.environ.get("SNAPPY_B200_CHUNK_MIB", "default"), dt * 1e3, n / dt / 1e9, got))
