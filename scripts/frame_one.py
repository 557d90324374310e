#!/usr/bin/env python
"""Device timings of the framing path: CRC-32C pass, framed compress, framed (parallel) uncompress.
scripts/frame_one.py [--mib N] [--mode exact|fast] [--reps K]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import __graft_entry__ as g
import corpus

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=1024)
ap.add_argument("--mode", default="exact")
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
pkg = g.load_package()
ctx = pkg.Context(0)
n = a.mib << 20
src = torch.from_numpy(corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)).cuda()


def timed(fn):
    fn(); ctx.sync()
    t0 = time.perf_counter()
    for _ in range(a.reps):
        fn()
    ctx.sync()
    return (time.perf_counter() - t0) / a.reps * 1e3


crcs = ctx.crc32c_chunks_device(src)
ms = timed(lambda: ctx.crc32c_chunks_device(src, out=crcs))
print("crc32c %d MiB in 64 KiB chunks: %.3f ms  %.1f GB/s" % (a.mib, ms, n / ms / 1e6))
dst, out_n = ctx.frame_compress_device(src, mode=a.mode)
ms = timed(lambda: ctx.frame_compress_device(src, dst=dst, out_n=out_n, mode=a.mode))
total = int(out_n.item())
print("frame compress %s: %.3f ms  %.2f GB/s in, out/in %.4f" % (a.mode, ms, n / ms / 1e6, total / n))
back = torch.empty(n, dtype=torch.uint8, device="cuda")
ms = timed(lambda: ctx.frame_uncompress_device(dst, total, back))
print("frame uncompress: %.3f ms  %.2f GB/s out (%.2f GB/s in+out); identical: %s" % (
    ms, n / ms / 1e6, (n + total) / ms / 1e6, bool(torch.equal(back, src))))
