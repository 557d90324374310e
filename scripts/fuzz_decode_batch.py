#!/usr/bin/env python
"""Differential run of the batch decompressor (snappy_b200_uncompress_batch_device: page-shaped batches take the
kernel that stages stream and page in shared memory, larger ones the global-memory kernel): `count` streams of
random shapes, a third of them mutated (flipped bytes, truncated, extended), every verdict and every accepted
byte against the reference decompressor.

    python scripts/fuzz_decode_batch.py [--seconds 90] [--seed 1]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=90.0)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    import numpy as np
    import torch
    import corpus
    import oracle_lib
    import fuzz_parity
    import __graft_entry__ as g
    pkg = g.load_package()
    ctx = pkg.Context(0)
    chk = oracle_lib.checker()
    text = corpus.fixture("all-mtg-cards.txt")
    rng = np.random.default_rng(a.seed)
    t0 = time.time()
    runs = fails = streams = refused = 0
    while time.time() - t0 < a.seconds:
        length = int(rng.choice([int(rng.integers(1, 300)), int(rng.integers(300, 4097)), 4096, int(rng.integers(4097, 8193)),
                                 int(rng.integers(8193, 70000))]))
        count = int(rng.choice([1, 5, 33, 150]))
        out_cap = length + int(rng.integers(0, 3)) * 8                 # capacity per stream (>= the honest length)
        blob = fuzz_parity.make_input(rng, length * count, corpus, text)
        comps, wants = [], []
        for i in range(count):
            one = blob[i * length:(i + 1) * length]
            comp = bytearray(chk.compress(one))
            how = int(rng.integers(0, 9))
            if how == 1:
                for _ in range(int(rng.integers(1, 4))):
                    comp[int(rng.integers(0, len(comp)))] = int(rng.integers(0, 256))
            elif how == 2:
                del comp[int(rng.integers(0, len(comp))):]
            elif how == 3:
                comp += bytes(rng.integers(0, 256, size=int(rng.integers(1, 6)), dtype=np.uint8))
            comp = bytes(comp)
            try:
                n = chk.uncompressed_length(comp)
                want = chk.uncompress(comp) if n <= out_cap else None   # more than the slot holds: refused
            except ValueError:
                want = None
            comps.append(comp)
            wants.append(want)
        in_stride = max(len(c) for c in comps) + int(rng.integers(0, 5))
        in_stride = max(in_stride, 1)
        buf = np.zeros(in_stride * count + 64, dtype=np.uint8)
        for i, c in enumerate(comps):
            buf[i * in_stride:i * in_stride + len(c)] = np.frombuffer(c, dtype=np.uint8)
        src = torch.from_numpy(buf).cuda()
        sizes = torch.tensor([len(c) for c in comps], dtype=torch.int32).cuda()
        dst, lens = ctx.uncompress_batch_device(src, in_stride, sizes, count, out_cap)
        ctx.sync()
        out = dst.cpu().numpy()
        ln = lens.cpu().numpy().astype(np.int64)
        runs += 1
        for i in range(count):
            streams += 1
            got = None if ln[i] < 0 else bytes(out[i * out_cap:i * out_cap + ln[i]])
            refused += got is None
            if got != wants[i]:
                fails += 1
                print("MISMATCH length=%d count=%d i=%d cap=%d stream=%d B: reference %s, device %s (seed %d run %d)" % (
                    length, count, i, out_cap, len(comps[i]), "refuses" if wants[i] is None else "%d B" % len(wants[i]),
                    "refuses" if got is None else "%d B" % len(got), a.seed, runs), flush=True)
                break
    print("fuzz-decode-batch: %d batches, %d streams (%d refused), %d mismatches, %.0f s, seed %d" % (
        runs, streams, refused, fails, time.time() - t0, a.seed))
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
