#!/usr/bin/env python
"""Differential run on the GPU: random inputs through every exact engine, the fast mode and the batch entry, every
stream compared byte for byte with the CPU checker (compiled reference Snappy when present, else the port; the fast
mode against its sequential model, oracle/snappy_port.c port_fast_compress).

    python scripts/fuzz_parity.py [--seconds 120] [--seed 1]

Inputs: structured random data (tests/corpus.py, after snappy_unittest.cc's RandomizedTest), long incompressible
runs (the skip schedule), text with planted far copies, and mixtures; lengths from 1 byte to a few fragments;
batches with ragged strides of 1 .. 20,000-byte buffers.  Prints one line per failure and a summary; exit 1 on
any mismatch."""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def make_input(rng, n, corpus, text):
    import numpy as np
    kind = int(rng.integers(0, 6))
    if n > (1 << 20) and kind in (0, 5):            # the structured generator is a Python loop per byte
        kind = int(rng.integers(1, 5))
    if kind == 0:
        return corpus.structured_random(n, int(rng.integers(0, 1 << 30)))
    if kind == 1:                                   # incompressible, small alphabet now and then
        alpha = int(rng.choice([2, 4, 16, 93, 256]))
        return bytes(rng.integers(0, alpha, size=n, dtype=np.uint8))
    if kind == 2:                                   # text from a random place
        at = int(rng.integers(0, len(text)))
        return (text * (n // len(text) + 2))[at:at + n]
    if kind == 3:                                   # long misses, then a hit, again and again
        out = bytearray()
        seed_bytes = bytes(rng.integers(0, 256, size=64, dtype=np.uint8))
        while len(out) < n:
            out += bytes(rng.integers(0, 256, size=int(rng.integers(1, 3000)), dtype=np.uint8))
            out += seed_bytes[:int(rng.integers(4, 64))]
        return bytes(out[:n])
    if kind == 4:                                   # periodic with a random period and noise
        per = int(rng.integers(1, 300)) if rng.integers(0, 3) else int(rng.choice([32, 64, 96, 128, 256, 512, 1024, 2048, 4096]))
        base = bytes(rng.integers(0, 256, size=per, dtype=np.uint8))
        arr = np.frombuffer((base * (n // per + 2))[:n], dtype=np.uint8).copy()
        for _ in range(int(rng.integers(0, 8))):
            arr[int(rng.integers(0, n))] ^= 0x55
        return arr.tobytes()
    a = corpus.structured_random(n // 2 + 1, int(rng.integers(0, 1 << 30)))
    return (a + text[:n])[:n]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=120.0)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    import numpy as np
    import torch
    import corpus
    import oracle_lib
    import __graft_entry__ as g
    pkg = g.load_package()
    ctx = pkg.Context(0)
    chk = oracle_lib.checker()
    text = corpus.fixture("all-mtg-cards.txt")
    rng = np.random.default_rng(a.seed)
    t0 = time.time()
    runs = fails = 0
    port = oracle_lib.port()
    modes = ["exact", "exact_lane", "exact_pipe", "exact_gtab", "fast", "fast"]      # fast: against its sequential model
    L = pkg.lib()
    na_dst = torch.zeros(pkg.max_compressed_length(1 << 20) + 64, dtype=torch.uint8, device="cuda")
    na_n = torch.zeros(1, dtype=torch.int64, device="cuda")
    while time.time() - t0 < a.seconds:
        if rng.integers(0, 8) == 0:                 # the no-allocation entry (inputs up to 1 MiB, its own scratch)
            n = int(rng.choice([int(rng.integers(0, 300)), int(rng.integers(300, 70000)), int(rng.integers(70000, (1 << 20) + 1))]))
            data = make_input(rng, n, corpus, text) if n else b""
            src = torch.frombuffer(bytearray(data) if n else bytearray(1), dtype=torch.uint8).cuda()
            torch.cuda.synchronize()
            st = L.snappy_b200_compress_device_noalloc(src.data_ptr(), n, na_dst.data_ptr(), na_dst.numel(), na_n.data_ptr(), None)
            torch.cuda.synchronize()
            runs += 1
            if st != 0 or bytes(na_dst[:int(na_n.item())].cpu().numpy()) != chk.compress(data):
                fails += 1
                print("MISMATCH noalloc n=%d status=%d seed=%d run=%d" % (n, st, a.seed, runs), flush=True)
            continue
        if rng.integers(0, 3) < 2:                  # one stream, one engine
            n = int(rng.choice([int(rng.integers(1, 300)), int(rng.integers(300, 70000)), int(rng.integers(70000, 400000))]))
            data = make_input(rng, n, corpus, text)
            mode = modes[int(rng.integers(0, len(modes)))]
            want = port.fast_compress(data) if mode == "fast" else chk.compress(data)
            src = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
            dst, out_n = ctx.compress_device(src, mode=mode)
            ctx.sync()
            got = bytes(dst[:int(out_n.item())].cpu().numpy())
            runs += 1
            if got != want:
                fails += 1
                print("MISMATCH stream mode=%s n=%d seed=%d run=%d" % (mode, n, a.seed, runs), flush=True)
        else:                                       # a batch with a ragged stride
            length = int(rng.choice([int(rng.integers(1, 300)), int(rng.integers(257, 2100)), int(rng.integers(2000, 9000)),
                                     int(rng.integers(4097, 20000))]))
            count = int(rng.choice([1, 3, 40, 200])) if length < 4097 or rng.integers(0, 2) else 1800
            stride = length + int(rng.integers(0, 9))
            blob = make_input(rng, stride * count, corpus, text)
            src = torch.frombuffer(bytearray(blob), dtype=torch.uint8).cuda()
            mode = ["exact", "exact", "exact_gtab", "fast"][int(rng.integers(0, 4))]
            dst, ostride, sizes = ctx.compress_batch_device(src, length, stride, count, mode=mode)
            ctx.sync()
            out = dst.cpu().numpy()
            sz = sizes.cpu().numpy()
            runs += 1
            for i in (range(count) if count <= 200 else range(0, count, 9)):
                one = blob[i * stride:i * stride + length]
                want = port.fast_compress(one) if mode == "fast" else chk.compress(one)
                if bytes(out[i * ostride:i * ostride + sz[i]]) != want:
                    fails += 1
                    print("MISMATCH batch mode=%s length=%d stride=%d count=%d i=%d seed=%d run=%d" % (
                        mode, length, stride, count, i, a.seed, runs), flush=True)
                    break
    print("fuzz: %d runs, %d mismatches, %.0f s, seed %d" % (runs, fails, time.time() - t0, a.seed))
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
