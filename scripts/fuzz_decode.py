#!/usr/bin/env python
"""Differential run of the device decompressor: valid streams (reference Snappy, our exact and fast modes) and
mutated ones (flipped bytes, truncations, appended bytes) through snappy_b200_uncompress; the verdict (accepted /
refused) and, when accepted, every byte must equal the reference decompressor's (oracle/_ref when present, else the
port).  Streams of 128 KiB and more take the block-parallel decoder and its fallbacks (decode_par.cuh).

    python scripts/fuzz_decode.py [--seconds 90] [--seed 1]"""
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
    ap.add_argument("--verbose", action="store_true", help="print every step before it runs (to localise a fault)")
    a = ap.parse_args()
    import numpy as np
    import corpus
    import oracle_lib
    import fuzz_parity
    import __graft_entry__ as g
    pkg = g.load_package()
    ctx = pkg.Context(0)
    chk = oracle_lib.checker()
    port = oracle_lib.port()
    text = corpus.fixture("all-mtg-cards.txt")
    rng = np.random.default_rng(a.seed)
    t0 = time.time()
    runs = fails = refused = 0
    while time.time() - t0 < a.seconds:
        n = int(rng.choice([int(rng.integers(0, 300)), int(rng.integers(300, 70000)), int(rng.integers(70000, 600000))]))
        data = fuzz_parity.make_input(rng, n, corpus, text) if n else b""
        which = int(rng.integers(0, 3))
        if a.verbose:
            print("run %d: compress n=%d which=%d" % (runs, n, which), flush=True)
        comp = chk.compress(data) if which == 0 else ctx.compress(data, mode="exact" if which == 1 else "fast")
        buf = bytearray(comp)
        how = int(rng.integers(0, 5))
        if how == 1 and buf:                                       # flip a few bytes
            for _ in range(int(rng.integers(1, 4))):
                buf[int(rng.integers(0, len(buf)))] = int(rng.integers(0, 256))
        elif how == 2 and buf:                                     # truncate
            del buf[int(rng.integers(0, len(buf))):]
        elif how == 3:                                             # append
            buf += bytes(rng.integers(0, 256, size=int(rng.integers(1, 9)), dtype=np.uint8))
        elif how == 4 and len(buf) > 8:                            # flip one bit in an element tag region
            at = int(rng.integers(0, len(buf)))
            buf[at] ^= 1 << int(rng.integers(0, 8))
        blob = bytes(buf)
        if a.verbose:
            print("run %d: uncompress how=%d stream=%d B" % (runs, how, len(blob)), flush=True)
        try:                                                       # what the reference says
            if chk.uncompressed_length(blob) > (64 << 20):
                continue                                           # a mutated preamble asking for gigabytes: skipped
            want = chk.uncompress(blob)
        except ValueError:
            want = None
        try:
            got = ctx.uncompress(blob)
        except pkg.SnappyB200Error as e:
            got = None
            if e.status not in (1, 2):                             # INVALID_INPUT / BUFFER_TOO_SMALL are verdicts
                print("ERROR status=%d run=%d" % (e.status, runs), flush=True)
                fails += 1
        runs += 1
        refused += got is None
        if got != want:
            fails += 1
            print("MISMATCH n=%d which=%d how=%d stream=%d B: reference %s, device %s (seed %d run %d)" % (
                n, which, how, len(blob), "refuses" if want is None else "%d B" % len(want),
                "refuses" if got is None else "%d B" % len(got), a.seed, runs), flush=True)
    print("fuzz-decode: %d runs (%d refused), %d mismatches, %.0f s, seed %d" % (runs, refused, fails, time.time() - t0, a.seed))
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
