#!/usr/bin/env python
"""Differential run of the chunk-pipelined host entry (snappy_b200_compress, inputs of 24 MiB and more: H2D, parse,
emit and D2H of neighbouring chunks overlap; 112 MiB and more go to the global-table parser): random shapes and
sizes, exact mode against the reference Snappy and fast mode against its model, whole streams compared.

    python scripts/fuzz_host.py [--seconds 90] [--seed 1]"""
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
    runs = fails = 0
    while time.time() - t0 < a.seconds:
        n = int(rng.choice([int(rng.integers(24 << 20, 40 << 20)), int(rng.integers(40 << 20, 112 << 20)),
                            int(rng.integers(112 << 20, 160 << 20))])) + int(rng.integers(0, 65536))
        data = fuzz_parity.make_input(rng, n, corpus, text)
        mode = "exact" if rng.integers(0, 3) else "fast"
        want = chk.compress(data) if mode == "exact" else port.fast_compress(data)
        got = ctx.compress(data, mode=mode)
        runs += 1
        if got != want:
            fails += 1
            print("MISMATCH host entry mode=%s n=%d: %d vs %d bytes (seed %d run %d)" % (mode, n, len(got), len(want), a.seed, runs), flush=True)
    print("fuzz-host: %d runs, %d mismatches, %.0f s, seed %d" % (runs, fails, time.time() - t0, a.seed))
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
