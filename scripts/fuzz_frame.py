#!/usr/bin/env python
"""Differential run of the framing format (tests/snappy/framing_format.txt): the device writer against the CPU
restatement byte for byte, and the device reader on valid, re-chunked and mutated framed streams (flipped bytes,
truncations, inserted padding / skippable / reserved chunks, uncompressed chunks, repeated identifiers) against the
restatement's verdict and bytes (tests/oracle_lib.py frame_compress / frame_uncompress).

    python scripts/fuzz_frame.py [--seconds 90] [--seed 1]"""
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
    runs = fails = refused = 0

    def chunk(typ, body):
        return bytes([typ, len(body) & 0xff, (len(body) >> 8) & 0xff, (len(body) >> 16) & 0xff]) + body

    while time.time() - t0 < a.seconds:
        n = int(rng.choice([int(rng.integers(0, 300)), int(rng.integers(300, 70000)), int(rng.integers(70000, 500000))]))
        data = fuzz_parity.make_input(rng, n, corpus, text) if n else b""
        want_framed = oracle_lib.frame_compress(data)
        got_framed = ctx.frame_compress(data)
        runs += 1
        if got_framed != want_framed:
            fails += 1
            print("MISMATCH writer n=%d (seed %d run %d)" % (n, a.seed, runs), flush=True)
            continue
        # a stream for the reader: ours, or hand-built from chunks of random sizes and kinds
        how = int(rng.integers(0, 6))
        if how == 0:
            blob = want_framed
        elif how in (1, 2):
            out = bytearray(oracle_lib.STREAM_ID)
            pos = 0
            while pos < len(data):
                k = int(rng.integers(1, 65537))
                part = data[pos:pos + k]
                crc = port.crc32c_masked(part).to_bytes(4, "little")
                kind = int(rng.integers(0, 6))
                if kind == 0:
                    out += chunk(0x01, crc + part)                      # uncompressed data
                else:
                    out += chunk(0x00, crc + chk.compress(part))
                if kind == 2:
                    out += chunk(0xfe, bytes(int(rng.integers(0, 40))))  # padding
                if kind == 3:
                    out += chunk(int(rng.integers(0x80, 0xfe)), bytes(rng.integers(0, 256, size=int(rng.integers(0, 30)), dtype=np.uint8)))
                if kind == 4:
                    out += oracle_lib.STREAM_ID                         # a repeated identifier
                if how == 2 and kind == 5:
                    out += chunk(int(rng.integers(0x02, 0x80)), b"xyz")  # reserved unskippable: must be refused
                pos += k
            blob = bytes(out)
        else:
            buf = bytearray(want_framed)
            if how == 3 and buf:
                for _ in range(int(rng.integers(1, 4))):
                    buf[int(rng.integers(0, len(buf)))] = int(rng.integers(0, 256))
            elif how == 4 and buf:
                del buf[int(rng.integers(0, len(buf))):]
            else:
                buf += bytes(rng.integers(0, 256, size=int(rng.integers(1, 12)), dtype=np.uint8))
            blob = bytes(buf)
        try:
            want = oracle_lib.frame_uncompress(blob)
        except ValueError:
            want = None
        try:
            got = ctx.frame_uncompress(blob)
        except pkg.SnappyB200Error as e:
            got = None
            if e.status not in (1, 2):
                print("ERROR status=%d run=%d" % (e.status, runs), flush=True)
                fails += 1
        refused += got is None
        if got != want:
            fails += 1
            print("MISMATCH reader n=%d how=%d stream=%d B: restatement %s, device %s (seed %d run %d)" % (
                n, how, len(blob), "refuses" if want is None else "%d B" % len(want),
                "refuses" if got is None else "%d B" % len(got), a.seed, runs), flush=True)
    print("fuzz-frame: %d runs (%d refused), %d mismatches, %.0f s, seed %d" % (runs, refused, fails, time.time() - t0, a.seed))
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
