"""Test inputs: committed fixtures (tests/golden/data, xz-compressed copies of the reference's
data files made by tests/golden/make_golden.py) and seeded synthetic generators.

Never reads /root/reference (it does not exist on the GPU box).
"""
import json
import lzma
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
DATA = os.path.join(GOLDEN, "data")

THISTLE = (b"If Tid the thistle sifter lifts his sieve of sifted thistles, then he'll shift the "
           b"thistles in his sifting sieve to live unsifted.")  # tests/compression.c:43, swcompression.cc:43

SWEEP_SIZES = (10, 20, 50, 100, 200, 500, 1000, 2000, 5000, 10000, 20000, 50000)  # DataGenerator.scala:27

_cache = {}


def fixture(name: str) -> bytes:
    """Bytes of a committed fixture, e.g. 'fox-in-socks.txt', 'all-mtg-cards.txt'."""
    if name not in _cache:
        with lzma.open(os.path.join(DATA, name + ".xz"), "rb") as f:
            _cache[name] = f.read()
    return _cache[name]


def manifest() -> dict:
    with open(os.path.join(GOLDEN, "manifest.json")) as f:
        return json.load(f)


def java_random_ascii(n: int = 100000, seed: int = 4444) -> bytes:
    """data/randomASCII.txt regenerated: scala.util.Random(4444).nextInt(93)+32
    (src/test/scala/dataset/DataGenerator.scala:12-18); java.util.Random is a 48-bit LCG."""
    mask = (1 << 48) - 1
    s = (seed ^ 0x5DEECE66D) & mask
    out = bytearray(n)
    bound = 93
    for i in range(n):
        while True:
            s = (s * 0x5DEECE66D + 0xB) & mask
            bits = s >> 17                      # next(31)
            val = bits % bound
            if bits - val + (bound - 1) < (1 << 31):
                break
        out[i] = val + 32
    return bytes(out)


def benchmark_file(kind: str, n: int) -> bytes:
    """benchmark/benchmark-data/<kind>-<n>.txt (DataGenerator.scala:44-67)."""
    if kind == "random":
        return fixture("randomASCII.txt")[:n]
    if kind == "repeating":
        return b"a" * n
    if kind == "real":
        return fixture("real-50000.txt")[:n]
    raise KeyError(kind)


def random_ascii(n: int, seed: int = 4444) -> np.ndarray:
    """n bytes uniform over the 93 symbols 0x20..0x7C (config 3 scaled up)."""
    rng = np.random.default_rng(seed)
    return (rng.integers(0, 93, size=n, dtype=np.uint8) + 32).astype(np.uint8)


def repeat_to(data: bytes, n: int) -> np.ndarray:
    """Plain concatenation of `data` to exactly n bytes."""
    src = np.frombuffer(data, dtype=np.uint8)
    reps = -(-n // src.size)
    return np.tile(src, reps)[:n].copy()


def structured_random(n: int, seed: int) -> bytes:
    """Random mix of literal runs, short/long back-references and byte runs, in the spirit of
    snappy_unittest.cc:686-730 (RandomizedTest); exercises skip, chained copies and long copies."""
    rng = np.random.default_rng(seed)
    out = bytearray()
    while len(out) < n:
        kind = rng.integers(0, 10)
        if kind < 3 or len(out) < 8:
            k = int(rng.integers(1, 40))
            alpha = int(rng.integers(2, 96))
            out += bytes((rng.integers(0, alpha, size=k) + 32).astype(np.uint8))
        elif kind < 8:
            off = int(rng.integers(1, min(len(out), 70000) + 1))
            k = int(rng.integers(4, 24)) if kind < 7 else int(rng.integers(24, 400))
            for _ in range(k):
                out.append(out[-off])
        elif kind == 8:
            out += bytes([int(rng.integers(0, 256))]) * int(rng.integers(1, 200))
        else:
            k = int(rng.integers(1, 300))
            out += bytes(rng.integers(0, 256, size=k, dtype=np.uint8))
    return bytes(out[:n])
