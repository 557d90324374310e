"""The lane parser's per-fragment routine (csrc/parse_lane.cuh), compiled for the HOST by
tests/host_lane/lane_host.cc, against the CPU oracle: byte-identical streams on every fixture, KAT and
edge size.  Pins the logic the CUDA kernel instantiates without needing a GPU (the `-m gpu` tests then
pin the kernel itself through the C-ABI)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import corpus

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_lane", "lane_host.cc")
SO = os.path.join(HERE, "host_lane", "liblane_host.so")


@pytest.fixture(scope="module")
def lane():
    core = os.path.join(os.path.dirname(HERE), "compression-accelerator_b200", "csrc", "parse_lane.cuh")
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(core)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", SO, SRC],
                       check=True)
    lib = C.CDLL(SO)
    lib.lane_host_compress.restype = C.c_size_t
    lib.lane_host_compress.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_int, C.c_uint32]

    def compress(data: bytes, pf: int = 0, epoch0: int = 1) -> bytes:
        out = C.create_string_buffer(32 + len(data) + len(data) // 6 + 16)
        n = lib.lane_host_compress(data, len(data), out, pf, epoch0)
        assert n, "lane parser: records do not cover the fragment or the running size is wrong"
        return out.raw[:n]
    return compress


@pytest.mark.parametrize("pf", [0, 1, 2, 32, 34])
def test_fixtures(lane, oracle, pf):
    for name in sorted(corpus.manifest()["files"]):
        data = corpus.fixture(name)
        assert lane(data, pf) == oracle.compress(data), name


def test_small_lengths_and_kats(lane, oracle):
    rng = np.random.default_rng(7)
    base = bytes(rng.integers(97, 101, size=300, dtype=np.uint8))
    for n in list(range(0, 80)) + [255, 256, 257, 299]:
        assert lane(base[:n]) == oracle.compress(base[:n]), n
    for data in (corpus.THISTLE, b"a" * 4096, b"a" * 65536, b"a" * 70000, b"abc" * 100, b"abcdabcdabcdab"):
        assert lane(data) == oracle.compress(data)
    for kind in ("random", "real", "repeating"):
        for n in corpus.SWEEP_SIZES:
            data = corpus.benchmark_file(kind, n)
            assert lane(data, 2) == oracle.compress(data), (kind, n)


@pytest.mark.parametrize("seed", range(12))
def test_structured_random(lane, oracle, seed):
    n = [1, 14, 15, 16, 100, 4096, 65535, 65536, 65537, 70000, 200000, 300001][seed]
    data = corpus.structured_random(n, seed)
    assert lane(data, (0, 1, 2, 32, 34)[seed % 5]) == oracle.compress(data)


def test_epoch_wrap_and_stale_entries(lane, oracle):
    """Stale entries of earlier fragments (and of epochs about to wrap) must read as empty."""
    data = corpus.fixture("all-mtg-cards.txt")[:65536 * 6 + 999]
    want = oracle.compress(data)
    for epoch0 in (1, 65533, 65535):
        assert lane(data, 34, epoch0) == want, epoch0
