"""The small-buffer kernel's per-buffer routine (csrc/parse_tiny.cuh), compiled for the HOST by
tests/host_lane/tiny_host.cc, against the CPU oracle: byte-identical streams for every length and data
family the kernel serves (and beyond: the routine itself handles one whole fragment)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import corpus

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_lane", "tiny_host.cc")
SO = os.path.join(HERE, "host_lane", "libtiny_host.so")


@pytest.fixture(scope="module")
def tiny():
    core = os.path.join(os.path.dirname(HERE), "compression-accelerator_b200", "csrc", "parse_tiny.cuh")
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(core)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", SO, SRC], check=True)
    lib = C.CDLL(SO)
    lib.tiny_host_compress.restype = C.c_size_t
    lib.tiny_host_compress.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p]

    def compress(data: bytes) -> bytes:
        out = C.create_string_buffer(32 + len(data) + len(data) // 6 + 16)
        n = lib.tiny_host_compress(data, len(data), out)
        return out.raw[:n]
    return compress


def test_every_small_length(tiny, oracle):
    rng = np.random.default_rng(11)
    fams = {"abcd": bytes(rng.integers(97, 101, size=2100, dtype=np.uint8)), "text": corpus.fixture("alice29.txt")[5000:7100],
            "a": b"a" * 2100, "random": corpus.fixture("randomASCII.txt")[:2100], "html": corpus.fixture("html")[:2100]}
    for name, base in fams.items():
        for n in list(range(0, 300)) + [511, 512, 513, 1000, 1023, 1024, 1025, 2000, 2047, 2048]:
            assert tiny(base[:n]) == oracle.compress(base[:n]), (name, n)


def test_sweep_points_and_kats(tiny, oracle):
    for kind in ("random", "real", "repeating"):
        for n in corpus.SWEEP_SIZES:
            data = corpus.benchmark_file(kind, n)
            assert tiny(data) == oracle.compress(data), (kind, n)
    for data in (corpus.THISTLE, b"abcdabcdabcdab", b"abc" * 100, b"a" * 4096, b"a" * 65536, corpus.fixture("fox-in-socks.txt")):
        assert tiny(data) == oracle.compress(data)


@pytest.mark.parametrize("seed", range(8))
def test_structured_random(tiny, oracle, seed):
    n = [15, 16, 100, 777, 1500, 2048, 20000, 65536][seed]
    data = corpus.structured_random(n, 100 + seed)
    assert tiny(data) == oracle.compress(data)
