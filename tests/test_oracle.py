"""Pins the oracle: the C restatement (oracle/snappy_port.c) against the compiled reference
(oracle/_ref), the committed golden vectors and the reference's known-answer tables.
CPU only."""
import hashlib

import numpy as np
import pytest

import corpus
import oracle_lib


def fnv1a32(b: bytes) -> str:
    h = 0x811C9DC5
    for c in b:
        h = ((h ^ c) * 0x01000193) & 0xFFFFFFFF
    return "%08x" % h


MAN = corpus.manifest()
need_ref = pytest.mark.skipif(not oracle_lib.have_ref(), reason="oracle/_ref not built")


@pytest.mark.parametrize("name", sorted(MAN["files"]))
def test_port_matches_golden_files(name):
    data = corpus.fixture(name)
    ent = MAN["files"][name]
    assert len(data) == ent["n"]
    assert hashlib.sha256(data).hexdigest() == ent["input_sha256"]
    comp = oracle_lib.port().compress(data)
    assert len(comp) == ent["compressed_n"]
    assert hashlib.sha256(comp).hexdigest() == ent["sha256"]
    assert fnv1a32(comp) == ent["fnv1a32"]
    assert oracle_lib.port().uncompress(comp, len(data)) == data


# FNV-1a-32 goldens recorded independently at survey time
SURVEY_GOLDENS = {
    "fox-in-socks.txt": (2407, "81ceb610"),
    "all-mtg-cards.txt": (1334237, "9cf82c71"),
    "randomASCII.txt": (100009, "044623a9"),
    "alice29.txt": (88034, "039d479a"),
    "html": (22843, "e7416f8b"),
    "fireworks.jpeg": (123034, "d0faf4f5"),
}


@pytest.mark.parametrize("name", sorted(SURVEY_GOLDENS))
def test_manifest_agrees_with_survey_goldens(name):
    n, h = SURVEY_GOLDENS[name]
    assert MAN["files"][name]["compressed_n"] == n
    assert MAN["files"][name]["fnv1a32"] == h


@pytest.mark.parametrize("name", sorted(MAN["kats"]))
def test_port_matches_kats(name):
    ent = MAN["kats"][name]
    data = bytes.fromhex(ent["input_hex"]) if "input_hex" in ent else {
        "a4096": b"a" * 4096, "a65536": b"a" * 65536, "a70000": b"a" * 70000,
        "ab_period3_300": b"abc" * 100}[name]
    comp = oracle_lib.port().compress(data)
    assert len(comp) == ent["compressed_n"]
    assert fnv1a32(comp) == ent["fnv1a32"]
    if "hex" in ent:
        assert comp.hex() == ent["hex"]


def test_thistle_kat_from_survey():
    """130-byte driver string of tests/compression.c:43 -> 112 bytes."""
    want = ("8201b0496620546964207468652074686973746c6520736966746572206c6966747320686973207369"
            "657665206f6620051a012d052904732c0137306e206865276c6c207368696674011100200d480c7320"
            "696e0d3f106674696e670d4740746f206c69766520756e7369667465642e")
    assert oracle_lib.port().compress(corpus.THISTLE).hex() == want


def test_edge_sizes_from_survey():
    p = oracle_lib.port()
    assert p.compress(b"") == b"\x00"
    assert len(p.compress(b"abcdabcdabcdab")) == 16
    c = p.compress(b"a" * 4096)
    assert len(c) == 196 and c[:2] == b"\x80\x20"
    assert len(p.compress(b"a" * 65536)) == 3077
    assert len(p.compress(b"a" * 70000)) == 3289


@pytest.mark.parametrize("kind", ["random", "real", "repeating"])
def test_sweep_lengths(kind):
    """benchmark/benchmark-data size sweep: port == oracle bytes (fnv), and the lengths
    published in benchmark/sw_results.csv are reproduced for the 33 rows the vendored Snappy
    reproduces (the 3 'real' >= 10000 rows differ by 2-3 bytes in the reference's own table;
    BASELINE.md section 1.2)."""
    for n in corpus.SWEEP_SIZES:
        key = "%s-%d" % (kind, n)
        comp = oracle_lib.port().compress(corpus.benchmark_file(kind, n))
        assert len(comp) == MAN["sweep"][key]["compressed_n"], key
        assert fnv1a32(comp) == MAN["sweep"][key]["fnv1a32"], key
        if not (kind == "real" and n >= 10000):
            assert len(comp) == MAN["sw_results_csv"][key], key


def test_java_random_regenerates_randomascii():
    """DataGenerator.scala:12-18 restated; must reproduce data/randomASCII.txt exactly."""
    assert corpus.java_random_ascii(100000, 4444) == corpus.fixture("randomASCII.txt")


def test_table_entries():
    te = oracle_lib.port().lib.port_table_entries
    assert te(0) == 256 and te(255) == 256 and te(256) == 256 and te(257) == 512
    assert te(4096) == 4096 and te(4097) == 8192 and te(4602) == 8192 and te(4664) == 8192
    assert te(16384) == 16384 and te(16385) == 16384 and te(65536) == 16384


@need_ref
@pytest.mark.parametrize("seed", range(12))
def test_port_vs_ref_structured_random(seed):
    n = [1, 14, 15, 16, 100, 4096, 65535, 65536, 65537, 70000, 200000, 300001][seed]
    data = corpus.structured_random(n, seed)
    a = oracle_lib.port().compress(data)
    b = oracle_lib.ref().compress(data)
    assert a == b
    assert oracle_lib.ref().uncompress(a) == data
    assert oracle_lib.port().uncompress(a, n) == data


@need_ref
def test_port_vs_ref_all_small_lengths():
    rng = np.random.default_rng(7)
    base = bytes(rng.integers(97, 101, size=600, dtype=np.uint8))
    for n in range(0, 600):
        d = base[:n]
        assert oracle_lib.port().compress(d) == oracle_lib.ref().compress(d), n


@need_ref
def test_ref_roundtrip_and_validate():
    r = oracle_lib.ref()
    data = corpus.fixture("alice29.txt")
    c = r.compress(data)
    assert r.is_valid(c) and r.uncompressed_length(c) == len(data) and r.uncompress(c) == data
    assert not r.is_valid(c[:-3])


def test_batch_entry_matches_single():
    data = corpus.fixture("all-mtg-cards.txt")[:4096 * 8]
    p = oracle_lib.port()
    out, ostride, sizes = p.compress_batch(data, 4096, 4096, 8)
    for i in range(8):
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == p.compress(data[i * 4096:(i + 1) * 4096])
        assert bytes(out[i * ostride:i * ostride + 2]) == b"\x80\x20"
