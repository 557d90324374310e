#!/usr/bin/env python
"""Generate tests/golden/ from the reference (run in the build container only).

Inputs are the reference's own data files (/root/reference/data, benchmark/benchmark-data,
tests/snappy/testdata) stored xz-compressed as test fixtures; expected outputs come from the
reference's own Snappy (oracle/_ref/libsnappy_ref.so = tests/snappy compiled unmodified).
Also records the known-answer compressed lengths of benchmark/sw_results.csv.

    python tests/golden/make_golden.py
"""
import csv
import hashlib
import json
import lzma
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib  # noqa: E402
import corpus  # noqa: E402

REF = "/root/reference"
DATA = os.path.join(HERE, "data")


def fnv1a32(b: bytes) -> str:
    h = 0x811C9DC5
    for c in b:
        h = ((h ^ c) * 0x01000193) & 0xFFFFFFFF
    return "%08x" % h


def store(name: str, data: bytes):
    with lzma.open(os.path.join(DATA, name + ".xz"), "wb", preset=9 | lzma.PRESET_EXTREME) as f:
        f.write(data)


def main():
    os.makedirs(DATA, exist_ok=True)
    oracle_lib.build()
    ref = oracle_lib.ref()
    files = {
        "fox-in-socks.txt": "data/fox-in-socks.txt",
        "all-mtg-cards.txt": "data/all-mtg-cards.txt",
        "randomASCII.txt": "data/randomASCII.txt",
        "real-50000.txt": "benchmark/benchmark-data/real-50000.txt",
        "alice29.txt": "tests/snappy/testdata/alice29.txt",
        "html": "tests/snappy/testdata/html",
        "geo.protodata": "tests/snappy/testdata/geo.protodata",
        "fireworks.jpeg": "tests/snappy/testdata/fireworks.jpeg",
    }
    man = {"_generator": "tests/golden/make_golden.py", "_oracle": "oracle/_ref/libsnappy_ref.so "
           "(reference tests/snappy @ d58cd61, built by oracle/Makefile)", "files": {}, "kats": {},
           "sweep": {}, "sw_results_csv": {}}
    for name, rel in files.items():
        data = open(os.path.join(REF, rel), "rb").read()
        store(name, data)
        comp = ref.compress(data)
        assert ref.uncompress(comp) == data
        man["files"][name] = {"source": rel, "n": len(data), "compressed_n": len(comp),
                              "fnv1a32": fnv1a32(comp), "sha256": hashlib.sha256(comp).hexdigest(),
                              "input_sha256": hashlib.sha256(data).hexdigest()}
    # real-10000 / real-20000 must be prefixes of real-50000 so one fixture serves all three
    r50 = open(os.path.join(REF, "benchmark/benchmark-data/real-50000.txt"), "rb").read()
    for n in (10000, 20000):
        assert open(os.path.join(REF, "benchmark/benchmark-data/real-%d.txt" % n), "rb").read() == r50[:n]
    # small known-answer vectors, full bytes in hex
    kats = {
        "empty": b"",
        "one": b"x",
        "thistle": corpus.THISTLE,
        "lt15": b"abcdabcdabcdab",
        "exact15": b"abcdabcdabcdabc",
        "a4096": b"a" * 4096,
        "a65536": b"a" * 65536,
        "a70000": b"a" * 70000,
        "ab_period3_300": b"abc" * 100,
        "zeros_100": bytes(100),
    }
    for k, data in kats.items():
        comp = ref.compress(data)
        ent = {"n": len(data), "compressed_n": len(comp), "fnv1a32": fnv1a32(comp)}
        if len(comp) <= 4096:
            ent["hex"] = comp.hex()
        if len(data) <= 256:
            ent["input_hex"] = data.hex()
        man["kats"][k] = ent
    # size sweep of benchmark/benchmark-data (oracle on the actual files) + sw_results.csv lengths
    csv_len = {}
    with open(os.path.join(REF, "benchmark/sw_results.csv")) as f:
        for row in csv.DictReader(f, delimiter=";"):
            csv_len[(row["type"], int(row["length"]))] = int(row["compressed-length"])
    for kind in ("random", "real", "repeating"):
        for n in corpus.SWEEP_SIZES:
            data = open(os.path.join(REF, "benchmark/benchmark-data/%s-%d.txt" % (kind, n)), "rb").read()
            assert data == corpus.benchmark_file(kind, n), (kind, n)
            comp = ref.compress(data)
            man["sweep"]["%s-%d" % (kind, n)] = {"compressed_n": len(comp), "fnv1a32": fnv1a32(comp)}
            key = ("repeat" if kind == "repeating" else kind, n)
            if key in csv_len:
                man["sw_results_csv"]["%s-%d" % (kind, n)] = csv_len[key]
    # the reference's own corrupted streams (snappy_unittest.cc:585-600 CorruptedTest: each must be refused
    # by IsValidCompressedBuffer and Uncompress, and must not announce a crazy length)
    man["bad"] = {}
    for i in (1, 2, 3):
        rel = "tests/snappy/testdata/baddata%d.snappy" % i
        data = open(os.path.join(REF, rel), "rb").read()
        store("baddata%d.snappy" % i, data)
        try:
            ulen = ref.uncompressed_length(data)
        except ValueError:
            ulen = None
        assert not ref.is_valid(data)
        man["bad"]["baddata%d.snappy" % i] = {"source": rel, "n": len(data), "announced_length": ulen,
                                              "reference_is_valid": False,
                                              "input_sha256": hashlib.sha256(data).hexdigest()}
    with open(os.path.join(HERE, "manifest.json"), "w") as f:
        json.dump(man, f, indent=1, sort_keys=True)
    print("wrote", len(man["files"]), "files,", len(man["kats"]), "kats,", len(man["sweep"]), "sweep points")


if __name__ == "__main__":
    main()
