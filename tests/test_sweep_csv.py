"""Benchmark tooling parity: the size sweep in the reference's CSV shape,
parsed the way benchmark/csv_scan.py:4-36 parses hw_results.csv / sw_results.csv."""
import pytest

import corpus
import __graft_entry__ as g

g.load_package()
from compression_accelerator_b200 import sweep_csv

KIND = {"random": "random", "real": "real", "repeat": "repeating"}


def _data(typ, n):
    return corpus.benchmark_file(KIND[typ], n)


def test_csv_shape_with_oracle_lengths():
    """Host logic only: rows, header and csv_scan's derived columns, lengths from the CPU oracle."""
    import oracle_lib
    chk = oracle_lib.checker()
    calls = []

    def run(data):
        calls.append(len(data))
        return 1000 + len(calls), len(chk.compress(data))

    rows = sweep_csv.sweep_rows(run, _data, repeats=2)
    assert len(rows) == 36 and len(calls) == 72
    text = sweep_csv.to_csv(rows)
    lines = text.strip().split("\n")
    assert lines[0] == "type;length;cycles;compressed-length"       # hw_results.csv:1
    assert lines[1].startswith("random;10;") and lines[-1].startswith("repeat;50000;")
    res = sweep_csv.scan(text)
    assert set(res) == {"random", "real", "repeat"}
    assert all(len(res[t]) == 12 for t in res)
    ratio, eff = res["repeat"][50000]
    assert ratio > 20 and eff > 0
    assert res["random"][10000][0] < 1.0                              # incompressible: stream is longer
    # best-of: the lower cycle count of the two runs of a point is kept
    assert rows[0][2] == 1001


@pytest.mark.gpu
def test_csv_from_device(pkg, oracle):
    dev = pkg.RoccDevice()
    try:
        rows = sweep_csv.sweep_rows(sweep_csv.device_runner(dev), _data, repeats=2)
    finally:
        dev.close()
    for typ, n, cycles, clen in rows:
        assert clen == len(oracle.compress(_data(typ, n))), (typ, n)
        assert cycles > 0
    res = sweep_csv.scan(sweep_csv.to_csv(rows))
    assert len(res["real"]) == 12


def test_committed_bench_sweep_csv():
    """profiles/r02_size_sweep.csv (written from bench.py's size sweep: every point's buffer replicated to
    >= 1 GiB through the batch entry) parses the way benchmark/csv_scan.py:9-27 parses the reference's files,
    covers the reference's 36 points, and carries the reference Snappy's compressed lengths."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "profiles", "r02_size_sweep.csv")).read()
    assert text.splitlines()[0] == "type;length;cycles;compressed-length"
    res = sweep_csv.scan(text)
    assert sorted(res) == ["random", "real", "repeat"]                      # csv_scan.py:22-27 spellings
    man = json.load(open(os.path.join(root, "tests", "golden", "manifest.json")))["sweep"]
    for typ, rows in res.items():
        assert sorted(rows) == sorted(corpus.SWEEP_SIZES)
        for n, (ratio, cyc_per_byte) in rows.items():
            want = man["%s-%d" % ("repeating" if typ == "repeat" else typ, n)]["compressed_n"]
            assert abs(ratio - n / want) < 1e-12 and cyc_per_byte > 0
