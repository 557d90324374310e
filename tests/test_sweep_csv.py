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
