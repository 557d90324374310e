"""One fresh process that drives every exact engine (lane, pipelined, global-table, window, page, small-buffer),
the fast mode, the batch / framing / decode / no-allocation / host entries on small inputs with tight strides and
checks every result against the oracle (scripts/all_engines_check.py), with the fused framed reader forced."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_all_engines_fresh_process():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "all_engines_check.py")], capture_output=True,
                       text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    for part in ("device entry: ok", "batch entry: ok", "framing / decode: ok", "noalloc / host entry: ok"):
        assert part in r.stdout
