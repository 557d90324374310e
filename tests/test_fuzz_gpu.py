"""Differential run (scripts/fuzz_parity.py): random inputs of many shapes through every exact engine and the batch
entry for a fixed time and seed, every stream byte for byte against the CPU checker.  A 150-second run of the same
script (seed 7: 20,558 calls, no mismatch) is recorded in DESIGN.md section 2; this is the short form."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_differential_run_fresh_process():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "fuzz_parity.py"), "--seconds", "12", "--seed", "3"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert " 0 mismatches" in r.stdout
