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


@pytest.mark.gpu
def test_differential_decode_run_fresh_process():
    """scripts/fuzz_decode.py: valid and mutated streams (flipped bytes, truncations, appended bytes) through the device
    decompressor; verdict and bytes against the reference decompressor.  A 120-second run (seed 5: 21,301 streams,
    10,275 of them refused by both, no mismatch) is recorded in DESIGN.md section 7a."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "fuzz_decode.py"), "--seconds", "10", "--seed", "2"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert " 0 mismatches" in r.stdout


@pytest.mark.gpu
def test_differential_framing_run_fresh_process():
    """scripts/fuzz_frame.py: the device framing writer against the restatement byte for byte; the device reader on valid,
    hand-chunked (uncompressed / padding / skippable / reserved chunks, repeated identifiers) and mutated framed
    streams against the restatement's verdict and bytes.  An 80-second run (seed 3: 13,064 streams, 7,364 refused by
    both, no mismatch) is recorded in DESIGN.md section 7b."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "fuzz_frame.py"), "--seconds", "10", "--seed", "4"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert " 0 mismatches" in r.stdout


@pytest.mark.gpu
def test_differential_batch_decode_run_fresh_process():
    """scripts/fuzz_decode_batch.py: batches of valid and mutated streams through snappy_b200_uncompress_batch_device
    (staged page kernel and global-memory kernel); every verdict and accepted byte against the reference decompressor.
    A 90-second run (seed 9: 694,011 streams in 14,470 batches, 175,690 refused by both, no mismatch) is recorded in
    DESIGN.md section 7a."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "fuzz_decode_batch.py"), "--seconds", "10", "--seed", "6"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert " 0 mismatches" in r.stdout
