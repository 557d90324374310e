"""GPU parity of the lane parser (csrc/parse_lane.cuh: one thread per fragment, fat table entries in
global memory, epochs instead of clears) through the C-ABI, against the CPU oracle, byte for byte.
`mode="exact_lane"` forces the kernel whatever the size; `mode="exact"` picks it for large streams."""
import hashlib

import numpy as np
import pytest

import corpus

pytestmark = pytest.mark.gpu
MAN = corpus.manifest()
LANE = "exact_lane"


@pytest.mark.parametrize("name", sorted(MAN["files"]))
def test_lane_golden_files(ctx, name):
    data = corpus.fixture(name)
    ent = MAN["files"][name]
    comp = ctx.compress(data, mode=LANE)
    assert len(comp) == ent["compressed_n"]
    assert hashlib.sha256(comp).hexdigest() == ent["sha256"]


def test_lane_small_lengths_and_kats(ctx, oracle):
    rng = np.random.default_rng(7)
    base = bytes(rng.integers(97, 101, size=300, dtype=np.uint8))
    for n in list(range(0, 80)) + [255, 256, 257, 299]:
        assert ctx.compress(base[:n], mode=LANE) == oracle.compress(base[:n]), n
    for data in (corpus.THISTLE, b"a" * 4096, b"a" * 65536, b"a" * 70000, b"abc" * 100, b"abcdabcdabcdab"):
        assert ctx.compress(data, mode=LANE) == oracle.compress(data)
    for kind in ("random", "real", "repeating"):
        for n in corpus.SWEEP_SIZES:
            data = corpus.benchmark_file(kind, n)
            assert ctx.compress(data, mode=LANE) == oracle.compress(data), (kind, n)


@pytest.mark.parametrize("seed", range(12))
def test_lane_structured_random(ctx, oracle, seed):
    n = [1, 14, 15, 16, 100, 4096, 65535, 65536, 65537, 70000, 200000, 300001][seed]
    data = corpus.structured_random(n, seed)
    assert ctx.compress(data, mode=LANE) == oracle.compress(data)


def test_lane_stale_entries_between_calls(ctx, oracle):
    """The table pool is never cleared between calls: entries of earlier calls must read as empty."""
    a = corpus.fixture("all-mtg-cards.txt")[:65536 * 9 + 5]
    b = corpus.structured_random(65536 * 7 + 123, 99)
    c = corpus.fixture("alice29.txt")
    for data in (a, b, a, c, b, c, a):
        assert ctx.compress(data, mode=LANE) == oracle.compress(data)


def test_lane_unaligned_device_source(ctx, oracle):
    import torch
    data = corpus.fixture("alice29.txt")
    buf = torch.zeros(len(data) + 64, dtype=torch.uint8, device="cuda")
    for shift in (1, 2, 3, 8, 17):
        src = buf[shift:shift + len(data)]
        src.copy_(torch.frombuffer(bytearray(data), dtype=torch.uint8))
        torch.cuda.synchronize()
        dst, out_n = ctx.compress_device(src, mode=LANE)
        ctx.sync()
        n = int(out_n.item())
        assert bytes(dst[:n].cpu().numpy()) == oracle.compress(data), shift


@pytest.mark.parametrize("length", [1, 14, 15, 100, 256, 1000, 4096, 5000, 8193, 16384, 50000, 65536])
def test_lane_batch_sizes(ctx, oracle, length):
    """Batch entry (each buffer its own stream, table sized from the buffer length) on the lane parser."""
    text = corpus.fixture("alice29.txt")
    count = 40
    stride = length + 3
    buf = np.frombuffer((text * 20)[:stride * count], dtype=np.uint8).copy()
    out, ostride, sizes = ctx.compress_batch(buf, length, stride, count, mode=LANE)
    for i in range(count):
        want = oracle.compress(bytes(buf[i * stride:i * stride + length]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


def test_lane_host_entry_chunk_pipeline(ctx, oracle):
    """Host entry >= 24 MiB: chunked H2D / parse / emit / D2H with one lane-parser launch per chunk."""
    n = (70 << 20) + 12345
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    host[5 << 20:(5 << 20) + (3 << 20)] = corpus.random_ascii(3 << 20, 3)
    host[(40 << 20) + 77:(40 << 20) + 77 + (2 << 20)] = 97
    data = host.tobytes()
    comp = ctx.compress(data, mode=LANE)
    assert comp == oracle.compress(data)


def test_host_entry_tail_first_hybrid(ctx, oracle):
    """Host entry, exact mode, >= 4096 fragments: the tail of the input crosses PCIe first and goes to the
    lane parser, the head follows in order on the pipelined parser; the stream is still the reference's."""
    n = (300 << 20) + 4321
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    host[(290 << 20):(290 << 20) + (2 << 20)] = corpus.random_ascii(2 << 20, 5)     # literal-heavy stretch in the tail
    data = host.tobytes()
    comp = ctx.compress(data)
    want = oracle.compress(data)
    assert len(comp) == len(want)
    assert hashlib.sha256(comp).hexdigest() == hashlib.sha256(want).hexdigest()


def test_large_stream_every_exact_engine(ctx, oracle):
    """256 MiB of concatenated MTG text through the default exact mode, the lane parser and the pipelined
    parser: each WHOLE stream equals the reference's, byte for byte."""
    import torch
    n = 256 << 20
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    src = torch.from_numpy(host).cuda()
    want = hashlib.sha256(oracle.compress(host.tobytes())).hexdigest()
    for mode in ("exact", LANE, "exact_pipe"):
        dst, out_n = ctx.compress_device(src, mode=mode)
        ctx.sync()
        total = int(out_n.item())
        assert hashlib.sha256(dst[:total].cpu().numpy().tobytes()).hexdigest() == want, mode


_WAVES = r"""
import hashlib, os, sys
sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import corpus, oracle_lib
import __graft_entry__ as g
pkg = g.load_package()
ctx = pkg.Context(0)
orc = oracle_lib.checker()
for n in (65536 * 100 + 777, 65536 * 33, 65536 * 32 + 1, 65536 * 31 + 65535):
    data = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n).tobytes()
    want = orc.compress(data)
    assert ctx.compress(data, mode="exact_lane") == want, n
assert ctx.compress(b"", mode="exact_lane") == orc.compress(b"")
print("waves: ok")
"""


def test_lane_waves_fresh_process():
    """Calls whose tables exceed the pool cap run in equal waves over one pool (capi.cu lane_parse_all).  The cap is
    20 GiB (only the forced mode gets there: exact mode sends such streams to the global-table parser); here it is
    lowered to 16 MiB = 32 fragments a wave, so that a 6.5 MiB stream takes four waves."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, SNAPPY_B200_LANE_POOL_MIB="16")
    r = subprocess.run([sys.executable, "-c", _WAVES], capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    assert "waves: ok" in r.stdout
