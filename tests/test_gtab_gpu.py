"""GPU parity of the global-table window parser (parse_gtab_kernel: one warp per fragment, uint16 table in
global memory, input read in place, speculative look-ups two windows ahead) through the C-ABI, against the
CPU oracle, byte for byte.  `mode="exact_gtab"` forces it; `mode="exact"` picks it between ~112 and ~832 MiB
and for the chunks of the host entry."""
import hashlib

import numpy as np
import pytest

import corpus

pytestmark = pytest.mark.gpu
MAN = corpus.manifest()
GTAB = "exact_gtab"


@pytest.mark.parametrize("name", sorted(MAN["files"]))
def test_gtab_golden_files(ctx, name):
    data = corpus.fixture(name)
    ent = MAN["files"][name]
    comp = ctx.compress(data, mode=GTAB)
    assert len(comp) == ent["compressed_n"]
    assert hashlib.sha256(comp).hexdigest() == ent["sha256"]


def test_gtab_small_lengths_and_kats(ctx, oracle):
    rng = np.random.default_rng(7)
    base = bytes(rng.integers(97, 101, size=300, dtype=np.uint8))
    for n in list(range(0, 80)) + [255, 256, 257, 299]:
        assert ctx.compress(base[:n], mode=GTAB) == oracle.compress(base[:n]), n
    for data in (corpus.THISTLE, b"a" * 4096, b"a" * 65536, b"a" * 70000, b"abc" * 100, b"abcdabcdabcdab"):
        assert ctx.compress(data, mode=GTAB) == oracle.compress(data)
    for kind in ("random", "real", "repeating"):
        for n in corpus.SWEEP_SIZES:
            data = corpus.benchmark_file(kind, n)
            assert ctx.compress(data, mode=GTAB) == oracle.compress(data), (kind, n)


@pytest.mark.parametrize("seed", range(12))
def test_gtab_structured_random(ctx, oracle, seed):
    n = [1, 14, 15, 16, 100, 4096, 65535, 65536, 65537, 70000, 200000, 300001][seed]
    data = corpus.structured_random(n, seed)
    assert ctx.compress(data, mode=GTAB) == oracle.compress(data)


def test_gtab_tables_are_reused_between_fragments_and_calls(ctx, oracle):
    """More fragments than tables would need a big input; the persistent grid's reuse of a table (zeroed per
    fragment, sized from THAT fragment's length) is exercised by calls of different shapes in a row."""
    a = corpus.fixture("all-mtg-cards.txt")
    b = corpus.structured_random(65536 * 7 + 123, 99)
    c = corpus.fixture("alice29.txt")
    for data in (a, b, c[:5000], a[:65536 * 3 + 17], c, b):
        assert ctx.compress(data, mode=GTAB) == oracle.compress(data)


def test_gtab_unaligned_device_source(ctx, oracle):
    import torch
    data = corpus.fixture("alice29.txt")
    buf = torch.zeros(len(data) + 64, dtype=torch.uint8, device="cuda")
    for shift in (1, 2, 3, 8, 17):
        src = buf[shift:shift + len(data)]
        src.copy_(torch.frombuffer(bytearray(data), dtype=torch.uint8))
        dst, out_n = ctx.compress_device(src, mode=GTAB)
        n = int(out_n.item())
        assert bytes(dst[:n].cpu().numpy()) == oracle.compress(data), shift


@pytest.mark.parametrize("length", [1, 15, 100, 1000, 4096, 8193, 16384, 50000, 65536])
def test_gtab_batch_sizes(ctx, oracle, length):
    text = corpus.fixture("alice29.txt")
    count = 40
    stride = length + 3
    buf = np.frombuffer((text * 20)[:stride * count], dtype=np.uint8).copy()
    out, ostride, sizes = ctx.compress_batch(buf, length, stride, count, mode=GTAB)
    for i in range(count):
        want = oracle.compress(bytes(buf[i * stride:i * stride + length]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


@pytest.mark.parametrize("length", [4097, 8192, 20000])
def test_batch_of_large_buffers_takes_gtab_by_default(ctx, oracle, length):
    """Exact mode, buffers over 4 KiB, 1,792 and more of them: the batch entry picks the global-table parser."""
    text = corpus.fixture("all-mtg-cards.txt")
    count = 2000
    buf = np.frombuffer((text * 14)[:length * count], dtype=np.uint8).copy()
    out, ostride, sizes = ctx.compress_batch(buf, length, length, count)
    for i in list(range(0, count, 97)) + [count - 1]:
        want = oracle.compress(bytes(buf[i * length:(i + 1) * length]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


def test_mid_size_stream_and_host_entry_take_gtab_by_default(ctx, oracle):
    """300 MiB (4,800 fragments): the device entry and the chunked host entry both pick the global-table parser
    in exact mode; whole streams equal the reference's (sha256)."""
    import torch
    n = (300 << 20) + 4321
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    host[(290 << 20):(290 << 20) + (2 << 20)] = corpus.random_ascii(2 << 20, 5)
    host[(100 << 20) + 5:(100 << 20) + 5 + (1 << 20)] = 97
    data = host.tobytes()
    want = hashlib.sha256(oracle.compress(data)).hexdigest()
    assert hashlib.sha256(ctx.compress(data)).hexdigest() == want                    # host entry, chunked
    dst, out_n = ctx.compress_device(torch.from_numpy(host).cuda())
    total = int(out_n.item())
    assert hashlib.sha256(dst[:total].cpu().numpy().tobytes()).hexdigest() == want   # device entry
