"""GPU parity tests proper: the CUDA path, called through the C-ABI, against the CPU oracle
(bit-exact: this is byte/integer work) and against the committed golden vectors."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import corpus

pytestmark = pytest.mark.gpu
MAN = corpus.manifest()


def roundtrip(oracle, comp: bytes, data: bytes):
    assert oracle.uncompress(comp, len(data)) == data


@pytest.mark.parametrize("name", sorted(MAN["files"]))
def test_golden_files_exact(ctx, oracle, name):
    """Every reference data file: GPU stream == golden sha256 (made by the reference's Snappy)."""
    data = corpus.fixture(name)
    ent = MAN["files"][name]
    comp = ctx.compress(data)
    assert len(comp) == ent["compressed_n"]
    assert hashlib.sha256(comp).hexdigest() == ent["sha256"]
    roundtrip(oracle, comp, data)


@pytest.mark.parametrize("name", sorted(MAN["kats"]))
def test_kats_exact(ctx, name):
    ent = MAN["kats"][name]
    data = bytes.fromhex(ent["input_hex"]) if "input_hex" in ent else {
        "a4096": b"a" * 4096, "a65536": b"a" * 65536, "a70000": b"a" * 70000,
        "ab_period3_300": b"abc" * 100}[name]
    comp = ctx.compress(data)
    assert len(comp) == ent["compressed_n"]
    if "hex" in ent:
        assert comp.hex() == ent["hex"]


def test_thistle_through_rocc_driver_contract(pkg, oracle):
    """tests/compression.c:41-50: 130-byte string into a caller-sized 256-byte array."""
    dest = C.create_string_buffer(256)
    n = pkg.rocc_compress(corpus.THISTLE, 130, dest)
    assert n == 112
    assert dest.raw[:n] == oracle.compress(corpus.THISTLE)


@pytest.mark.parametrize("kind", ["random", "real", "repeating"])
def test_benchmark_sweep_exact(ctx, oracle, kind):
    """benchmark/benchmark-data size sweep (DataGenerator.scala:27): lengths as sw_results.csv."""
    for n in corpus.SWEEP_SIZES:
        data = corpus.benchmark_file(kind, n)
        comp = ctx.compress(data)
        assert comp == oracle.compress(data), (kind, n)
        assert len(comp) == MAN["sweep"]["%s-%d" % (kind, n)]["compressed_n"]


def test_all_small_lengths(ctx, oracle):
    rng = np.random.default_rng(7)
    base = bytes(rng.integers(97, 101, size=300, dtype=np.uint8))
    for n in list(range(0, 80)) + [255, 256, 257, 299]:
        assert ctx.compress(base[:n]) == oracle.compress(base[:n]), n


@pytest.mark.parametrize("seed", range(12))
def test_structured_random_exact(ctx, oracle, seed):
    n = [1, 14, 15, 16, 100, 4096, 65535, 65536, 65537, 70000, 200000, 300001][seed]
    data = corpus.structured_random(n, seed)
    comp = ctx.compress(data)
    assert comp == oracle.compress(data)
    roundtrip(oracle, comp, data)


def test_unaligned_device_source(ctx, oracle):
    """Source not 16-byte aligned: the TMA path must fall back to ordinary loads."""
    import torch
    data = corpus.fixture("alice29.txt")
    buf = torch.zeros(len(data) + 64, dtype=torch.uint8, device="cuda")
    for shift in (1, 3, 8, 17):
        src = buf[shift:shift + len(data)]
        src.copy_(torch.frombuffer(bytearray(data), dtype=torch.uint8))
        torch.cuda.synchronize()
        dst, out_n = ctx.compress_device(src)
        ctx.sync()
        n = int(out_n.item())
        assert bytes(dst[:n].cpu().numpy()) == oracle.compress(data), shift


def test_batch_pages_exact(ctx, oracle):
    """Config 4 shape: independent 4 KiB buffers, each its own stream (header 80 20)."""
    text = corpus.fixture("all-mtg-cards.txt")
    count = 512
    pages = bytearray(text[:4096 * (count - 64)])
    pages += bytes(corpus.random_ascii(4096 * 32, 1)) + b"a" * (4096 * 32)
    out, ostride, sizes = ctx.compress_batch(np.frombuffer(bytes(pages), dtype=np.uint8), 4096, 4096, count)
    ref_out, ref_stride, ref_sizes = oracle.compress_batch(bytes(pages), 4096, 4096, count)
    assert ostride == ref_stride and (sizes == ref_sizes).all()
    for i in range(count):
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == bytes(ref_out[i * ostride:i * ostride + sizes[i]]), i
        assert bytes(out[i * ostride:i * ostride + 2]) == b"\x80\x20"


@pytest.mark.parametrize("length", [1, 14, 15, 100, 255, 256, 257, 300, 512, 1000, 2048, 4096, 5000, 8192, 8193, 12000, 16384, 50000,
                                    65536])
def test_batch_size_sweep(ctx, oracle, length):
    text = corpus.fixture("alice29.txt")
    count = 16
    stride = length + 3          # ragged stride: sources are not 16-byte aligned
    buf = np.frombuffer((text * 8)[:stride * count], dtype=np.uint8).copy()
    out, ostride, sizes = ctx.compress_batch(buf, length, stride, count)
    for i in range(count):
        want = oracle.compress(bytes(buf[i * stride:i * stride + length]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


@pytest.mark.parametrize("length", [513, 600, 1000, 4096, 4097, 8192])
def test_batch_incompressible_pages_emitted_by_the_parser(ctx, oracle, length):
    """A page that parses to ONE literal is written by the parse kernel itself (from shared memory); its
    neighbours go through the size / emit kernels.  Mixed batch, ragged stride, every stream vs the oracle."""
    count = 48
    stride = length + 5
    rnd = corpus.random_ascii(stride * count, 77)
    txt = np.frombuffer((corpus.fixture("alice29.txt") * 4)[:stride * count], dtype=np.uint8)
    buf = rnd.copy()
    for i in range(count):
        if i % 3 == 1:                                        # every third buffer is text
            buf[i * stride:(i + 1) * stride] = txt[i * stride:(i + 1) * stride]
        elif i % 3 == 2:                                      # and every third is random bytes of all 256 values
            buf[i * stride:(i + 1) * stride] = np.random.default_rng(i).integers(0, 256, stride, dtype=np.uint8)
    out, ostride, sizes = ctx.compress_batch(buf, length, stride, count)
    for i in range(count):
        want = oracle.compress(bytes(buf[i * stride:i * stride + length]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


@pytest.mark.parametrize("length,count", [(513, 96), (600, 96), (777, 96), (1000, 96), (1500, 96), (2000, 96), (2048, 96),
                                          (3000, 96), (4096, 96), (6000, 96), (8192, 96),
                                          (4100, 1800), (5000, 1800), (20000, 1800), (50000, 1800)])
def test_batch_short_buffers_emitted_by_the_parsing_warp(ctx, oracle, length, count):
    """Exact mode, buffers of 513 .. 2,048 bytes: the warp that parsed a buffer also writes its elements
    (kernels.cu warp_emit_job), 32 sequences a round -- no size / emit launch.  Mixed batch: text, runs (copies of
    68 bytes and more, split as EmitCopy does), literals of 1 .. 16 bytes (written by their lane), 17 .. 60,
    61 .. 256 and longer (written by the whole warp, 1-, 2- and 3-byte tags), more than 32 sequences per buffer,
    a ragged stride; every stream against the oracle.  The cases with 1,800 buffers of more than 4 KiB go through the
    global-table parser, whose warps emit from the buffer in place (global memory) instead of a staged copy."""
    stride = length + 7
    rng = np.random.default_rng(length)
    text = corpus.fixture("all-mtg-cards.txt")
    txt = np.frombuffer((text * (stride * count // len(text) + 2))[5000:5000 + stride * count], dtype=np.uint8)
    buf = txt.copy()
    for i in range(count):
        b = buf[i * stride:(i + 1) * stride]
        k = i % 8
        if k == 1:                                            # random ASCII with planted repeats: literals of many lengths
            b[:] = corpus.random_ascii(stride, i)
            pos = 0
            for gap in (1, 5, 16, 17, 40, 60, 61, 100, 256, 257, 300):
                pos += gap + 8
                if pos + 8 < length:
                    b[pos:pos + 8] = b[0:8]
        elif k == 2:
            b[:] = 97                                         # one run: a single long copy
        elif k == 3:
            b[:] = np.tile(rng.integers(0, 256, 70, dtype=np.uint8), stride // 70 + 1)[:stride]   # period 70
        elif k == 4:
            b[:] = rng.integers(0, 256, stride, dtype=np.uint8)                                   # one literal
        elif k == 5:
            b[:] = np.tile(np.frombuffer(b"abcdefgh", dtype=np.uint8), stride // 8 + 1)[:stride]
            b[length // 2:length // 2 + 200] = rng.integers(0, 256, 200, dtype=np.uint8)
    src = buf[:(count - 1) * stride + length].copy()
    import torch
    dsrc = torch.from_numpy(src).cuda()
    dst, ostride, sizes = ctx.compress_batch_device(dsrc, length, stride, count)
    ctx.sync()
    out = dst.cpu().numpy()
    sizes = sizes.cpu().numpy()
    for i in (range(count) if count <= 96 else list(range(0, count, 7)) + [count - 1]):
        want = oracle.compress(bytes(src[i * stride:i * stride + length]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


@pytest.mark.parametrize("count", [1, 2, 5, 7])
def test_batch_few_pages(ctx, oracle, count):
    """Page jobs share a CTA (several warps, one job each): counts that leave warps of the last CTA idle."""
    text = corpus.fixture("all-mtg-cards.txt")
    for length in (1024, 4096):
        buf = np.frombuffer(text[1000:1000 + length * count], dtype=np.uint8).copy()
        out, ostride, sizes = ctx.compress_batch(buf, length, length, count)
        for i in range(count):
            want = oracle.compress(bytes(buf[i * length:(i + 1) * length]))
            assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, (length, i)


def test_fragment_range_entry(ctx, oracle):
    """Multi-GPU building block: payload of a run of whole fragments, no header."""
    import torch
    data = corpus.fixture("all-mtg-cards.txt")[:65536 * 5 + 1234]
    full = oracle.compress(data)
    hdr = 3
    parts = []
    for lo, hi in ((0, 65536 * 2), (65536 * 2, 65536 * 5), (65536 * 5, len(data))):
        src = torch.frombuffer(bytearray(data[lo:hi]), dtype=torch.uint8).cuda()
        dst, out_n = ctx.compress_device(src, header=False)
        ctx.sync()
        parts.append(bytes(dst[:int(out_n.item())].cpu().numpy()))
    assert full[:hdr] + b"".join(parts) == full


def test_large_stream_whole_bytes(ctx, oracle):
    """256 MiB of concatenated MTG text: the WHOLE device stream equals the reference's, byte for byte
    (sha256 of both), and decodes back to the input through the CPU decompressor."""
    import torch
    n = 256 << 20
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    src = torch.from_numpy(host).cuda()
    dst, out_n = ctx.compress_device(src)
    ctx.sync()
    total = int(out_n.item())
    comp = dst[:total].cpu().numpy().tobytes()
    assert comp[:5] == bytes([0x80, 0x80, 0x80, 0x80, 0x01])
    want = oracle.compress(host.tobytes())
    assert len(comp) == len(want)
    assert hashlib.sha256(comp).hexdigest() == hashlib.sha256(want).hexdigest()
    assert oracle.uncompress(comp, n) == host.tobytes()


def test_host_entry_chunk_pipeline(ctx, oracle):
    """Inputs of 24 MiB and more go through the chunked H2D/parse/emit/D2H pipeline of the host
    entry; the stream must still be the oracle's, byte for byte (odd length, ragged last chunk)."""
    n = (70 << 20) + 12345
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    host[5 << 20:(5 << 20) + (3 << 20)] = corpus.random_ascii(3 << 20, 3)      # a literal-heavy stretch
    host[(40 << 20) + 77:(40 << 20) + 77 + (2 << 20)] = 97                      # and a long run
    data = host.tobytes()
    comp = ctx.compress(data)
    want = oracle.compress(data)
    assert len(comp) == len(want)
    assert comp == want


def test_rocc_device_protocol(pkg, oracle):
    """The Scala tester's poke/step/peek protocol (Test.scala:17-53) over the C-ABI device object,
    for the benchmark size sweep; prints the `type;length;cycles;compressed-length` rows of
    benchmark/hw_results.csv and checks stream bytes, which the reference harness never did."""
    dev = pkg.RoccDevice()
    try:
        rows = []
        for kind in ("random", "real", "repeating"):
            for n in corpus.SWEEP_SIZES:
                data = corpus.benchmark_file(kind, n)
                dev.load_mem(data, 0)
                cycles, clen, dst = dev.run_compress(n)
                want = oracle.compress(data)
                assert clen == len(want), (kind, n)
                assert dev.read_mem(dst, clen) == want, (kind, n)
                assert clen == MAN["sweep"]["%s-%d" % (kind, n)]["compressed_n"]
                rows.append("%s;%d;%d;%d" % (kind, n, cycles, clen))
        assert len(rows) == 36
    finally:
        dev.close()
