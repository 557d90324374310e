"""FAST mode (parallel parse).  CPU part: the sequential model of the mode (oracle/snappy_port.c,
port_fast_compress) round-trips through the reference decompressor and stays within the stated
2 % of the reference's compressed size.  GPU part: the CUDA kernel's stream equals the model's,
byte for byte (the mode is deterministic), through the C-ABI."""
import numpy as np
import pytest

import corpus
import oracle_lib

TOL = 1.02          # north_star: compression-ratio loss of the fast mode versus the reference <= 2 %
MAN = corpus.manifest()


def inputs():
    yield "thistle", corpus.THISTLE
    for name in sorted(MAN["files"]):
        yield name, corpus.fixture(name)
    yield "a70000", b"a" * 70000
    yield "abc300", b"abc" * 100
    yield "random1m", bytes(corpus.random_ascii(1 << 20, 5))
    for seed, n in enumerate([1, 3, 4, 5, 31, 32, 33, 100, 4096, 65535, 65536, 65537, 200000]):
        yield "struct%d" % n, corpus.structured_random(n, 100 + seed)
    yield "empty", b""
    # purely periodic data, period a multiple of the 32-position window: every copy is 1,024 bytes long and the
    # windows inside it are marked without being evaluated first -- their evaluators may then find candidates of a
    # LATER window in the ring (an out-of-range read in a full 64 KiB fragment before the fix; scripts/fuzz_decode.py)
    for per, n in ((32, 65536), (64, 65536), (128, 190589), (256, 131072), (96, 65536), (60, 65536)):
        base = bytes(np.random.default_rng(per).integers(0, 256, size=per, dtype=np.uint8))
        yield "periodic%d_%d" % (per, n), (base * (n // per + 2))[:n]


@pytest.mark.parametrize("name,data", list(inputs()), ids=[n for n, _ in inputs()])
def test_model_roundtrip_and_ratio(oracle, name, data):
    port = oracle_lib.port()
    comp = port.fast_compress(data)
    assert oracle.uncompress(comp, len(data)) == data
    ref = oracle.compress(data)
    if len(data) >= 4096:
        assert len(comp) <= TOL * len(ref), (len(comp), len(ref))


@pytest.mark.gpu
@pytest.mark.parametrize("name,data", list(inputs()), ids=[n for n, _ in inputs()])
def test_gpu_fast_equals_model(ctx, oracle, name, data):
    comp = ctx.compress(data, mode="fast")
    want = oracle_lib.port().fast_compress(data)
    assert len(comp) == len(want)
    assert comp == want
    assert oracle.uncompress(comp, len(data)) == data


@pytest.mark.gpu
def test_gpu_fast_small_lengths(ctx, oracle):
    rng = np.random.default_rng(11)
    base = bytes(rng.integers(97, 100, size=400, dtype=np.uint8))
    port = oracle_lib.port()
    for n in list(range(0, 70)) + [127, 128, 129, 255, 256, 257, 399]:
        assert ctx.compress(base[:n], mode="fast") == port.fast_compress(base[:n]), n


@pytest.mark.gpu
def test_gpu_fast_batch_pages(ctx, oracle):
    """Config 4 shape in fast mode: every 4 KiB page its own stream, equal to the model's."""
    text = corpus.fixture("all-mtg-cards.txt")
    count = 256
    pages = bytearray(text[:4096 * (count - 32)]) + bytes(corpus.random_ascii(4096 * 16, 2)) + b"a" * (4096 * 16)
    buf = np.frombuffer(bytes(pages), dtype=np.uint8)
    out, ostride, sizes = ctx.compress_batch(buf, 4096, 4096, count, mode="fast")
    port = oracle_lib.port()
    for i in range(count):
        want = port.fast_compress(bytes(pages[i * 4096:(i + 1) * 4096]))
        assert bytes(out[i * ostride:i * ostride + sizes[i]]) == want, i


@pytest.mark.gpu
def test_gpu_fast_large_stream(ctx, oracle):
    """128 MiB through the chunk-pipelined host entry in fast mode: equals the model, round-trips,
    and is no larger than 1.02 x the reference's stream."""
    n = (128 << 20) + 777
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    data = host.tobytes()
    comp = ctx.compress(data, mode="fast")
    assert comp == oracle_lib.port().fast_compress(data)
    assert oracle.uncompress(comp, n) == data
    assert len(comp) <= TOL * len(oracle.compress(data))
