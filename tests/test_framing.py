"""Snappy framing format (tests/snappy/framing_format.txt): CRC-32C, the framed
writer and the parallel, CRC-checked reader.  The reference vendors only the specification, so the
checker is a CPU restatement (tests/oracle_lib.py frame_compress / frame_uncompress, CRC-32C in
oracle/snappy_port.c) pinned by the RFC 3720 B.4 vectors the specification points at (:43-44)."""
import os

import numpy as np
import pytest

import corpus
import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# RFC 3720 B.4 "CRC Examples": 32 bytes of zeroes, of ones, incrementing, decrementing
RFC_VECTORS = [
    (bytes(32), 0x8A9136AA),
    (bytes([0xff] * 32), 0x62A8AB43),
    (bytes(range(32)), 0x46DD794E),
    (bytes(range(31, -1, -1)), 0x113FDB5C),
]


def _unmask(m):
    r = (m - 0xa282ead8) & 0xffffffff
    return ((r >> 17) | (r << 15)) & 0xffffffff


def test_crc32c_oracle_rfc3720_vectors():
    p = oracle_lib.port()
    for data, want in RFC_VECTORS:
        assert p.crc32c(data) == want
        assert _unmask(p.crc32c_masked(data)) == want
    assert p.crc32c(b"") == 0
    assert p.crc32c(b"123456789") == 0xE3069283                      # the usual check value


def test_framing_restatement_round_trip():
    data = corpus.fixture("alice29.txt")
    framed = oracle_lib.frame_compress(data)
    assert framed[:10] == oracle_lib.STREAM_ID
    assert framed[10] == 0x00                                         # first chunk: compressed data
    assert oracle_lib.frame_uncompress(framed) == data
    # chunk structure: 3 chunks (152,089 bytes), lengths add up to the stream
    pos, chunks = 10, 0
    while pos < len(framed):
        pos += 4 + int.from_bytes(framed[pos + 1:pos + 4], "little")
        chunks += 1
    assert pos == len(framed) and chunks == 3
    bad = bytearray(framed); bad[30] ^= 1
    with pytest.raises(ValueError):
        oracle_lib.frame_uncompress(bytes(bad))


# ---------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_crc32c_device_vectors(ctx):
    import torch
    for data, want in RFC_VECTORS + [(b"123456789", 0xE3069283), (b"a", oracle_lib.port().crc32c(b"a"))]:
        t = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
        got = ctx.crc32c_chunks_device(t)
        ctx.sync()
        assert _unmask(int(got[0].item()) & 0xffffffff) == want, data[:8]


@pytest.mark.gpu
@pytest.mark.parametrize("n,chunk", [(65536, 65536), (200000, 65536), (70001, 4096), (1 << 20, 65536),
                                     (12345, 1000), (5, 3), (65536 * 3 + 1, 65535)])
def test_crc32c_device_chunks_equal_oracle(ctx, n, chunk):
    import torch
    data = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n + 3)[3:]          # unaligned source
    t = torch.from_numpy(np.ascontiguousarray(corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n + 3))).cuda()[3:]
    got = ctx.crc32c_chunks_device(t, chunk)
    ctx.sync()
    got = got.cpu().numpy().astype(np.uint32)
    p = oracle_lib.port()
    raw = data.tobytes()
    want = [p.crc32c_masked(raw[i:i + chunk]) for i in range(0, n, chunk)]
    assert got.tolist() == want


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["alice29.txt", "html", "fireworks.jpeg", "all-mtg-cards.txt"])
def test_frame_compress_equals_restatement(ctx, oracle, name):
    data = corpus.fixture(name)
    got = ctx.frame_compress(data)
    assert got == oracle_lib.frame_compress(data, oracle.compress)
    assert ctx.frame_uncompress(got) == data
    assert oracle_lib.frame_uncompress(got) == data


@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 14, 65535, 65536, 65537, 131072, 200001])
def test_frame_edge_sizes(ctx, oracle, n):
    data = corpus.random_ascii(n).tobytes() if n % 2 else (b"ab" * (n // 2))
    got = ctx.frame_compress(data)
    assert got == oracle_lib.frame_compress(data, oracle.compress)
    assert ctx.frame_uncompress(got) == data


@pytest.mark.gpu
def test_frame_fast_mode_round_trips(ctx):
    data = corpus.fixture("all-mtg-cards.txt")
    got = ctx.frame_compress(data, mode="fast")
    assert oracle_lib.frame_uncompress(got) == data
    assert ctx.frame_uncompress(got) == data


@pytest.mark.gpu
def test_frame_reader_accepts_every_chunk_type(ctx):
    """Uncompressed chunks (:94-106), padding (:109-118), skippable (:131-138), repeated identifier (:76-78)."""
    p = oracle_lib.port()
    a, b = b"hello framing " * 100, corpus.fixture("html")[:5000]

    def chunk(typ, body):
        return bytes([typ]) + len(body).to_bytes(3, "little") + body

    s = oracle_lib.STREAM_ID
    s += chunk(0x01, p.crc32c_masked(a).to_bytes(4, "little") + a)
    s += chunk(0xfe, bytes(37))
    s += chunk(0x80, b"skip me")
    s += oracle_lib.STREAM_ID
    s += chunk(0x00, p.crc32c_masked(b).to_bytes(4, "little") + p.compress(b))
    s += chunk(0xfd, b"")
    assert ctx.frame_uncompress(s) == a + b
    assert oracle_lib.frame_uncompress(s) == a + b


@pytest.mark.gpu
def test_frame_reader_rejects_bad_streams(ctx, pkg):
    p = oracle_lib.port()
    data = corpus.fixture("alice29.txt")
    good = ctx.frame_compress(data)
    cases = {}
    x = bytearray(good); x[14] ^= 0x40; cases["crc"] = bytes(x)                  # stored CRC flipped
    x = bytearray(good); x[40] ^= 0x01; cases["payload"] = bytes(x)              # payload bit flipped
    cases["no id"] = good[10:]
    cases["truncated"] = good[:-5]
    cases["reserved"] = good + bytes([0x02, 0, 0, 0])
    cases["bad id"] = good + bytes([0xff, 6, 0, 0]) + b"sNaPpX"
    big = b"z" * 70000
    cases["chunk too large"] = oracle_lib.STREAM_ID + bytes([0x01]) + (len(big) + 4).to_bytes(3, "little") + \
        p.crc32c_masked(big).to_bytes(4, "little") + big
    for name, s in cases.items():
        with pytest.raises(pkg.SnappyB200Error):
            ctx.frame_uncompress(s)


@pytest.mark.gpu
def test_frame_device_entries_large(ctx):
    """64 MiB through the device entries: chunk count, CRCs, parallel decode, byte equality."""
    import torch
    n = 64 << 20
    host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    src = torch.from_numpy(host).cuda()
    dst, out_n = ctx.frame_compress_device(src)
    ctx.sync()
    total = int(out_n.item())
    back = torch.empty(n, dtype=torch.uint8, device="cuda")
    got = ctx.frame_uncompress_device(dst, total, back)
    assert got == n and torch.equal(back, src)
    # the first chunk is what the restatement writes
    first_len = 4 + int.from_bytes(bytes(dst[11:14].cpu().numpy()), "little")
    want = oracle_lib.frame_compress(host[:65536].tobytes())
    assert bytes(dst[:10 + first_len].cpu().numpy()) == want


# ---- a pin from OUTSIDE the restatement --------------------------------------------------------
# Derived by hand from the specification text alone (tests/snappy/framing_format.txt) and one published
# constant, the CRC-32C check value of the nine ASCII digits "123456789" = 0xE3069283 (the catalogued
# check of CRC-32C/iSCSI, the algorithm RFC 3720 12.1 defines and the specification points at, :43-44):
#   mask (:46-54):  x >> 15 = 0x0001C60D;  x << 17 = 0x25060000;  or = 0x2507C60D;  + 0xA282EAD8 = 0xC78AB0E5
#   stored little-endian (:58): e5 b0 8a c7
#   stream identifier (:68-75): ff 06 00 00 73 4e 61 50 70 59
#   uncompressed chunk (:94-106): type 01, length = 4 (CRC) + 9 = 0x00000d LE -> 01 0d 00 00, CRC, data
#   compressed chunk (:80-91): type 00, body = the data's own Snappy stream = varint(9) 09, literal tag
#     (9-1)<<2 = 20 (format_description.txt:56-64), nine bytes -> 11 bytes, length 4 + 11 = 0x00000f
HAND_CRC = "e5b08ac7"
HAND_UNCOMPRESSED = "ff060000734e61507059" + "010d0000" + HAND_CRC + "313233343536373839"
HAND_COMPRESSED = "ff060000734e61507059" + "000f0000" + HAND_CRC + "0920" + "313233343536373839"


def test_hand_derived_vector_pins_the_restatement():
    """The CPU restatement (the checker of the GPU framing tests) against the hand-derived bytes."""
    assert oracle_lib.port().crc32c(b"123456789") == 0xE3069283
    assert oracle_lib.port().crc32c_masked(b"123456789").to_bytes(4, "little").hex() == HAND_CRC
    assert oracle_lib.frame_compress(b"123456789").hex() == HAND_COMPRESSED
    assert oracle_lib.frame_uncompress(bytes.fromhex(HAND_UNCOMPRESSED)) == b"123456789"
    assert oracle_lib.frame_uncompress(bytes.fromhex(HAND_COMPRESSED)) == b"123456789"


@pytest.mark.gpu
def test_hand_derived_vector_on_the_gpu(ctx):
    """Writer emits exactly the hand-derived compressed-chunk stream; reader accepts both chunk forms."""
    assert ctx.frame_compress(b"123456789").hex() == HAND_COMPRESSED
    assert ctx.frame_uncompress(bytes.fromhex(HAND_UNCOMPRESSED)) == b"123456789"
    assert ctx.frame_uncompress(bytes.fromhex(HAND_COMPRESSED)) == b"123456789"
    x = bytearray.fromhex(HAND_UNCOMPRESSED)
    x[14] ^= 1                                            # a CRC bit: must be refused
    with pytest.raises(Exception):
        ctx.frame_uncompress(bytes(x))


@pytest.mark.gpu
def test_frame_reader_many_small_chunks(ctx):
    """More chunks than the reader's first index guess (2 * (dst_cap >> 16) + 1024): the walk must keep
    counting after the index is full, and the second pass decodes the stream (frame.cuh)."""
    p = oracle_lib.port()
    text = corpus.fixture("alice29.txt")
    parts, s = [], bytearray(oracle_lib.STREAM_ID)
    for i in range(3000):
        piece = text[i * 40:i * 40 + 40]
        body = p.compress(piece)
        s += bytes([0x00]) + (4 + len(body)).to_bytes(3, "little") + p.crc32c_masked(piece).to_bytes(4, "little") + body
        parts.append(piece)
    want = b"".join(parts)
    assert oracle_lib.frame_uncompress(bytes(s)) == want
    assert ctx.frame_uncompress(bytes(s)) == want


@pytest.mark.gpu
def test_bytearray_input(ctx, oracle):
    """bytearray arguments are passed by address (no temporary copy may be freed under the C call)."""
    data = bytearray(corpus.fixture("fox-in-socks.txt"))
    framed = ctx.frame_compress(data)
    assert framed == oracle_lib.frame_compress(bytes(data), oracle.compress)
    assert ctx.frame_uncompress(bytearray(framed)) == bytes(data)
    assert ctx.compress(data) == oracle.compress(bytes(data))


@pytest.mark.gpu
@pytest.mark.parametrize("fused", ["0", "1"])
def test_frame_reader_both_paths(pkg, fused):
    """The framed reader has two device paths (separate walk kernels + decode; walk and decode fused in one
    kernel, the default for large streams).  Both must decode every chunk type, survive an index that is too
    small on the first pass, and refuse malformed streams -- run in a fresh process with the path forced."""
    import subprocess
    import sys
    code = r"""
import os, sys
sys.path.insert(0, os.path.join(%r, "tests")); sys.path.insert(0, %r)
import corpus, oracle_lib
import __graft_entry__ as g
pkg = g.load_package(); ctx = pkg.Context(0); p = oracle_lib.port()
def chunk(t, body): return bytes([t]) + len(body).to_bytes(3, "little") + body
text = corpus.fixture("alice29.txt")
a, b = b"hello framing " * 100, corpus.fixture("html")[:5000]
s = oracle_lib.STREAM_ID + chunk(0x01, p.crc32c_masked(a).to_bytes(4, "little") + a) + chunk(0xfe, bytes(37)) + chunk(0x80, b"skip me")
s += oracle_lib.STREAM_ID + chunk(0x00, p.crc32c_masked(b).to_bytes(4, "little") + p.compress(b)) + chunk(0xfd, b"")
assert ctx.frame_uncompress(s) == a + b
many, parts = bytearray(oracle_lib.STREAM_ID), []
for i in range(3000):
    piece = text[i * 40:i * 40 + 40]
    many += chunk(0x00, p.crc32c_masked(piece).to_bytes(4, "little") + p.compress(piece)); parts.append(piece)
assert ctx.frame_uncompress(bytes(many)) == b"".join(parts)
big = (text * 40)[:5 << 20]
good = ctx.frame_compress(big)
assert ctx.frame_uncompress(good) == big
for bad in (good[10:], good[:-5], good + bytes([0x02, 0, 0, 0]), good[:3000] + bytes([good[3000] ^ 1]) + good[3001:],
            good[:70000] + bytes([0x03, 9, 0, 0]) + good[70004:]):
    try:
        ctx.frame_uncompress(bad)
    except pkg.SnappyB200Error:
        continue
    raise AssertionError("a malformed framed stream was accepted")
print("both-paths ok")
""" % (ROOT, ROOT)
    env = dict(os.environ, SNAPPY_B200_FRAME_FUSED=fused)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "both-paths ok" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])
