"""Device decompressor behind uncompress(source, compressed_length, dest) (tests/compression.c:32-39).
Round trips (streams from the reference Snappy, from our exact and fast modes), the driver contract,
the page-shaped batch, and malformed streams in the spirit of snappy_unittest.cc:536- (truncation,
bad offsets, length lies): rejected, never out of bounds."""
import ctypes as C

import numpy as np
import pytest

import corpus
import oracle_lib

pytestmark = pytest.mark.gpu
MAN = corpus.manifest()


def cases():
    yield "thistle", corpus.THISTLE
    yield "empty", b""
    yield "one", b"x"
    for name in ("fox-in-socks.txt", "alice29.txt", "html", "fireworks.jpeg", "randomASCII.txt"):
        yield name, corpus.fixture(name)
    yield "a70000", b"a" * 70000
    yield "abc", b"abc" * 5000
    yield "struct", corpus.structured_random(150000, 9)


@pytest.mark.parametrize("name,data", list(cases()), ids=[n for n, _ in cases()])
def test_roundtrip_reference_streams(ctx, oracle, pkg, name, data):
    comp = oracle.compress(data)                 # the reference's own stream
    assert ctx.uncompress(comp) == data
    assert ctx.uncompress(ctx.compress(data)) == data
    assert ctx.uncompress(ctx.compress(data, mode="fast")) == data


def test_rocc_driver_roundtrip(pkg):
    """main() of tests/compression.c:41-61: compress into a 256-byte array, uncompress, compare."""
    compressed = C.create_string_buffer(256)
    recovered = C.create_string_buffer(256)
    n = pkg.rocc_compress(corpus.THISTLE, 130, compressed)
    assert n == 112
    assert pkg.rocc_uncompress(compressed, n, recovered) == 0
    assert recovered.raw[:130] == corpus.THISTLE
    assert pkg.rocc_uncompress(b"\x05\xff\xff", 3, recovered) != 0        # garbage: non-zero = failure


def test_copy4_and_long_literal_forms(ctx):
    """Elements our compressor never emits but the format allows: copy-4 and 3/4-byte literal lengths."""
    lit = bytes(range(256)) * 300                                # 76800 bytes
    n = len(lit) + 10
    hdr = bytearray()
    v = n
    while v >= 128:
        hdr.append((v & 0x7f) | 0x80); v >>= 7
    hdr.append(v)
    L = len(lit) - 1
    stream = bytes(hdr) + bytes([62 << 2, L & 0xff, (L >> 8) & 0xff, L >> 16]) + lit
    stream += bytes([3 | ((10 - 1) << 2)]) + (70000).to_bytes(4, "little")   # copy-4: len 10, offset 70000
    want = lit + lit[len(lit) - 70000:len(lit) - 70000 + 10]
    assert ctx.uncompress(stream) == want


def test_malformed_streams_rejected(ctx, oracle, pkg):
    good = oracle.compress(corpus.fixture("fox-in-socks.txt"))
    bad = [
        b"",                                   # no preamble
        b"\x80\x80\x80\x80\x80\x01",           # varint too long
        good[:-1], good[:len(good) // 2],      # truncated
        good + b"\x00",                        # trailing garbage element (1-byte literal without data / too long)
        b"\x0a" + bytes([0x01 | (0 << 2), 0x05]),        # copy before any output (offset 5 > produced 0)
        b"\x08\x0cabcd" + bytes([0x01, 0x00]),           # copy with offset 0
        b"\x04\x0cabcd\x00z",                             # more output than the preamble promises
        b"\x10\x0cabcd",                                  # less output than the preamble promises
        b"\x05" + bytes([61 << 2, 0xff, 0xff]),           # literal longer than the input
    ]
    for s in bad:
        with pytest.raises(pkg.SnappyB200Error) as ei:
            ctx.uncompress(s)
        assert ei.value.status == pkg.api.INVALID_INPUT if hasattr(pkg, "api") else ei.value.status == 1


def test_batch_pages_roundtrip(ctx, oracle):
    """Config 4 shape: 2048 pages compressed on the device, decompressed on the device, compared."""
    import torch
    text = corpus.fixture("all-mtg-cards.txt")
    count = 2048
    pages = np.frombuffer((text * 3)[:4096 * count], dtype=np.uint8).copy()
    pages[4096 * 100:4096 * 110] = 97
    src = torch.from_numpy(pages).cuda()
    comp, ostride, sizes = ctx.compress_batch_device(src, 4096, 4096, count)
    out, lens = ctx.uncompress_batch_device(comp, ostride, sizes, count, 4096)
    ctx.sync()
    assert (lens.cpu().numpy() == 4096).all()
    assert bytes(out.cpu().numpy()) == pages.tobytes()
    # one corrupted page is flagged, the others are unaffected
    comp2 = comp.clone()
    comp2[ostride * 7 + 3] = 0xff
    comp2[ostride * 7 + 4] = 0xff
    out2, lens2 = ctx.uncompress_batch_device(comp2, ostride, sizes, count, 4096)
    ctx.sync()
    l2 = lens2.cpu().numpy()
    assert l2[7] == -1 or l2[7] == 4096          # 0xffffffff, unless the damage happened to stay decodable
    assert (np.delete(l2, 7) == 4096).all()


def _pack(streams, stride):
    """streams (list of bytes) at a fixed stride -> (uint8 array, int32 sizes)."""
    buf = np.zeros(stride * len(streams) + 64, dtype=np.uint8)
    for i, s in enumerate(streams):
        buf[i * stride:i * stride + len(s)] = np.frombuffer(s, dtype=np.uint8)
    return buf, np.array([len(s) for s in streams], dtype=np.int32)


def test_staged_page_kernel_malformed_and_fallback(ctx, oracle):
    """The page-shaped batch takes decode_staged_kernel (streams and pages staged in shared memory).
    Every malformed form must be flagged there too, a stream whose preamble promises more than the
    page capacity is rejected, empty slots are errors, and neighbours are unaffected."""
    import torch
    page = 4096
    text = corpus.fixture("alice29.txt")
    good = [oracle.compress(text[i * page:(i + 1) * page]) for i in range(24)]
    fox = oracle.compress(corpus.fixture("fox-in-socks.txt")[:page])
    bad = [
        b"",                                              # empty slot
        b"\x80\x80\x80\x80\x80\x01",                      # varint too long
        fox[:-1], fox[:len(fox) // 2],                    # truncated
        fox + b"\x00",                                    # trailing element without data
        b"\x0a" + bytes([0x01, 0x05]),                    # copy before any output
        b"\x08\x0cabcd" + bytes([0x01, 0x00]),            # offset 0
        b"\x04\x0cabcd\x00z",                             # more output than promised
        b"\x10\x0cabcd",                                  # less output than promised
        b"\x05" + bytes([61 << 2, 0xff, 0xff]),           # literal longer than the input
        oracle.compress(text[:page + 1]),                 # decodes to 4097 bytes: larger than the page capacity
        b"\x06\x0cabcd" + bytes([0x03 | (1 << 2)]) + (2).to_bytes(4, "little"),   # copy-4 form, valid: "abcdcd"
    ]
    streams = good[:8] + bad + good[8:]
    stride = oracle.max_compressed_length(page)
    buf, sizes = _pack(streams, stride)
    src = torch.from_numpy(buf).cuda()
    out, lens = ctx.uncompress_batch_device(src, stride, torch.from_numpy(sizes).cuda(), len(streams), page)
    ctx.sync()
    lens = lens.cpu().numpy()
    out = out.cpu().numpy()
    for i, s in enumerate(streams):
        if 8 <= i < 8 + len(bad) - 1:
            assert lens[i] == -1, (i, lens[i])
        elif i == 8 + len(bad) - 1:
            assert lens[i] == 6 and bytes(out[i * page:i * page + 6]) == b"abcdcd"
        else:
            j = i if i < 8 else i - len(bad)
            assert lens[i] == page and bytes(out[i * page:(i + 1) * page]) == text[j * page:(j + 1) * page], i


@pytest.mark.parametrize("page,skew", [(1024, 0), (4096, 3), (8192, 1), (100, 5), (16384, 0)])
def test_batch_page_sizes_and_alignments(ctx, oracle, page, skew):
    """Staged kernel up to 8 KiB pages, global-memory kernel above; source and destination unaligned."""
    import torch
    count = 64
    text = corpus.fixture("all-mtg-cards.txt")
    pages = [text[i * page:(i + 1) * page] for i in range(count)]
    stride = oracle.max_compressed_length(page)
    buf, sizes = _pack([oracle.compress(p) for p in pages], stride)
    src = torch.from_numpy(np.concatenate([np.zeros(skew, dtype=np.uint8), buf])).cuda()[skew:]
    dst = torch.zeros(page * count + skew + 16, dtype=torch.uint8, device="cuda")[skew:]
    out, lens = ctx.uncompress_batch_device(src, stride, torch.from_numpy(sizes).cuda(), count, page, dst=dst)
    ctx.sync()
    assert (lens.cpu().numpy() == page).all()
    assert bytes(out[:page * count].cpu().numpy()) == b"".join(pages)


# ---- one large raw stream: block-parallel path with on-device fallback (csrc/decode_par.cuh) ----
def _varint(n):
    out = bytearray()
    while n >= 128:
        out.append((n & 0x7f) | 0x80); n >>= 7
    out.append(n)
    return bytes(out)


@pytest.mark.parametrize("name,n", [("all-mtg-cards.txt", 3084856), ("alice29.txt", 152089 * 4 + 17),
                                    ("fireworks.jpeg", 123093 * 3), ("zeros", 5 << 20), ("mtg", 32 << 20)])
def test_large_stream_parallel_decode(ctx, oracle, name, n):
    if name == "zeros":
        data = bytes(n)
    elif name == "mtg":
        data = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n).tobytes()
    else:
        data = corpus.repeat_to(corpus.fixture(name), n).tobytes()
    comp = oracle.compress(data)
    assert len(comp) >= 128 << 10                      # takes the block-parallel path
    assert ctx.uncompress(comp) == data
    assert ctx.uncompress(ctx.compress(data, mode="fast")) == data


def test_large_stream_not_block_structured_falls_back(ctx, oracle):
    """Valid streams the fast path must refuse: a copy reaching across a 64 KiB boundary, an element
    straddling one, one huge literal.  The device-side fallback decodes them all the same."""
    base = corpus.fixture("all-mtg-cards.txt")[:400000]
    # (a) literal of everything, then copies that reach far back across block boundaries
    lit = base
    L = len(lit) - 1
    body = bytes([63 << 2]) + L.to_bytes(4, "little") + lit
    want = bytearray(lit)
    for off, ln in ((70000, 64), (200000, 33), (65536, 64), (5, 60)):
        body += bytes([3 | ((ln - 1) << 2)]) + off.to_bytes(4, "little")        # copy-4
        for _ in range(ln):
            want.append(want[-off])
    stream = _varint(len(want)) + body
    assert len(stream) >= 128 << 10
    assert ctx.uncompress(stream) == bytes(want)
    # (b) block-structured stream shifted by a 1000-byte literal in front: elements straddle boundaries
    tail = oracle.compress(base)
    hdr_len = len(_varint(len(base)))
    pre = b"x" * 1000
    stream = _varint(len(base) + 1000) + bytes([61 << 2]) + (999).to_bytes(2, "little") + pre + tail[hdr_len:]
    assert ctx.uncompress(stream) == pre + base


def test_large_stream_malformed_is_reported(ctx, oracle, pkg):
    data = corpus.fixture("all-mtg-cards.txt")[:1500000]
    good = oracle.compress(data)
    cases = [good[:-1], good[:len(good) // 2], good + b"\x00"]
    x = bytearray(good); x[300000] ^= 0xff; x[300001] ^= 0xff; x[300002] ^= 0xff
    cases.append(bytes(x))                                   # may stay decodable: then the bytes must differ or be flagged
    for i, s in enumerate(cases[:3]):
        with pytest.raises(pkg.SnappyB200Error):
            ctx.uncompress(s)
    try:
        got = ctx.uncompress(cases[3])
        assert len(got) == len(data)
    except pkg.SnappyB200Error:
        pass
    # wrong preamble: promises more than the elements deliver
    bad = _varint(len(data) + 5) + good[len(_varint(len(data))):]
    with pytest.raises(pkg.SnappyB200Error):
        ctx.uncompress(bad)


def test_large_stream_device_entry_1gib_class(ctx):
    """256 MiB through the device entries: compress, block-parallel uncompress, compare on the device."""
    import torch
    n = 256 << 20
    src = torch.from_numpy(corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)).cuda()
    comp, out_n = ctx.compress_device(src)
    ctx.sync()
    total = int(out_n.item())
    back = torch.zeros(n, dtype=torch.uint8, device="cuda")
    got = torch.zeros(1, dtype=torch.int64, device="cuda")
    ctx.uncompress_device(comp, total, back, got)
    ctx.sync()
    assert int(got.item()) == n and torch.equal(back, src)


def test_large_stream_mixed_content(ctx, oracle):
    """Text, incompressible and constant regions interleaved: long literals (which the speculative
    walks refuse) next to dense short elements, runs of the true chain spanning several segments."""
    rng = np.random.default_rng(7)
    text = corpus.fixture("all-mtg-cards.txt")
    parts = []
    for i in range(24):
        kind = i % 4
        ln = int(rng.integers(30000, 400000))
        if kind == 0:
            parts.append(text[(i * 77777) % 2000000:][:ln])
        elif kind == 1:
            parts.append(rng.integers(0, 256, size=ln, dtype=np.uint8).tobytes())
        elif kind == 2:
            parts.append(bytes([i]) * ln)
        else:
            parts.append(corpus.random_ascii(ln, seed=i).tobytes())
    data = b"".join(parts)
    for comp in (oracle.compress(data), ctx.compress(data, mode="fast")):
        assert len(comp) >= 128 << 10
        assert ctx.uncompress(comp) == data


@pytest.mark.parametrize("name", sorted(MAN.get("bad", {})))
def test_reference_corrupted_files_refused(ctx, pkg, name):
    """tests/snappy/testdata/baddata{1,2,3}.snappy (snappy_unittest.cc:585-600): the reference refuses them
    (IsValidCompressedBuffer / Uncompress false) and so must every decode path here: the host entry, the
    device entry and the batch entry (one warp per stream, a good stream beside the bad one)."""
    import torch
    data = corpus.fixture(name)
    ent = MAN["bad"][name]
    assert len(data) == ent["n"]
    with pytest.raises(pkg.SnappyB200Error) as ei:
        ctx.uncompress(data)
    assert ei.value.status == 1                                     # SNAPPY_B200_INVALID_INPUT
    # device entry: out_n = -1 (all ones)
    src = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    dst = torch.zeros(ent["announced_length"] + 64, dtype=torch.uint8, device="cuda")
    got = torch.zeros(1, dtype=torch.int64, device="cuda")
    ctx.uncompress_device(src, len(data), dst, got)
    ctx.sync()
    assert int(got.item()) == -1
    # batch entry (one warp per stream; a good stream beside it still decodes)
    good = ctx.compress(b"neighbour " * 50)
    stride = max(len(data), len(good)) + 16
    buf = torch.zeros(2 * stride, dtype=torch.uint8, device="cuda")
    buf[:len(data)] = src
    buf[stride:stride + len(good)] = torch.frombuffer(bytearray(good), dtype=torch.uint8).cuda()
    sizes = torch.tensor([len(data), len(good)], dtype=torch.int32, device="cuda")
    out, lens = ctx.uncompress_batch_device(buf, stride, sizes, 2, ent["announced_length"] + 64)
    ctx.sync()
    assert int(lens[0].item()) == -1 and int(lens[1].item()) == 500


def test_reference_zero_offset_and_four_byte_offset(ctx, pkg):
    """snappy_unittest.cc:945-959 (copy with offset 0 must fail) and :732-764 (a hand-emitted copy with a
    4-byte offset must decode)."""
    for s in (b"\x40\x12\x00\x00", b"\x05\x12\x00\x00"):
        with pytest.raises(pkg.SnappyB200Error):
            ctx.uncompress(s)
    f1, f2 = b"012345689abcdefghijklmnopqrstuvwxyz", b"some other string"
    n2 = 100000 // len(f2)
    src = f1 + f2 * n2
    length = 2 * len(f1) + n2 * len(f2)

    def varint(v):
        out = bytearray()
        while v >= 128:
            out.append((v & 0x7f) | 0x80); v >>= 7
        out.append(v)
        return bytes(out)

    def literal(b):
        return bytes([(len(b) - 1) << 2]) + b                       # both fragments are shorter than 60 bytes

    comp = varint(length) + literal(f1) + literal(f2) * n2
    comp += bytes([3 | ((len(f1) - 1) << 2)]) + len(src).to_bytes(4, "little")      # copy-4: offset = everything so far
    assert ctx.uncompress(comp) == src + f1


def test_rocc_uncompress_operands_are_checked(pkg):
    """the UNCOMPRESS command of the RoCC object must refuse operands outside its memory and a
    stream whose preamble announces more than fits behind rs2."""
    dev = pkg.RoccDevice(1 << 20)
    try:
        comp = pkg.default_context().compress(corpus.THISTLE)
        dev.load_mem(comp, 0)

        def run(rs1, length, rs2):
            dev.poke_cmd(dev.SET_LENGTH, length, 0)
            dev.step(1)
            dev.poke_cmd(dev.UNCOMPRESS, rs1, rs2)
            dev.step(1)
            return dev.peek_resp_data()
        assert run(0, len(comp), 4096) == 0 and dev.read_mem(4096, 130) == corpus.THISTLE
        assert run(0, len(comp), (1 << 20) + 5) != 0                 # destination outside the memory
        assert run((1 << 20) - 3, len(comp), 4096) != 0              # source runs off the end
        assert run(0, len(comp), (1 << 20) - 64) != 0                # 130 bytes do not fit behind rs2
        assert run(0, 1 << 30, 4096) != 0                            # poked length larger than the memory
        big = b"\xff\xff\xff\xff\x0f" + b"\x00" * 16                 # preamble announces 4 GiB - 1
        dev.load_mem(big, 0)
        assert run(0, len(big), 4096) != 0
    finally:
        dev.close()
