"""ctypes loaders for the CPU checkers under oracle/ (TEST INFRASTRUCTURE ONLY).

``port`` = oracle/libsnappy_port.so, our C restatement (oracle/snappy_port.c).
``ref``  = oracle/_ref/libsnappy_ref.so, the reference's own tests/snappy sources compiled
           unmodified by oracle/Makefile (prebuilt file travels to the GPU box).

Nothing under compression-accelerator_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE, "libsnappy_port.so")
REF_SO = os.path.join(ORACLE, "_ref", "libsnappy_ref.so")


def build():
    """(Re)build the checkers; the reference build is skipped when /root/reference is absent."""
    subprocess.run(["make", "-s", "-C", ORACLE], check=True, stdout=subprocess.DEVNULL)


class _Lib:
    def __init__(self, path, prefix):
        self.path, self.prefix = path, prefix
        self.lib = C.CDLL(path)
        f = getattr(self.lib, prefix + "_compress")
        f.restype, f.argtypes = C.c_size_t, [C.c_char_p, C.c_size_t, C.c_void_p]
        f = getattr(self.lib, prefix + "_max_compressed_length")
        f.restype, f.argtypes = C.c_size_t, [C.c_size_t]
        f = getattr(self.lib, prefix + "_bench")
        f.restype = C.c_double
        f.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        f = getattr(self.lib, prefix + "_compress_batch")
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]

    def max_compressed_length(self, n):
        return getattr(self.lib, self.prefix + "_max_compressed_length")(n)

    def compress(self, data: bytes) -> bytes:
        out = C.create_string_buffer(self.max_compressed_length(len(data)) + 16)
        n = getattr(self.lib, self.prefix + "_compress")(data, len(data), out)
        return out.raw[:n]

    def compress_batch(self, buf, length, stride, count):
        """Each of `count` buffers of `length` bytes -> own stream; returns list of bytes."""
        import numpy as np
        arr = np.frombuffer(buf, dtype=np.uint8)
        ostride = self.max_compressed_length(length)
        out = np.zeros(ostride * count + 16, dtype=np.uint8)
        sizes = np.zeros(count, dtype=np.uint32)
        getattr(self.lib, self.prefix + "_compress_batch")(
            arr.ctypes.data, length, stride, count, out.ctypes.data, ostride, sizes.ctypes.data)
        return out, ostride, sizes

    def bench(self, arr, slice_bytes, threads, reps):
        """Time `reps` passes over numpy uint8 array `arr`; returns (seconds, produced_bytes)."""
        produced = C.c_size_t(0)
        secs = getattr(self.lib, self.prefix + "_bench")(
            arr.ctypes.data, arr.size, slice_bytes, threads, reps, C.byref(produced))
        return secs, produced.value


class Port(_Lib):
    def __init__(self):
        super().__init__(PORT_SO, "port")
        self.lib.port_uncompress.restype = C.c_int
        self.lib.port_uncompress.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                             C.POINTER(C.c_size_t)]
        self.lib.port_table_entries.restype = C.c_uint32
        self.lib.port_table_entries.argtypes = [C.c_uint32]
        self.lib.port_fast_compress.restype = C.c_size_t
        self.lib.port_fast_compress.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p]

    def crc32c(self, data: bytes) -> int:
        """CRC-32C (RFC 3720 12.1), oracle/snappy_port.c port_crc32c."""
        self.lib.port_crc32c.restype = C.c_uint32
        self.lib.port_crc32c.argtypes = [C.c_char_p, C.c_size_t]
        return self.lib.port_crc32c(data, len(data))

    def crc32c_masked(self, data: bytes) -> int:
        self.lib.port_crc32c_masked.restype = C.c_uint32
        self.lib.port_crc32c_masked.argtypes = [C.c_char_p, C.c_size_t]
        return self.lib.port_crc32c_masked(data, len(data))

    def fast_compress(self, data: bytes) -> bytes:
        """Sequential model of libsnappy_b200's FAST mode (oracle/snappy_port.c, port_fast_compress):
        defines that mode's stream; not a restatement of the reference."""
        out = C.create_string_buffer(self.max_compressed_length(len(data)) + 16)
        n = self.lib.port_fast_compress(data, len(data), out)
        return out.raw[:n]

    def uncompressed_length(self, data: bytes) -> int:
        """The varint32 preamble (snappy.cc:864-, GetUncompressedLength); ValueError when malformed."""
        v = shift = 0
        for i, b in enumerate(data[:5]):
            if shift == 28 and (b & 0x7f) > 15:
                break
            v |= (b & 0x7f) << shift
            if b < 128:
                return v
            shift += 7
        raise ValueError("bad header")

    def uncompress(self, data: bytes, cap: int = None) -> bytes:
        if cap is None:
            cap = self.uncompressed_length(data)
        out = C.create_string_buffer(cap + 16)
        n = C.c_size_t(0)
        rc = self.lib.port_uncompress(data, len(data), out, cap, C.byref(n))
        if rc != 0:
            raise ValueError("port_uncompress rc=%d" % rc)
        return out.raw[:n.value]


class Ref(_Lib):
    def __init__(self):
        super().__init__(REF_SO, "ref")
        self.lib.ref_uncompress.restype = C.c_int
        self.lib.ref_uncompress.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p]
        self.lib.ref_uncompressed_length.restype = C.c_int
        self.lib.ref_uncompressed_length.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        self.lib.ref_is_valid.restype = C.c_int
        self.lib.ref_is_valid.argtypes = [C.c_char_p, C.c_size_t]

    def uncompressed_length(self, data: bytes) -> int:
        n = C.c_size_t(0)
        if self.lib.ref_uncompressed_length(data, len(data), C.byref(n)) != 0:
            raise ValueError("bad header")
        return n.value

    def is_valid(self, data: bytes) -> bool:
        return bool(self.lib.ref_is_valid(data, len(data)))

    def uncompress(self, data: bytes, cap: int = None) -> bytes:
        n = self.uncompressed_length(data)
        out = C.create_string_buffer(n + 16)
        if self.lib.ref_uncompress(data, len(data), out) != 0:
            raise ValueError("ref_uncompress failed")
        return out.raw[:n]


_cache = {}


def port() -> Port:
    if "port" not in _cache:
        if not os.path.exists(PORT_SO):
            build()
        _cache["port"] = Port()
    return _cache["port"]


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def ref() -> Ref:
    if "ref" not in _cache:
        if not have_ref():
            build()
        _cache["ref"] = Ref()
    return _cache["ref"]


def checker():
    """The strongest checker available: the compiled reference if present, else the port."""
    return ref() if have_ref() else port()


# ---- Snappy framing format (tests/snappy/framing_format.txt), CPU restatement ------------------
STREAM_ID = bytes([0xff, 0x06, 0x00, 0x00, 0x73, 0x4e, 0x61, 0x50, 0x70, 0x59])   # :68-75
FRAME_CHUNK = 65536                                                               # :88-91


def frame_compress(data: bytes, compress=None) -> bytes:
    """Writer with the library's policy: identifier, then one compressed chunk (type 0x00) per 65,536
    input bytes: 3-byte LE length of (CRC + stream), masked CRC-32C of the uncompressed bytes
    (:38-57, :80-86), the chunk's own Snappy stream."""
    compress = compress or checker().compress
    out = bytearray(STREAM_ID)
    for i in range(0, len(data), FRAME_CHUNK):
        chunk = data[i:i + FRAME_CHUNK]
        body = compress(chunk)
        ln = 4 + len(body)
        out += bytes([0x00, ln & 0xff, (ln >> 8) & 0xff, (ln >> 16) & 0xff])
        out += port().crc32c_masked(chunk).to_bytes(4, "little")
        out += body
    return bytes(out)


def frame_uncompress(data: bytes) -> bytes:
    """Reader: every chunk type of the specification; raises ValueError on malformed input."""
    if data[:10] != STREAM_ID:
        raise ValueError("no stream identifier")
    out = bytearray()
    pos = 0
    while pos < len(data):
        if pos + 4 > len(data):
            raise ValueError("truncated header")
        typ = data[pos]
        ln = int.from_bytes(data[pos + 1:pos + 4], "little")
        body = data[pos + 4:pos + 4 + ln]
        if len(body) != ln:
            raise ValueError("truncated chunk")
        if typ == 0xff:
            if data[pos:pos + 10] != STREAM_ID:
                raise ValueError("bad identifier")
        elif typ in (0x00, 0x01):
            if ln < 4:
                raise ValueError("short chunk")
            crc = int.from_bytes(body[:4], "little")
            raw = port().uncompress(bytes(body[4:]), FRAME_CHUNK) if typ == 0x00 else bytes(body[4:])
            if len(raw) > FRAME_CHUNK or port().crc32c_masked(raw) != crc:
                raise ValueError("bad chunk")
            out += raw
        elif 0x02 <= typ <= 0x7f:
            raise ValueError("reserved unskippable chunk")
        pos += 4 + ln
    return bytes(out)
