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

    def fast_compress(self, data: bytes) -> bytes:
        """Sequential model of libsnappy_b200's FAST mode (oracle/snappy_port.c, port_fast_compress):
        defines that mode's stream; not a restatement of the reference."""
        out = C.create_string_buffer(self.max_compressed_length(len(data)) + 16)
        n = self.lib.port_fast_compress(data, len(data), out)
        return out.raw[:n]

    def uncompress(self, data: bytes, cap: int) -> bytes:
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
