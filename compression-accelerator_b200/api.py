"""ctypes binding of libsnappy_b200.so (include/snappy_b200.h) and the host-side mirror of the
reference's compress interface.

Mirrors, by name and argument meaning:
  * ``compress(source, uncompressed_length, dest) -> compressed_length`` of the RoCC driver
    (reference tests/compression.c:20-27)                      -> :func:`rocc_compress`
  * ``snappy::Compress(input, n, &output)`` as driven by tests/swcompression.cc:45
                                                               -> :func:`compress`
  * ``snappy::MaxCompressedLength`` (tests/snappy/snappy.cc:99-121) -> :func:`max_compressed_length`

There is no CPU path: if the library or a CUDA device is missing these calls raise.
torch is used only for device memory and streams (Context.*_device take torch uint8 tensors).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SNAPPY_B200_LIB") or os.path.join(_HERE, "libsnappy_b200.so")   # override: A/B builds

EXACT, FAST = 0, 1
_MODES = {"exact": EXACT, "fast": FAST, "exact_serial": 2, "exact_window": 3, "exact_pipe_smem": 4, "exact_lane": 5,
          "exact_pipe": 6, "exact_gtab": 7, EXACT: EXACT, FAST: FAST, 2: 2, 3: 3, 4: 4, 5: 5, 6: 6, 7: 7}
OK, INVALID_INPUT, BUFFER_TOO_SMALL, CUDA_ERROR, NO_DEVICE = range(5)
BLOCK = 65536            # tests/snappy/snappy.h:196 kBlockSize


class SnappyB200Error(RuntimeError):
    def __init__(self, status, text):
        super().__init__("libsnappy_b200 status %d: %s" % (status, text))
        self.status = status


_lib = None

# name -> (restype, argtypes); every symbol include/snappy_b200.h declares
SYMBOLS = {
    "snappy_b200_max_compressed_length": (C.c_size_t, [C.c_size_t]),
    "snappy_b200_last_error": (C.c_char_p, []),
    "snappy_b200_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "snappy_b200_destroy": (None, [C.c_void_p]),
    "snappy_b200_stream": (C.c_void_p, [C.c_void_p]),
    "snappy_b200_compress": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                       C.POINTER(C.c_size_t), C.c_int]),
    "snappy_b200_compress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                              C.c_size_t, C.c_void_p, C.c_int]),
    "snappy_b200_compress_fragments_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t,
                                                        C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]),
    # no-allocation entry (no counterpart in the reference; static device scratch, inputs up to 1 MiB)
    "snappy_b200_noalloc_max_input": (C.c_size_t, []),
    "snappy_b200_compress_device_noalloc": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                                      C.c_void_p, C.c_void_p]),
    "snappy_b200_compress_batch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                                    C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p,
                                                    C.c_int]),
    "snappy_b200_compress_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                             C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]),
    "snappy_b200_uncompressed_length": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "snappy_b200_uncompress": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                         C.POINTER(C.c_size_t)]),
    "snappy_b200_uncompress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                                C.c_void_p]),
    "snappy_b200_uncompress_batch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                                      C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]),
    # Snappy framing format (tests/snappy/framing_format.txt)
    "snappy_b200_frame_max_length": (C.c_size_t, [C.c_size_t]),
    "snappy_b200_frame_compress": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                             C.POINTER(C.c_size_t), C.c_int]),
    "snappy_b200_frame_uncompress": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                               C.POINTER(C.c_size_t)]),
    "snappy_b200_frame_uncompressed_length": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "snappy_b200_frame_compress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                                    C.c_size_t, C.c_void_p, C.c_int]),
    "snappy_b200_frame_uncompress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                                      C.c_size_t, C.POINTER(C.c_size_t)]),
    "snappy_b200_crc32c_chunks_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]),
    "snappy_b200_varint32": (C.c_size_t, [C.c_void_p, C.c_uint32]),
    "snappy_b200_sync": (C.c_int, [C.c_void_p]),
    "snappy_b200_host_alloc": (C.c_void_p, [C.c_size_t]),
    "snappy_b200_host_free": (None, [C.c_void_p]),
    "snappy_b200_launch_count": (C.c_ulonglong, [C.c_void_p]),
    "snappy_b200_last_kernel_ms": (C.c_float, [C.c_void_p]),
    "snappy_b200_stage_ms": (C.c_float, [C.c_void_p, C.c_int]),
    "snappy_b200_debug_stats": (C.c_int, [C.c_void_p, C.c_int]),
    "snappy_b200_rocc_compress": (C.c_ulong, [C.c_void_p, C.c_int, C.c_void_p]),
    "snappy_b200_rocc_uncompress": (C.c_ulong, [C.c_void_p, C.c_int, C.c_void_p]),
    # multi-GPU exchange over peer memory (csrc/peer.cu)
    "snappy_b200_peer_handle_bytes": (C.c_size_t, []),
    "snappy_b200_peer_last_error": (C.c_char_p, []),
    "snappy_b200_peer_create": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_void_p,
                                          C.POINTER(C.c_void_p)]),
    "snappy_b200_peer_connect": (C.c_int, [C.c_void_p, C.c_void_p]),
    "snappy_b200_peer_gather": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_ulonglong, C.c_ulonglong]),
    "snappy_b200_peer_reuse": (C.c_int, [C.c_void_p, C.c_ulonglong]),
    "snappy_b200_peer_wait": (C.c_int, [C.c_void_p, C.c_ulonglong]),
    "snappy_b200_peer_result": (C.c_int, [C.c_void_p, C.c_ulonglong, C.c_ulonglong, C.POINTER(C.c_ulonglong),
                                          C.POINTER(C.c_ulonglong)]),
    "snappy_b200_peer_archive": (C.c_void_p, [C.c_void_p, C.c_int]),
    "snappy_b200_peer_stream": (C.c_void_p, [C.c_void_p]),
    "snappy_b200_peer_launch_count": (C.c_ulonglong, [C.c_void_p]),
    "snappy_b200_peer_destroy": (None, [C.c_void_p]),
    # include/b200_rocc.h: the accelerator's top-level IO (ScratchpadTest.scala:23-51)
    "b200_rocc_new": (C.c_void_p, [C.c_size_t, C.c_int]),
    "b200_rocc_free": (None, [C.c_void_p]),
    "b200_rocc_load_mem": (C.c_long, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64]),
    "b200_rocc_load_mem_file": (C.c_long, [C.c_void_p, C.c_char_p, C.c_uint64]),
    "b200_rocc_mem": (C.c_void_p, [C.c_void_p]),
    "b200_rocc_mem_bytes": (C.c_size_t, [C.c_void_p]),
    "b200_rocc_poke_cmd": (None, [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_int]),
    "b200_rocc_step": (None, [C.c_void_p, C.c_uint64]),
    "b200_rocc_peek_cmd_ready": (C.c_int, [C.c_void_p]),
    "b200_rocc_peek_busy": (C.c_int, [C.c_void_p]),
    "b200_rocc_peek_resp_valid": (C.c_int, [C.c_void_p]),
    "b200_rocc_peek_resp_data": (C.c_uint64, [C.c_void_p]),
    "b200_rocc_peek_busy_cycles": (C.c_uint64, [C.c_void_p]),
    "b200_rocc_peek_interrupt": (C.c_int, [C.c_void_p]),
}


def lib():
    """Load libsnappy_b200.so (fails loudly if it was not built: run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SnappyB200Error(NO_DEVICE, "%s not built (python -c 'import __graft_entry__ as g; "
                                  "g.build()')" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def _check(status):
    if status != OK:
        raise SnappyB200Error(status, lib().snappy_b200_last_error().decode())


def debug_stats(reset=True):
    """Diagnostics counters of the pipelined parser (zeros unless built with SNAPPY_B200_STATS)."""
    buf = (C.c_ulonglong * 16)()
    lib().snappy_b200_debug_stats(buf, 1 if reset else 0)
    return list(buf)


def max_compressed_length(n: int) -> int:
    return lib().snappy_b200_max_compressed_length(n)


def varint32(n: int) -> bytes:
    buf = C.create_string_buffer(8)
    k = lib().snappy_b200_varint32(buf, n)
    return buf.raw[:k]


def _addr(buf):
    """A ctypes argument for the bytes of a bytes / bytearray / numpy array / tensor / ctypes buffer, without
    copying.  For a bytearray the result is a ctypes array SHARING its memory (the caller passes it straight
    to the C call, which keeps it alive for the call's duration); for the others an integer address of memory
    the caller's own reference keeps alive."""
    if isinstance(buf, bytearray):
        return (C.c_char * len(buf)).from_buffer(buf)
    if isinstance(buf, bytes):
        return C.cast(C.c_char_p(buf), C.c_void_p).value
    if hasattr(buf, "ctypes"):
        return buf.ctypes.data
    if hasattr(buf, "data_ptr"):
        return buf.data_ptr()
    return C.addressof(buf)


class Context:
    """One CUDA stream plus device scratch (snappy_b200_ctx).  Not thread-safe; make one per
    host thread, as one RoCC unit serves one command at a time (CompressionAccelerator.scala:72)."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        _check(lib().snappy_b200_create(C.byref(h), device))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            lib().snappy_b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except TypeError:            # interpreter shutdown: module globals are already gone
            pass

    @property
    def stream(self) -> int:
        return lib().snappy_b200_stream(self._h) or 0

    @property
    def launches(self) -> int:
        return lib().snappy_b200_launch_count(self._h)

    @property
    def last_kernel_ms(self) -> float:
        return lib().snappy_b200_last_kernel_ms(self._h)

    # ---- ordering against torch's streams ------------------------------------------------
    # The library works on its own non-blocking stream.  Every entry that takes torch tensors first makes
    # that stream wait for torch's current stream (so a src still being produced by a torch op, or a dst
    # torch has just allocated, is ready) and afterwards makes torch's current stream wait for the
    # library's work: `out_n.item()`, slicing, or freeing a tensor (the caching allocator reuses a block in
    # the order of the stream it was allocated on) is then ordered behind the library's kernels without an
    # explicit sync().
    def _ext(self):
        import torch
        if getattr(self, "_ext_stream", None) is None:
            self._ext_stream = torch.cuda.ExternalStream(self.stream, device=torch.device("cuda", self.device))
        return self._ext_stream

    def _enter(self, *tensors):
        import torch
        cur = torch.cuda.current_stream(self.device)
        ext = self._ext()
        if cur.cuda_stream != ext.cuda_stream:
            ext.wait_stream(cur)

    def _exit(self):
        import torch
        cur = torch.cuda.current_stream(self.device)
        ext = self._ext()
        if cur.cuda_stream != ext.cuda_stream:
            cur.wait_stream(ext)

    def stage_ms(self, stage: int) -> float:
        """Device ms of stage 0 (parse kernel) / 1 (scan + gather) of the last device call."""
        return lib().snappy_b200_stage_ms(self._h, stage)

    def sync(self):
        _check(lib().snappy_b200_sync(self._h))

    # ---- host buffers -------------------------------------------------------------------
    def compress(self, data, mode="exact") -> bytes:
        """snappy::Compress(input, n, &output) (tests/swcompression.cc:45): host bytes in, the
        compressed stream out."""
        import numpy as np
        n = len(data)
        cap = max_compressed_length(n)
        out = np.empty(cap + 16, dtype=np.uint8)             # not zero-filled, and only the stream is copied out
        out_n = C.c_size_t(0)
        src = data if isinstance(data, bytes) else bytes(data)
        _check(lib().snappy_b200_compress(self._h, src, n, out.ctypes.data, cap, C.byref(out_n), _MODES[mode]))
        return out[:out_n.value].tobytes()

    def compress_into(self, src_addr: int, n: int, dst_addr: int, dst_cap: int, mode="exact") -> int:
        """Raw-pointer flavour (pinned or pageable host memory); returns the stream length."""
        out_n = C.c_size_t(0)
        _check(lib().snappy_b200_compress(self._h, src_addr, n, dst_addr, dst_cap, C.byref(out_n),
                                          _MODES[mode]))
        return out_n.value

    def compress_batch(self, buf, length: int, stride: int, count: int, mode="exact"):
        """`count` buffers of `length` bytes at `stride` -> (out uint8 array, out_stride, sizes);
        each buffer is its own Snappy stream (RawCompress per buffer, snappy.cc:1404-1414)."""
        import numpy as np
        arr = np.frombuffer(buf, dtype=np.uint8) if not hasattr(buf, "ctypes") else buf
        ostride = max_compressed_length(length)
        out = np.zeros(ostride * count + 16, dtype=np.uint8)
        sizes = np.zeros(count, dtype=np.uint32)
        _check(lib().snappy_b200_compress_batch(self._h, arr.ctypes.data, length, stride, count,
                                                out.ctypes.data, ostride, sizes.ctypes.data, _MODES[mode]))
        return out, ostride, sizes

    def uncompress(self, data: bytes) -> bytes:
        """snappy::Uncompress (tests/swcompression.cc:49): a raw stream in, the original bytes out;
        raises SnappyB200Error(INVALID_INPUT) for a malformed stream."""
        n = C.c_size_t(0)
        _check(lib().snappy_b200_uncompressed_length(data, len(data), C.byref(n)))
        import numpy as np
        out = np.empty(n.value + 16, dtype=np.uint8)
        got = C.c_size_t(0)
        _check(lib().snappy_b200_uncompress(self._h, data, len(data), out.ctypes.data, n.value, C.byref(got)))
        return out[:got.value].tobytes()

    def uncompress_device(self, src, n, dst, out_n):
        """Decode the n-byte raw stream in device tensor `src` into `dst`; asynchronous.  out_n (a
        1-element int64 device tensor) receives the decoded length, or -1 for a malformed stream.
        Streams of 128 KiB and more are decoded block-parallel when they are block structured."""
        self._enter(src, dst, out_n)
        _check(lib().snappy_b200_uncompress_device(self._h, src.data_ptr(), n, dst.data_ptr(), dst.numel(),
                                                   out_n.data_ptr()))
        self._exit()
        return dst, out_n

    def uncompress_batch_device(self, src, in_stride, sizes, count, out_stride, dst=None, out_lens=None):
        """`count` streams at src + i*in_stride (sizes[i] bytes, int32/uint32 device tensor) ->
        (dst, out_lens): stream i decoded at dst + i*out_stride; out_lens[i] = -1 (0xffffffff) if malformed."""
        import torch
        if dst is None:
            dst = torch.empty(out_stride * count, dtype=torch.uint8, device=src.device)
        if out_lens is None:
            out_lens = torch.zeros(count, dtype=torch.int32, device=src.device)
        self._enter(src, sizes, dst, out_lens)
        _check(lib().snappy_b200_uncompress_batch_device(self._h, src.data_ptr(), in_stride, sizes.data_ptr(),
                                                         count, dst.data_ptr(), out_stride, out_lens.data_ptr()))
        self._exit()
        return dst, out_lens

    # ---- Snappy framing format (tests/snappy/framing_format.txt) ------------------------
    def frame_compress(self, data, mode="exact") -> bytes:
        """Framed stream (stream identifier + CRC-checked chunks of 65,536 bytes) of `data`."""
        n = len(data)
        cap = lib().snappy_b200_frame_max_length(n)
        out = C.create_string_buffer(cap + 16)
        got = C.c_size_t(0)
        _check(lib().snappy_b200_frame_compress(self._h, _addr(data) if n else None, n, out, cap,
                                                C.byref(got), _MODES[mode]))
        return out.raw[:got.value]

    def frame_uncompress(self, data) -> bytes:
        n = len(data)
        ulen = C.c_size_t(0)
        _check(lib().snappy_b200_frame_uncompressed_length(_addr(data), n, C.byref(ulen)))
        out = C.create_string_buffer(ulen.value + 16)
        got = C.c_size_t(0)
        _check(lib().snappy_b200_frame_uncompress(self._h, _addr(data), n, out, ulen.value, C.byref(got)))
        return out.raw[:got.value]

    def frame_compress_device(self, src, dst=None, out_n=None, mode="exact"):
        """Device tensors; asynchronous.  Returns (dst, out_n) like compress_device."""
        import torch
        n = src.numel()
        cap = lib().snappy_b200_frame_max_length(n)
        if dst is None:
            dst = torch.empty(cap, dtype=torch.uint8, device=src.device)
        if out_n is None:
            out_n = torch.zeros(1, dtype=torch.int64, device=src.device)
        self._enter(src, dst, out_n)
        _check(lib().snappy_b200_frame_compress_device(self._h, src.data_ptr(), n, dst.data_ptr(), dst.numel(),
                                                       out_n.data_ptr(), _MODES[mode]))
        self._exit()
        return dst, out_n

    def frame_uncompress_device(self, src, n, dst):
        """Decode the n-byte framed stream in device tensor `src` into `dst`; synchronous (every
        chunk's CRC is checked).  Returns the number of bytes written."""
        got = C.c_size_t(0)
        self._enter(src, dst)
        _check(lib().snappy_b200_frame_uncompress_device(self._h, src.data_ptr(), n, dst.data_ptr(), dst.numel(),
                                                         C.byref(got)))
        self._exit()
        return got.value

    def crc32c_chunks_device(self, src, chunk_len=65536, out=None):
        """Masked CRC-32C of consecutive chunk_len-byte chunks of a device tensor -> int32 device tensor."""
        import torch
        n = src.numel()
        chunks = (n + chunk_len - 1) // chunk_len
        if out is None:
            out = torch.zeros(max(chunks, 1), dtype=torch.int32, device=src.device)
        self._enter(src, out)
        _check(lib().snappy_b200_crc32c_chunks_device(self._h, src.data_ptr(), n, chunk_len, out.data_ptr()))
        self._exit()
        return out[:chunks]

    # ---- device tensors (torch uint8, on this context's device) -------------------------
    def compress_device(self, src, dst=None, out_n=None, mode="exact", header=True):
        """Compress a device-resident stream.  Returns (dst, out_n): dst a uint8 device tensor of
        capacity max_compressed_length(n), out_n a 1-element int64 device tensor with the
        stream length.  Asynchronous on self.stream.  header=False compresses a run of whole
        fragments without the varint preamble (multi-GPU shards)."""
        import torch
        n = src.numel()
        cap = max_compressed_length(n)
        if dst is None:
            dst = torch.empty(cap, dtype=torch.uint8, device=src.device)
        if out_n is None:
            out_n = torch.zeros(1, dtype=torch.int64, device=src.device)
        fn = lib().snappy_b200_compress_device if header else lib().snappy_b200_compress_fragments_device
        self._enter(src, dst, out_n)
        _check(fn(self._h, src.data_ptr(), n, dst.data_ptr(), dst.numel(), out_n.data_ptr(), _MODES[mode]))
        self._exit()
        return dst, out_n

    def compress_batch_device(self, src, length, stride, count, dst=None, sizes=None, mode="exact"):
        import torch
        ostride = max_compressed_length(length)
        if dst is None:
            dst = torch.empty(ostride * count, dtype=torch.uint8, device=src.device)
        if sizes is None:
            sizes = torch.zeros(count, dtype=torch.int32, device=src.device)
        self._enter(src, dst, sizes)
        _check(lib().snappy_b200_compress_batch_device(self._h, src.data_ptr(), length, stride, count,
                                                       dst.data_ptr(), ostride, sizes.data_ptr(),
                                                       _MODES[mode]))
        self._exit()
        return dst, ostride, sizes


_default = None


def default_context() -> Context:
    global _default
    if _default is None:
        _default = Context(0)
    return _default


def compress(data, mode="exact") -> bytes:
    """snappy::Compress on the default context (device 0)."""
    return default_context().compress(data, mode)


def rocc_compress(source, uncompressed_length: int, dest) -> int:
    """compress(source, uncompressed_length, dest) of tests/compression.c:20-27: writes the stream
    into the caller-sized `dest` (a writable ctypes / numpy buffer) and returns its length
    (0 on failure)."""
    src = source if isinstance(source, bytes) else _addr(source)
    return int(lib().snappy_b200_rocc_compress(src, uncompressed_length, _addr(dest)))


def rocc_uncompress(source, compressed_length: int, dest) -> int:
    """uncompress(source, compressed_length, dest) of tests/compression.c:32-39: 0 on success."""
    src = source if isinstance(source, bytes) else _addr(source)
    return int(lib().snappy_b200_rocc_uncompress(src, compressed_length, _addr(dest)))


class RoccDevice:
    """ScratchpadTestModule.io as an object (src/main/scala/example/ScratchpadTest.scala:23-51):
    ``poke_cmd`` / ``step`` / ``peek_*`` follow the PeekPokeTester calls of
    src/test/scala/example/Test.scala:17-53; ``load_mem`` is loadMemoryFromFile (:38)."""
    COMPRESS, UNCOMPRESS, SET_LENGTH = 0, 1, 2          # tests/compression.c:13-15

    def __init__(self, mem_bytes: int = 0, device: int = 0):
        self._h = lib().b200_rocc_new(mem_bytes, device)
        if not self._h:
            raise SnappyB200Error(NO_DEVICE, lib().snappy_b200_last_error().decode() or "b200_rocc_new failed")

    def close(self):
        if getattr(self, "_h", None):
            lib().b200_rocc_free(self._h)
            self._h = None

    __del__ = close

    def load_mem(self, data: bytes, addr: int = 0) -> int:
        return lib().b200_rocc_load_mem(self._h, data, len(data), addr)

    def read_mem(self, addr: int, n: int) -> bytes:
        return C.string_at(lib().b200_rocc_mem(self._h) + addr, n)

    def poke_cmd(self, funct: int, rs1: int = 0, rs2: int = 0, valid: bool = True):
        lib().b200_rocc_poke_cmd(self._h, funct, rs1, rs2, 1 if valid else 0)

    def step(self, n: int = 1):
        lib().b200_rocc_step(self._h, n)

    def peek_resp_valid(self) -> int:
        return lib().b200_rocc_peek_resp_valid(self._h)

    def peek_resp_data(self) -> int:
        return lib().b200_rocc_peek_resp_data(self._h)

    def peek_busy_cycles(self) -> int:
        return lib().b200_rocc_peek_busy_cycles(self._h)

    def peek_busy(self) -> int:
        return lib().b200_rocc_peek_busy(self._h)

    def run_compress(self, length: int):
        """The tester's whole protocol (Test.scala:27-53) for data preloaded at address 0:
        returns (busy_cycles, compressed_length, dst_address)."""
        dst = -(-length // 8) * 8 + 8                      # Test.scala:35
        c0 = self.peek_busy_cycles()
        self.poke_cmd(self.SET_LENGTH, length)             # :27-30
        self.step(1)
        self.poke_cmd(self.COMPRESS, 0, dst)               # :33-36
        self.step(1)
        self.poke_cmd(self.COMPRESS, 0, dst, valid=False)  # :39
        timeout = 0
        while not self.peek_resp_valid() and timeout < 50 * max(length, 1):   # :42-47
            self.step(1)
            timeout += 1
        self.step(20)                                      # :50
        return self.peek_busy_cycles() - c0, self.peek_resp_data(), dst
