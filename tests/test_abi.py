"""CPU-only: the C-ABI library loads and exports every symbol include/snappy_b200.h declares;
argument validation that needs no device; the product never falls back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            text = open(os.path.join(ROOT, "include", fn)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            names |= set(re.findall(r"\b((?:snappy_b200|b200_rocc)_\w+)\s*\(", text))
    return sorted(names)


def test_header_symbols_all_exported(pkg):
    L = C.CDLL(pkg.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(L, s), "libsnappy_b200.so does not export %s" % s
    # and the Python binding table covers the compress header completely
    hdr = [s for s in syms if s.startswith("snappy_b200_")]
    assert set(hdr) <= set(pkg.SYMBOLS), set(hdr) - set(pkg.SYMBOLS)


def test_max_compressed_length_and_varint(pkg):
    # tests/snappy/snappy.cc:99-121 ; stubs-internal.h:552-578
    for n in (0, 1, 4096, 65536, 3084856, (1 << 30)):
        assert pkg.max_compressed_length(n) == 32 + n + n // 6
    assert pkg.max_compressed_length(65536) == 76490
    assert pkg.varint32(0) == b"\x00" and pkg.varint32(4096) == b"\x80\x20"
    assert pkg.varint32(130) == b"\x82\x01" and pkg.varint32(3084856).hex() == "b8a4bc01"
    assert len(pkg.varint32(1 << 30)) == 5


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the product raises; it never routes to a CPU implementation."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("device present")
    with pytest.raises(pkg.SnappyB200Error) as ei:
        pkg.Context(0)
    assert ei.value.status == 4
    buf = C.create_string_buffer(256)
    assert pkg.rocc_compress(b"abc", 3, buf) == 0          # RoCC mirror: 0 = failure


def test_peer_exchange_argument_checks(pkg):
    """The peer-memory exchange entries reject bad arguments before touching a device."""
    L = pkg.lib()
    assert L.snappy_b200_peer_handle_bytes() == 128          # two cudaIpcMemHandle_t
    h = C.c_void_p()
    buf = C.create_string_buffer(128)
    assert L.snappy_b200_peer_create(None, 0, 0, 2, 0, 1 << 20, buf, C.byref(h)) == 1     # null context
    assert b"peer_create" in L.snappy_b200_peer_last_error()
    assert L.snappy_b200_peer_connect(None, buf) == 1
    assert L.snappy_b200_peer_gather(None, None, None, 0, 0) == 1
    assert L.snappy_b200_peer_archive(None, 0) is None
    L.snappy_b200_peer_destroy(None)                          # no-op


def test_product_does_not_reference_oracle():
    for dp, _, fns in os.walk(os.path.join(ROOT, "compression-accelerator_b200")):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".c", ".scala")) or fn == "Makefile":
                text = open(os.path.join(dp, fn), errors="ignore").read()
                assert "oracle" not in text.lower() or fn == "__init__.py" and False, os.path.join(dp, fn)
