"""The no-allocation entry point (snappy_b200_compress_device_noalloc; no counterpart in the reference):
called through the C-ABI, its stream must equal the existing device
entry's (and so the reference Snappy's) byte for byte -- integer / byte work, tolerance zero -- on calls in a
row that share the static scratch array, larger input first so stale scratch would show.

Collected last on purpose; passes alone in a fresh process and after the
other tests."""
import ctypes as C

import numpy as np
import pytest

import corpus

pytestmark = pytest.mark.gpu


def _inputs():
    text = corpus.fixture("all-mtg-cards.txt")
    mixed = bytearray(text[:1 << 20])
    mixed[300000:360000] = bytes(corpus.random_ascii(60000, 21))       # a literal-heavy stretch
    mixed[700000:780000] = b"a" * 80000                                  # a long run (64-byte copy splitting)
    return [("1MiB-mixed", bytes(mixed)), ("300000-text", text[:300000]), ("65537", text[1000:1000 + 65537]),
            ("65536", text[:65536]), ("100", text[:100]), ("14", b"abcdabcdabcdab"), ("1", b"x"), ("0", b""),
            ("1MiB-mixed-again", bytes(mixed))]


def _noalloc(pkg, torch, data, dst, out_n):
    L = pkg.lib()
    n = len(data)
    src = torch.frombuffer(bytearray(data) if n else bytearray(1), dtype=torch.uint8).cuda()
    stream = torch.cuda.current_stream().cuda_stream
    st = L.snappy_b200_compress_device_noalloc(src.data_ptr(), n, dst.data_ptr(), dst.numel(), out_n.data_ptr(),
                                               C.c_void_p(stream))
    assert st == 0, L.snappy_b200_last_error().decode()
    total = int(out_n.item())                                            # same stream: ordered behind the kernels
    return src, dst[:total].cpu().numpy().tobytes()


def test_noalloc_equals_existing_entry_on_calls_in_a_row(pkg, ctx, oracle):
    import torch
    L = pkg.lib()
    cap = pkg.max_compressed_length(1 << 20)
    assert L.snappy_b200_noalloc_max_input() == 1 << 20
    dst = torch.empty(cap, dtype=torch.uint8, device="cuda")
    out_n = torch.zeros(1, dtype=torch.int64, device="cuda")
    _noalloc(pkg, torch, b"warm-up call: loads the module and resolves the scratch symbol", dst, out_n)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for name, data in _inputs():                                         # calls in a row on one shared scratch array
        src, got = _noalloc(pkg, torch, data, dst, out_n)
        want_dst, want_n = ctx.compress_device(src[:len(data)])         # the existing entry point, same device input
        ctx.sync()
        want = want_dst[:int(want_n.item())].cpu().numpy().tobytes()
        assert got == want, name
        assert got == oracle.compress(data), name                        # and so the reference Snappy's
        # the same call again, at once: same bytes (nothing of the previous call leaks through the scratch)
        _, again = _noalloc(pkg, torch, data, dst, out_n)
        assert again == got, name
        del want_dst, want_n, src
    torch.cuda.synchronize()
    # the entry point itself allocates nothing: a cudaMalloc of its ~2 MiB scratch per call would show here
    # (the test's own tensors are freed into torch's cache, which keeps the driver-level figure steady)
    leaked = free_before - torch.cuda.mem_get_info()[0]
    assert leaked < (64 << 20), "device memory shrank by %d bytes over %d calls" % (leaked, 2 * len(_inputs()))


def test_noalloc_rejects_what_does_not_fit(pkg):
    import torch
    L = pkg.lib()
    n = (1 << 20) + 1
    src = torch.zeros(n, dtype=torch.uint8, device="cuda")
    dst = torch.empty(pkg.max_compressed_length(n), dtype=torch.uint8, device="cuda")
    out_n = torch.zeros(1, dtype=torch.int64, device="cuda")
    st = L.snappy_b200_compress_device_noalloc(src.data_ptr(), n, dst.data_ptr(), dst.numel(), out_n.data_ptr(), None)
    assert st == 1                                                        # SNAPPY_B200_INVALID_INPUT
    msg = L.snappy_b200_last_error().decode()
    assert "1048576" in msg and "maximum" in msg, msg
    # capacity is checked too, and null pointers
    st = L.snappy_b200_compress_device_noalloc(src.data_ptr(), 1000, dst.data_ptr(), 100, out_n.data_ptr(), None)
    assert st == 2                                                        # SNAPPY_B200_BUFFER_TOO_SMALL
    assert L.snappy_b200_compress_device_noalloc(src.data_ptr(), 1000, None, 1 << 20, out_n.data_ptr(), None) == 1
