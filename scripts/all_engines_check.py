#!/usr/bin/env python
"""Every engine and entry point once, on small inputs with tight strides, each result checked against the oracle
(written as a driver for `compute-sanitizer --tool memcheck`; that tool is closed on this pool, so it runs
plain: tests/test_all_engines.py runs it in a fresh process).  SNAPPY_B200_FRAME_FUSED=1 forces the fused
framed reader for the small framed stream."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("SNAPPY_B200_FRAME_FUSED", "1")
import numpy as np   # noqa: E402
import torch         # noqa: E402
import corpus        # noqa: E402
import oracle_lib    # noqa: E402
import __graft_entry__ as g   # noqa: E402

pkg = g.load_package()
ctx = pkg.Context(0)
chk = oracle_lib.checker()
text = corpus.fixture("all-mtg-cards.txt")
cases = [text[:65536 * 3 + 4321], corpus.structured_random(150001, 5), corpus.fixture("fireworks.jpeg"), b"a" * 70000, text[:15], b""]
for mode in ("exact", "exact_lane", "exact_pipe", "exact_gtab", "exact_window", "fast"):
    for data in cases:
        # exact-size device tensor: the allocation ends where the data ends
        src = torch.frombuffer(bytearray(data) if data else bytearray(1), dtype=torch.uint8).cuda()[:len(data)].clone()
        dst, out_n = ctx.compress_device(src, mode=mode)
        n = int(out_n.item())
        got = dst[:n].cpu().numpy().tobytes()
        if mode == "fast":
            assert chk.uncompress(got, len(data)) == data, (mode, len(data))
        else:
            assert got == chk.compress(data), (mode, len(data))
print("device entry: ok")
for length, count in ((10, 300), (100, 300), (512, 200), (1000, 64), (4096, 64), (5000, 1800), (20000, 1800)):
    stride = length                                               # tight: the last buffer ends the allocation
    base = (text * 13)[:stride * count]
    assert len(base) == stride * count
    buf = np.frombuffer(base, dtype=np.uint8).copy()
    if length == 4096:
        buf[:4096 * 8] = corpus.random_ascii(4096 * 8, 3)         # one-literal pages: emitted by the parse kernel
    for mode in ("exact", "exact_gtab"):
        out, ostride, sizes = ctx.compress_batch(buf, length, stride, count, mode=mode)
        for i in (0, 1, count // 2, count - 1):
            assert bytes(out[i * ostride:i * ostride + sizes[i]]) == chk.compress(bytes(buf[i * stride:i * stride + length])), (mode, length, i)
print("batch entry: ok")
data = text[:65536 * 5 + 77]
framed = ctx.frame_compress(data)
assert framed == oracle_lib.frame_compress(data, chk.compress) and ctx.frame_uncompress(framed) == data
assert ctx.uncompress(chk.compress(data)) == data
pages = np.frombuffer(text[:4096 * 64], dtype=np.uint8).copy()
out, ostride, sizes = ctx.compress_batch(pages, 4096, 4096, 64)
d_out = torch.from_numpy(out).cuda()
d_sizes = torch.from_numpy(sizes.astype(np.int32)).cuda()
back, lens = ctx.uncompress_batch_device(d_out, ostride, d_sizes, 64, 4096)
assert bytes(back.cpu().numpy()[:4096 * 64]) == pages.tobytes()
print("framing / decode: ok")
L = pkg.lib()
src = torch.frombuffer(bytearray(text[:300001]), dtype=torch.uint8).cuda()
dst = torch.empty(pkg.max_compressed_length(300001), dtype=torch.uint8, device="cuda")
out_n = torch.zeros(1, dtype=torch.int64, device="cuda")
import ctypes as C   # noqa: E402
assert L.snappy_b200_compress_device_noalloc(src.data_ptr(), 300001, dst.data_ptr(), dst.numel(), out_n.data_ptr(),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)) == 0
assert dst[:int(out_n.item())].cpu().numpy().tobytes() == chk.compress(text[:300001])
host = ctx.compress(text * 9)                                     # 27 MiB: the chunked host entry
assert host == chk.compress(text * 9)
print("noalloc / host entry: ok")
