#!/usr/bin/env python
"""Small driver for profiling: compress `--mib` MiB of a corpus on the device `--reps` times and
print per-stage device times.  Used under ncu (see profiles/README.md)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=float, default=18.5)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--mode", default="exact")
    ap.add_argument("--data", default="mtg", choices=["mtg", "random", "zeros", "alice"])
    ap.add_argument("--batch", type=int, default=0, help="if >0: batch path with buffers of this many bytes")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--decode", action="store_true", help="batch path: also time the device decompressor")
    a = ap.parse_args()
    import numpy as np
    import torch
    import corpus
    import __graft_entry__ as g
    pkg = g.load_package()
    n = int(a.mib * (1 << 20))
    if a.data == "mtg":
        host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
    elif a.data == "alice":
        host = corpus.repeat_to(corpus.fixture("alice29.txt"), n)
    elif a.data == "random":
        host = corpus.random_ascii(n)
    else:
        host = np.full(n, 97, dtype=np.uint8)
    ctx = pkg.Context(0)
    src = torch.from_numpy(host).cuda()
    if a.batch:
        count = n // a.batch
        for _ in range(a.reps):
            dst, ostride, sizes = ctx.compress_batch_device(src, a.batch, a.batch, count, mode=a.mode)
            ctx.sync()
            ms, ms1 = ctx.stage_ms(0), ctx.stage_ms(1)
            print("batch %d x %d B: parse %.3f ms  size+emit %.3f ms  %.2f GB/s in, out/in %.4f" % (
                count, a.batch, ms, ms1, count * a.batch / (ms + ms1) / 1e6, float(sizes.sum().item()) / (count * a.batch)))
        if a.decode:
            import time
            back, lens = ctx.uncompress_batch_device(dst, ostride, sizes, count, a.batch)
            ctx.sync()
            t0 = time.perf_counter()
            for _ in range(a.reps):
                ctx.uncompress_batch_device(dst, ostride, sizes, count, a.batch, dst=back, out_lens=lens)
            ctx.sync()
            ms = (time.perf_counter() - t0) / a.reps * 1e3
            comp = float(sizes.sum().item())
            print("decode %d x %d B: %.3f ms  %.2f GB/s out, %.2f GB/s in+out; identical: %s" % (
                count, a.batch, ms, count * a.batch / ms / 1e6, (count * a.batch + comp) / ms / 1e6,
                bool(torch.equal(back[:count * a.batch], src[:count * a.batch])) and bool((lens == a.batch).all())))
        return
    for _ in range(a.reps):
        dst, out_n = ctx.compress_device(src, mode=a.mode)
        ctx.sync()
        p, s = ctx.stage_ms(0), ctx.stage_ms(1)
        tot = int(out_n.item())
        print("%s %s %.1f MiB: parse %.3f ms  gather %.3f ms  %.2f GB/s in, out/in %.4f" % (
            a.data, a.mode, n / 2**20, p, s, n / (p + s) / 1e6, tot / n))
    if a.decode:
        import time
        back = torch.zeros(n, dtype=torch.uint8, device="cuda")
        got = torch.zeros(1, dtype=torch.int64, device="cuda")
        ctx.uncompress_device(dst, tot, back, got)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(a.reps):
            ctx.uncompress_device(dst, tot, back, got)
        ctx.sync()
        ms = (time.perf_counter() - t0) / a.reps * 1e3
        print("decode one stream of %.1f MiB: %.3f ms  %.2f GB/s out, %.2f GB/s in+out; identical: %s" % (
            n / 2**20, ms, n / ms / 1e6, (n + tot) / ms / 1e6, bool(int(got.item()) == n and torch.equal(back, src))))
    st = pkg.debug_stats()
    if st[0]:
        f = st[0]
        print('stats per fragment: cycles %.0f wait %.0f windows %.1f stale %.1f slow %.1f (hits %.1f, cycles %.0f) serial %.2f seqs %.0f slow-by-cause stale %.0f peer %.0f' % (
            st[1] / f, st[2] / f, st[3] / f, st[4] / f, st[5] / f, st[8] / f, st[6] / f, st[7] / f, st[9] / f, st[10] / f, st[11] / f))
    if st[0]:
        print('  phases per window: entry(wait+load+validate) %.0f  sequential %.0f (of which chase %.0f)' % (st[12] / st[3], st[13] / st[3], st[14] / st[3]))
    if a.check:
        import oracle_lib
        chk = oracle_lib.checker()
        comp = bytes(dst[:tot].cpu().numpy())
        if a.mode == "exact":
            want = chk.compress(host.tobytes())
            print("exact parity:", comp == want)
        print("roundtrip:", chk.uncompress(comp, n) == host.tobytes())


if __name__ == "__main__":
    main()
