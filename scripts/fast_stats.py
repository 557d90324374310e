#!/usr/bin/env python
"""Per-role cycle breakdown of the fast-mode kernel (library built with -DSNAPPY_B200_STATS)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, corpus
import __graft_entry__ as g
pkg = g.load_package()
n = int(float(sys.argv[1]) * (1 << 20)) if len(sys.argv) > 1 else 256 << 20
host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n)
ctx = pkg.Context(0)
src = torch.from_numpy(host).cuda()
ctx.compress_device(src, mode="fast"); ctx.sync(); pkg.debug_stats()
dst, out_n = ctx.compress_device(src, mode="fast"); ctx.sync()
st = pkg.debug_stats()
f = st[9] or 1
win = n / 32 / f
print("fragments %d  parse %.3f ms" % (f, ctx.stage_ms(0)))
print("per fragment (cycles): table warp total %.0f (waiting for chaser %.0f) | chaser total %.0f (waiting for records %.0f, windows visited %.0f)" % (
    st[0] / f, st[1] / f, st[7] / f, st[6] / f, st[8] / f))
print("per window (cycles, per evaluator warp): wait t_done %.0f, evaluate incl wait %.0f, wait answer %.0f, emit incl wait %.0f" % (
    st[2] / f / win, st[3] / f / win, st[4] / f / win, st[5] / f / win))
print("chaser per visited window (cycles): first shuffle %.0f, answer store %.0f, whole iteration %.0f" % (
    st[10] / max(st[8], 1), st[11] / max(st[8], 1), st[7] / max(st[8], 1)))
