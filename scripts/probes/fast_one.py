import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package(); ctx = pkg.Context(0)
per, n = int(sys.argv[1]), int(sys.argv[2])
base = bytes(np.random.default_rng(1).integers(0, 256, size=per, dtype=np.uint8))
data = (base * (n // per + 2))[:n]
got = ctx.compress(data, mode="fast")
print("ok", len(got))
