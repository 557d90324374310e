#!/usr/bin/env python
"""Per-kernel count of the SASS mnemonics that prove what a kernel uses (TMA bulk copies UBLKCP / UBLKPF,
mbarrier SYNCS, 128-bit atomics ATOMG.*.128, 256-bit stores STG.*.256, warp match MATCH, ...):
    python scripts/sass_mnemonics.py > profiles/r02_sass_mnemonics.txt
Static counts from `cuobjdump -sass compression-accelerator_b200/libsnappy_b200.so` (built in-tree)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "compression-accelerator_b200", "libsnappy_b200.so")
WATCH = ("UBLKCP", "UBLKPF", "UTMALDG", "SYNCS", "ATOMG", "ATOMS", "RED", "MATCH", "REDUX", "SHFL", "VOTE",
         "LDG", "STG", "LDS", "STS", "CCTL", "NANOSLEEP", "BAR", "HMMA", "UTC")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    per, cur = collections.OrderedDict(), None
    for ln in txt.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = per.setdefault(re.sub(r"\(.*", "", name), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", ln)
        if m and cur is not None:
            op = m.group(1)
            cur["total"] += 1
            for w in WATCH:
                if op.startswith(w):
                    key = w
                    if w in ("ATOMG", "STG", "LDG") and (".128" in op or ".256" in op):
                        key = w + (".256" if ".256" in op else ".128")
                    cur[key] += 1
                    break
    print("# static SASS mnemonic counts per kernel of libsnappy_b200.so (sm_100a); scripts/sass_mnemonics.py")
    for name, c in per.items():
        if name.startswith("void "):
            name = name[5:]
        rest = ", ".join("%s %d" % (k, v) for k, v in sorted(c.items()) if k != "total")
        print("%-48s %5d instr: %s" % (name[:48], c["total"], rest))


if __name__ == "__main__":
    main()
