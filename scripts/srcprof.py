#!/usr/bin/env python
"""Per-source-line profile: merge an `ncu --page source --csv` SASS dump with nvdisasm line info.

    ncu -i prof.ncu-rep --page source --csv > sass.csv
    python scripts/srcprof.py sass.csv <kernel-name-substring> [top N]

Needs the library built from the same sources as the profiled one (extracts its cubin)."""
import csv, os, re, subprocess, sys, tempfile, collections

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "compression-accelerator_b200", "libsnappy_b200.so")


def sass_lines(kernel_sub):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", SO], cwd=d, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    cub = [f for f in os.listdir(d) if f.startswith("kernels.") and f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
    out, cur, infn, in_block = [], ("?", 0), False, False
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            infn = kernel_sub in m.group(1)
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            if not in_block:                      # first line of a block = innermost inlined frame
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
            in_block = True
            continue
        in_block = False
        m = re.match(r"\s+(/\*[0-9a-f]+\*/)?\s+([A-Z@!][^;]*);", ln)
        if m and not ln.strip().startswith("."):
            out.append((cur, m.group(2).strip()))
    return out


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    h = rows[1]; data = rows[2:]
    ix = {k: i for i, k in enumerate(h)}
    sl = sass_lines(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    if len(sl) != len(data):
        print("warning: %d SASS lines from nvdisasm vs %d in the ncu dump" % (len(sl), len(data)))
    agg = collections.defaultdict(lambda: [0, 0, 0])
    tot = 0
    for (loc, _), r in zip(sl, data):
        s = int(r[ix['# Samples']] or 0); n = int(r[ix['Instructions Executed']] or 0)
        agg[loc][0] += s; agg[loc][1] += n; agg[loc][2] += 1
        tot += s
    print("total samples", tot)
    src = {}
    for (f, l), (s, n, k) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        if f not in src:
            p = os.path.join(ROOT, "compression-accelerator_b200", "csrc", f)
            src[f] = open(p).read().splitlines() if os.path.exists(p) else []
        text = src[f][l - 1].strip()[:90] if 0 < l <= len(src[f]) else ""
        print("%5.1f%% %7d smp %11d dyn %3d sass  %s:%d  %s" % (100.0 * s / max(tot, 1), s, n, k, f, l, text))


if __name__ == "__main__":
    main()
