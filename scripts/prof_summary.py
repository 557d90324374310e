#!/usr/bin/env python
"""Turn gpurun_out/{launches.csv,prof_*.ncu-rep} into the small text summaries kept under profiles/.

    python scripts/prof_summary.py <tag>            # e.g. r01_v2
"""
import collections
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
KEYS = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__shared_mem_per_block_dynamic",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_per_inst_issued.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__inst_executed_pipe_lsu.sum")


def launches(tag):
    path = os.path.join(OUT, "launches.csv")
    if not os.path.exists(path):
        return
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h, data = rows[hi], rows[hi + 1:]
    ix = {k: i for i, k in enumerate(h)}
    agg = collections.OrderedDict()
    for r in data:
        if len(r) < len(h):
            continue
        v = float(r[ix["Metric Value"]].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ix["Metric Unit"]], 1.0)
        a = agg.setdefault(r[ix["Kernel Name"]], [0, 0.0, r[ix["Grid Size"]], r[ix["Block Size"]]])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(ROOT, "profiles", tag + "_launches.csv"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none, command: python bench.py --steps 2 --warmup 3\n")
        f.write("# per-launch times are cold-cache and serialised: read the SHARES\n")
        f.write("kernel,launches,total_ms,share_pct,grid_of_last,block\n")
        for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write('"%s",%d,%.3f,%.2f,"%s","%s"\n' % (k, a[0], a[1], 100 * a[1] / tot, a[2], a[3]))


def full(tag, rep, name):
    path = os.path.join(OUT, rep)
    if not os.path.exists(path):
        return
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, u = rows[0], rows[1]
    with open(os.path.join(ROOT, "profiles", "%s_%s_ncu_full.csv" % (tag, name)), "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on; one row block per captured launch\n")
        for d in rows[2:]:
            f.write("kernel,%s\n" % d[h.index("Kernel Name")])
            for i, k in enumerate(h):
                if k in KEYS:
                    f.write("%s,%s,%s\n" % (k, d[i], u[i]))


if __name__ == "__main__":
    tag = sys.argv[1]
    launches(tag)
    for rep in sorted(os.listdir(OUT)):
        if rep.startswith("prof_") and rep.endswith(".ncu-rep"):
            full(tag, rep, rep[5:-8])
