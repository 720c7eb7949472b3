#!/usr/bin/env python
"""Condense ncu outputs brought back in gpurun_out/ into small text files for profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches.csv  > profiles/rNN_launches.txt
    python tools/ncu_summary.py full gpurun_out/prof.ncu-rep      > profiles/rNN_<kernel>_full.txt
"""
import csv
import re
import subprocess
import sys
from collections import OrderedDict

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"lts__t_sector_hit_rate\.pct|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|lts__t_bytes\.sum|"
    r"l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__t_sectors_pipe_lsu_mem_global_op_(ld|st|atom|red)\.sum|l1tex__t_requests_pipe_lsu_mem_global_op_(ld|st|atom|red)\.sum|"
    r"sm__warps_active\.avg\.pct_of_peak_sustained_active|launch__(registers_per_thread|grid_size|block_size|"
    r"shared_mem_per_block_dynamic|occupancy_limit_\w+|waves_per_multiprocessor)|smsp__cycles_active\.avg|"
    r"sm__inst_executed\.sum|smsp__average_warp\w*issue_stalled_\w+_per_issue_active\.ratio|"
    r"smsp__average_warps_issue_stalled_\w+|sm__cycles_elapsed\.max|smsp__inst_executed\.avg\.per_cycle_active)$")


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    ig, ib = hdr.index("Grid Size"), hdr.index("Block Size")
    agg = OrderedDict()
    for r in rows:
        name = re.sub(r"\(.*", "", r[ik])[:90]
        a = agg.setdefault(name, [0, 0.0, r[ig], r[ib]])
        a[0] += 1
        a[1] += float(r[iv].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    print(f"# {len(rows)} launches, {tot/1e3:.1f} us total device time (ncu: cold-cache, serialised -> compare shares)")
    print(f"{'kernel':92s} {'n':>4s} {'avg_us':>9s} {'share':>7s}  grid block")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:92s} {a[0]:4d} {a[1]/a[0]/1e3:9.2f} {100*a[1]/tot:6.1f}%  {a[2]} {a[3]}")
    ours = {k: a for k, a in agg.items() if "rpc::" in k}
    t2 = sum(a[1] for a in ours.values())
    print("\n# shares among this library's kernels only (the step = one of each)")
    for k, a in sorted(ours.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:92s} {a[0]:4d} {a[1]/a[0]/1e3:9.2f} {100*a[1]/t2:6.1f}%")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, unit = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"## {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
        for h, u, v in zip(hdr, unit, r):
            if KEEP.match(h):
                print(f"{h:90s} {v:>16s} {u}")
        print()


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
