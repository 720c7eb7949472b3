#!/usr/bin/env python
"""Per-source-line stall samples of an ncu report (needs -lineinfo and --import-source on):
    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]
Rows where the SASS address is '-' are the per-line totals of the cuda,sass correlated view."""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur, hdr, lines = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[2] == "-":
        lines.append((cur, r))
i_s = hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[i_s] or 0) for _, r in lines)
print(f"# {rep}: {tot} samples; top {top} source lines (share, samples, dominant stalls, file:line, source)")
for f, r in sorted(lines, key=lambda fr: -int(fr[1][i_s] or 0))[:top]:
    n = int(r[i_s] or 0)
    st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall), reverse=True)[:3]
    sts = " ".join(f"{nm}:{v}" for v, nm in st if v)
    print(f"{100*n/max(tot,1):5.1f}% {n:6d}  {sts:45s} {f}:{r[0]:>4s}  {r[1].strip()[:95]}")
