#!/usr/bin/env python3
"""Per-source-line instruction, active-lane and stall-sample totals of an .ncu-rep captured with --import-source on
(-lineinfo build).  usage: ncu_lines.py report.ncu-rep [top-n] [file-filter]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
flt = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows, fname, hdr = [], "", None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
        ci, cs, ct = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
    elif hdr and len(r) > ci and r[2] == "-" and r[0].isdigit():
        rows.append((fname, int(r[0]), r[1].strip(), int(r[ci] or 0), int(r[cs] or 0), int(r[ct] or 0)))
ti, ts, tt = sum(x[3] for x in rows), sum(x[4] for x in rows), sum(x[5] for x in rows)
print(f"total warp instructions {ti}, thread instructions {tt} ({tt / max(ti, 1):.1f} lanes), samples {ts}")
byfile = {}
for f, ln, src, n, s, t in rows:
    a = byfile.setdefault(f, [0, 0, 0])
    a[0] += n; a[1] += s; a[2] += t
for f, (n, s, t) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f}: {100 * n / ti:5.1f}% inst, {t / max(n, 1):5.1f} lanes, {100 * s / max(ts, 1):5.1f}% samples")
for f, ln, src, n, s, t in sorted([r for r in rows if flt in r[0]], key=lambda x: -x[3])[:top]:
    print(f"{100 * n / ti:5.1f}% inst {t / max(n, 1):5.1f} lanes {100 * s / max(ts, 1):5.1f}% smp  {f}:{ln}  {src[:100]}")
