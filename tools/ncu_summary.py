#!/usr/bin/env python3
"""Summarise an .ncu-rep (details page) into 'section | metric | value unit' lines for profiles/.
usage: ncu_summary.py report.ncu-rep [launch-id] [--all]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
lid = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("-") else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "details", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
I = {n: h.index(n) for n in ("ID", "Kernel Name", "Section Name", "Metric Name", "Metric Unit", "Metric Value")}
name = None
for r in rows[1:]:
    if r[I["ID"]] != lid:
        continue
    if name is None:
        name = r[I["Kernel Name"]]
        print("# kernel:", name)
    if r[I["Metric Name"]]:
        print(f'{r[I["Section Name"]]} | {r[I["Metric Name"]]} | {r[I["Metric Unit"]]} | {r[I["Metric Value"]]}')
