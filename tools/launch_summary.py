#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel: total time, share, count, average.
usage: launch_summary.py launches.csv ["header comment"]"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
h = next(r for r in rows if "Kernel Name" in r)
ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows:
    if r is h or len(r) <= iv or r[ik] == "Kernel Name":
        continue
    try:
        v = float(r[iv].replace(",", ""))
    except ValueError:
        continue
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[iu].replace("second", "s").replace("n", "n"), 1e-3 if r[iu].startswith("n") else 1.0)
    tot[r[ik]] += v
    cnt[r[ik]] += 1
T = sum(tot.values())
if len(sys.argv) > 2:
    print("#", sys.argv[2])
print(f"# {sum(cnt.values())} launches, {T / 1e3:.1f} ms total")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{v:12.1f} us {100 * v / T:5.1f}%  n={cnt[k]:4d} avg={v / cnt[k]:10.1f} us  {k[:110]}")
