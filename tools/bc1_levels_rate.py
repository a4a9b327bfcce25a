#!/usr/bin/env python3
"""Encode rate of the BC1 quality levels (row a15) on one GPU next to the reference's rgbcx at the same level on one host thread.
usage: python tools/bc1_levels_rate.py [size]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hkbcn  # noqa: E402
import refshim  # noqa: E402
import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
hkbcn.init(1)
t = synth.albedo_rgba8(256, 256, 0xB2000001, "mask").reshape(256, 256, 4)
img = np.ascontiguousarray(np.tile(t, (n // 256, n // 256, 1)))
sample = np.ascontiguousarray(t[:64, :64]).reshape(16, 4, 16, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64)
for level, opts in ((5, 0), (5, 3), (6, 0), (8, 0), (10, 0), (12, 3), (13, 0), (16, 0), (18, 0), (18, 3)):
    flags = ((level + 1) << 4) | (hkbcn.FLAG_BC1_3COLOR if opts & 1 else 0) | (hkbcn.FLAG_BC1_TRANSPARENT_BLACK if opts & 2 else 0)
    hkbcn.compress(54, img[:256, :256], 256, 256, flags)
    t0 = time.perf_counter()
    hkbcn.compress(54, img, n, n, flags)
    gpu = time.perf_counter() - t0
    t0 = time.perf_counter()
    refshim.encode_bc1_level_blocks(sample, sample.shape[0], level, bool(opts & 1), bool(opts & 2))
    cpu = time.perf_counter() - t0
    print(f"level {level:2d} options {opts}: GPU {n * n / gpu / 1e6:8.1f} Mpix/s (host call, {gpu * 1e3:7.1f} ms for {n}^2)   reference, 1 thread: {sample.shape[0] * 16 / cpu / 1e6:6.3f} Mpix/s")
