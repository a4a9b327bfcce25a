#!/usr/bin/env python3
"""hkbcn_compress_chain on pageable memory (BC7, 4096^2 chain) for different numbers of level-0 bands of the pinned piece
pipeline (HKBCN_PIPE_BANDS, read per call).  usage: python tools/pageable_ab.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
import hkbcn  # noqa: E402
import synth  # noqa: E402

n = 4096
L = hkbcn.lib()
hkbcn.init(1)
mips = L.hkbcn_calc_num_mips(n, n)
t = synth.albedo_rgba8(1024, 1024, 0xB2000004).reshape(1024, 1024, 4)
img = np.ascontiguousarray(np.tile(t, (4, 4, 1)))
chain = hkbcn.generate_mips(66, img, n, n, mips)
dst = np.zeros(L.hkbcn_chain_bytes(66, n, n, 1, mips), np.uint8)
px = dst.size
for bands in ("", "1", "2", "3", "4", "5", "6", "7"):
    if bands:
        os.environ["HKBCN_PIPE_BANDS"] = bands
    ts = []
    for i in range(12):
        t0 = time.perf_counter()
        hkbcn.check(L.hkbcn_compress_chain(66, chain.ctypes.data_as(C.c_void_p), dst.ctypes.data_as(C.c_void_p), n, n, mips, 0))
        ts.append(time.perf_counter() - t0)
    ms = float(np.mean(ts[2:])) * 1e3
    print(f"bands {bands or 'default'}: {ms:7.2f} ms  {px / ms / 1e3:7.1f} Mpix/s", flush=True)
