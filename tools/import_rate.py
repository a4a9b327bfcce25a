#!/usr/bin/env python3
"""hkbcn_import_chain (level 0 in host memory, device cascade, encode, compressed chain back) per format at BASELINE.json's
config sizes, pinned buffers, wall clock per call.  usage: python tools/import_rate.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import hkbcn  # noqa: E402
import synth  # noqa: E402

L = hkbcn.lib()
hkbcn.init(1)
CASES = [("BC1 2048^2 (config 1)", 54, 2048, 4, None), ("BC1_SRGB 2048^2 (config 1)", 55, 2048, 4, None), ("BC4 4096^2 (config 2)", 60, 4096, 1, hkbcn.MipConfig(3, 2, 0, 0)),
         ("BC5 4096^2 (config 2)", 62, 4096, 2, hkbcn.MipConfig(3, 2, 0, 0)), ("BC3 4096^2", 58, 4096, 4, None), ("BC7 4096^2 (config 3)", 66, 4096, 4, None)]
for name, fmt, n, ch, cfg in CASES:
    t = 512
    tile = (synth.roughness_r8(t, t, 3).reshape(t, t, 1) if ch == 1 else synth.normals_rg8(t, t, 4).reshape(t, t, 2) if ch == 2
            else synth.albedo_rgba8(t, t, 0xB2000004, "mask" if fmt == 58 else "opaque").reshape(t, t, 4))
    img = np.ascontiguousarray(np.tile(tile, (n // t, n // t, 1)))
    mips = L.hkbcn_calc_num_mips(n, n)
    nbytes = L.hkbcn_chain_bytes(fmt, n, n, 1, mips)
    src = torch.from_numpy(img.ravel().copy()).pin_memory()
    dst = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
    ts = []
    for i in range(12):
        t0 = time.perf_counter()
        hkbcn.check(L.hkbcn_import_chain(fmt, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), n, n, mips, C.byref(cfg) if cfg else None, 0))
        ts.append(time.perf_counter() - t0)
    ms = float(np.mean(ts[2:])) * 1e3
    px = sum(max(n >> i, 4) ** 2 for i in range(mips))
    print(f"{name:28s} {ms:8.3f} ms  {px / ms / 1e3:9.1f} Mpix/s  ({img.nbytes / 1e6:.1f} MB up, {nbytes / 1e6:.1f} MB down)", flush=True)
