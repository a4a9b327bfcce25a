#!/usr/bin/env python3
"""Parity at sizes beyond the configs (TEST INFRASTRUCTURE, uses oracle/): BC7 and BC3 at 8192^2, BC1 at 16384^2 (2^24 blocks)
through hkbcn_compress / hkbcn_decompress against the reference on all host threads.  usage: python tools/big_surface_check.py"""
import sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import hkbcn, refshim, synth
t = synth.albedo_rgba8(1024, 1024, 0xB2000005, "mask").reshape(1024, 1024, 4)
for fmt, n in ((66, 8192), (54, 16384), (58, 8192)):
    img = np.ascontiguousarray(np.tile(t, (n // 1024, n // 1024, 1)))
    img[1024:2048] = img[1024:2048, ::-1]
    got = hkbcn.compress(fmt, img, n, n)
    want = refshim.compress(fmt, img, n, n)
    bb = 8 if fmt == 54 else 16
    d = (got.reshape(-1, bb) != want.reshape(-1, bb)).any(axis=1).sum()
    print(f"fmt {fmt} {n}^2: {got.size // bb} blocks, {d} differ", flush=True)
    back = hkbcn.decompress(fmt, got, n, n)
    print("  decode equal:", bool(np.array_equal(np.asarray(back).ravel(), np.asarray(refshim.decompress(fmt, want, n, n)).ravel())), flush=True)
