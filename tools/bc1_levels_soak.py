#!/usr/bin/env python3
"""Parity soak of the BC1 quality levels (row a15): many seeded blocks of several kinds, every level 5..18 x the four
combinations of Encode_BC1's 3-colour options, GPU (hkbcn_encode_blocks) against the reference (rgbcx::encode_bc1).
usage: python tools/bc1_levels_soak.py [blocks per kind] [seed]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hkbcn  # noqa: E402
import refshim  # noqa: E402

per = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rng = np.random.default_rng(seed)


def kinds(n):
    out = []
    px = rng.integers(0, 256, (n, 16, 4), dtype=np.uint8)                       # noise
    out.append(px)
    a, b = rng.integers(0, 256, (n, 1, 3)), rng.integers(0, 256, (n, 1, 3))     # two colours + ramp between them
    t = rng.integers(0, 5, (n, 16, 1)) / 4.0
    g = np.zeros((n, 16, 4), np.uint8); g[..., :3] = np.clip(a + (b - a) * t + rng.integers(-2, 3, (n, 16, 3)), 0, 255); out.append(g)
    base = rng.integers(0, 256, (n, 1, 3))                                      # near-solid
    s = np.zeros((n, 16, 4), np.uint8); s[..., :3] = np.clip(base + rng.integers(-1, 2, (n, 16, 3)) * (rng.random((n, 16, 1)) < 0.2), 0, 255); out.append(s)
    d = np.zeros((n, 16, 4), np.uint8); d[..., :3] = rng.integers(0, 12, (n, 16, 3)); out.append(d)   # dark
    m = rng.integers(0, 256, (n, 16, 4), dtype=np.uint8)                        # coloured with black holes
    m[rng.random((n, 16)) < 0.3, :3] = 0; out.append(m)
    gr = np.zeros((n, 16, 4), np.uint8); gr[..., :3] = rng.integers(0, 256, (n, 16, 1)); out.append(gr)  # greyscale
    q = np.zeros((n, 16, 4), np.uint8); q[..., :3] = (rng.integers(0, 32, (n, 16, 3)) * 8) | 4; out.append(q)  # on the 565 lattice's midpoints
    return np.concatenate(out).reshape(-1, 64)


hkbcn.init(1)
blocks = kinds(per)
n = blocks.shape[0]
bad_total = 0
for level in range(5, 19):
    sub = blocks if level < 13 else blocks[:: 4 if level < 16 else 8]
    for allow3, black in ((False, False), (True, False), (False, True), (True, True)):
        if level == 5 and not (allow3 or black):
            continue
        flags = ((level + 1) << 4) | (hkbcn.FLAG_BC1_3COLOR if allow3 else 0) | (hkbcn.FLAG_BC1_TRANSPARENT_BLACK if black else 0)
        got = hkbcn.encode_blocks(54, sub, sub.shape[0], flags)
        want = refshim.encode_bc1_level_blocks(sub, sub.shape[0], level, allow3, black)
        bad = np.flatnonzero((got.reshape(-1, 8) != want.reshape(-1, 8)).any(axis=1))
        bad_total += bad.size
        print(f"level {level:2d} 3colour {int(allow3)} black {int(black)}: {sub.shape[0]:6d} blocks, {bad.size} differ{'' if not bad.size else ' first ' + str(bad[:4])}", flush=True)
print("TOTAL MISMATCHES", bad_total)
sys.exit(1 if bad_total else 0)
