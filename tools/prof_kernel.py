#!/usr/bin/env python3
"""Tiny driver for ncu captures: one format, one surface size, a few encodes (or decodes) through the C ABI.
usage: prof_kernel.py <fmt-id> <size> [reps] [decode|decode32|mask]   (HKBCN_PROF_FLAGS=<int>: flags of the encode, e.g. a quality level)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
import hkbcn  # noqa: E402
import synth  # noqa: E402

fmt, n = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
decode = len(sys.argv) > 4 and sys.argv[4] in ("decode", "decode32")
f32 = len(sys.argv) > 4 and sys.argv[4] == "decode32"
hkbcn.init()
bpp = hkbcn.lib().hkbcn_src_texel_bytes(fmt, 0)
if decode:
    if fmt == 66:
        blocks = hkbcn.compress(66, synth.albedo_rgba8(n, n, 4), n, n)
    elif fmt in (64, 65):
        # BC6H encode is slow: a 256^2 sky encoded once, its blocks tiled up to n x n
        t = 256
        b = np.asarray(hkbcn.compress(fmt, synth.sky_rgba32f(t, t, 6), t, t)).reshape(t // 4, t // 4, 16)
        blocks = np.ascontiguousarray(np.tile(b, (n // t, n // t, 1))).ravel()
    else:
        blocks = synth.random_blocks((n // 4) ** 2, 1, 5)[: (n // 4) ** 2 * hkbcn.block_bytes(fmt)]
    for _ in range(reps):
        out = hkbcn.decompress(fmt, blocks, n, n, flags=(1 << 8) if f32 else 0) if f32 else hkbcn.decompress(fmt, blocks, n, n)
else:
    if bpp == 4:
        img = synth.albedo_rgba8(n, n, 0xB2000004, "mask" if len(sys.argv) > 4 and sys.argv[4] == "mask" else "opaque")
    elif bpp == 1:
        img = synth.roughness_r8(n, n, 2)
    elif bpp == 2:
        img = synth.normals_rg8(n, n, 3)
    else:
        img = synth.sky_rgba32f(n, n, 6)
    for _ in range(reps):
        out = hkbcn.compress(fmt, img, n, n, int(os.environ.get("HKBCN_PROF_FLAGS", "0")))
print("ok", out[:16].tolist(), "launches", hkbcn.lib().hkbcn_launch_count())
