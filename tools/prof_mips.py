#!/usr/bin/env python3
"""ncu driver for the device mip cascade: one hkbcn_generate_mips of a 4096^2 RGBA8 surface (Mitchell / wrap, alpha-weighted).
usage: prof_mips.py [size]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
import hkbcn  # noqa: E402
import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
hkbcn.init()
mips = hkbcn.lib().hkbcn_calc_num_mips(n, n)
chain = hkbcn.generate_mips(66, synth.albedo_rgba8(n, n, 0xB2000004), n, n, mips)
print("ok", chain.size, "launches", hkbcn.lib().hkbcn_launch_count())
