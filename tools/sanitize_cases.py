#!/usr/bin/env python3
"""One small invocation of every kernel of the library through the C ABI -- the program tools/sanitize.sh runs under
compute-sanitizer (memcheck / racecheck / synccheck, one tool per run).  Sizes are tiny: the tools slow kernels 10-100x."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hkbcn  # noqa: E402
import synth  # noqa: E402

hkbcn.init()
L = hkbcn.lib()
rgba = synth.albedo_rgba8(64, 48, 5, "mask")
r8 = synth.roughness_r8(64, 48, 2)
rg8 = synth.normals_rg8(64, 48, 3)
hdr = synth.sky_rgba32f(32, 16, 6)
done = []
for fmt, src in ((54, rgba), (56, rgba), (58, rgba), (60, r8), (62, rg8), (66, rgba)):
    b = hkbcn.compress(fmt, src, 64, 48)
    hkbcn.decompress(fmt, b, 64, 48)
    done.append(f"fmt{fmt}")
b = hkbcn.compress(54, rgba[:, :22].copy(), 22, 48)  # ragged width: plain-load staging path
for fmt in (64, 65):
    b = hkbcn.compress(fmt, hdr, 32, 16)
    hkbcn.decompress(fmt, b, 32, 16)
    hkbcn.decompress(fmt, b, 32, 16, flags=hkbcn.FLAG_DST_F32)
    done.append(f"fmt{fmt}")
hkbcn.compress(64, hdr.astype(np.float16), 32, 16, hkbcn.FLAG_SRC_F16)
# chains, slices, the mip cascade (tile kernels and, non-square, the general ones), import
sq = synth.albedo_rgba8(64, 64, 7, "mask")
mips = L.hkbcn_calc_num_mips(64, 64)
hkbcn.import_chain(66, sq, 64, 64, mips)
hkbcn.import_chain(55, sq, 64, 64, mips)
hkbcn.import_chain(60, synth.roughness_r8(64, 16, 2), 64, 16, L.hkbcn_calc_num_mips(64, 16), hkbcn.MipConfig(1, 2, 0, 0))
hkbcn.generate_mips(64, synth.sky_rgba32f(32, 32, 6), 32, 32, 4, hkbcn.MipConfig(4, 3, 0, 0))
faces = np.concatenate([synth.sky_rgba32f(16, 16, 6, f).ravel() for f in range(6)])
hkbcn.compress_slices(64, faces, 16, 16, 6, 1)
blk = synth.random_blocks(6 * 16 * 3, 1, 3)
hkbcn.decompress_slices(66, blk[: L.hkbcn_chain_bytes(66, 16, 16, 6, 3)], 16, 16, 6, 3)
# packers, bake, environment probe
n = np.random.default_rng(1).normal(size=(1000, 3)).astype(np.float32)
for v in range(6):
    p = hkbcn.pack_normals(v, n)
    hkbcn.unpack_normals(v, p, 1000)
hkbcn.normalize_vectors(np.abs(n))
nm = hkbcn.normalize_vectors(np.random.default_rng(2).random((32 * 32, 3), dtype=np.float32))
chain = hkbcn.generate_mips(60, synth.roughness_r8(32, 32, 2), 32, 32, 4, hkbcn.MipConfig(3, 2, 0, 0))
hkbcn.bake_roughness(nm, chain, 32, 32, 4, 1)
hkbcn.bake_roughness(nm, chain, 32, 32, 4, 2)
cube = np.stack([synth.sky_rgba32f(16, 16, 6, f) for f in range(6)])
hkbcn.envprobe_prefilter(cube, 16, 2)
hkbcn.irradiance(cube, 16, 4)
print("sanitize_cases: ok,", L.hkbcn_launch_count(), "launches")
