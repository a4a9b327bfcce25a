#!/usr/bin/env python3
"""Device-resident timing of the mip cascade (hkbcn_generate_mips_dev) per format: whole chain and the first level alone.
usage: python tools/mips_sweep.py [size] [noise|smooth]   (HKBCN_MIP_GATHER=1: the one-thread-per-texel kernels, for comparison)
noise (default): independent random bytes -- the worst case for the kernels' shared-memory decode tables (32 lanes, 32 random
entries: bank conflicts); smooth: a 64 x 64 random image enlarged bilinearly plus +-2 LSB of noise, what a photograph looks like
to those tables (neighbouring texels read neighbouring entries)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
import torch  # noqa: E402
import hkbcn  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
content = sys.argv[2] if len(sys.argv) > 2 else "noise"
L = hkbcn.lib()
hkbcn.init(1)
st = torch.cuda.current_stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
gen = torch.Generator(device="cuda").manual_seed(3)
for name, fmt, bpp, cfg in (("RGBA8 (BC7, alpha-weighted, Mitchell/wrap)", 66, 4, None), ("RGBA8 sRGB (BC1_SRGB)", 55, 4, None),
                            ("R8 (BC4, triangle)", 60, 1, hkbcn.MipConfig(3, 2, 0, 0)), ("RG8 (BC5, triangle)", 62, 2, hkbcn.MipConfig(3, 2, 0, 0)),
                            ("RGBA32F (BC6H)", 64, 16, None)):
    for mips in (2, L.hkbcn_calc_num_mips(n, n)):
        nbytes = L.hkbcn_chain_src_bytes(fmt, n, n, mips, 0)
        if bpp == 16:
            chain = torch.rand(nbytes // 4, dtype=torch.float32, device="cuda", generator=gen).view(torch.uint8)
        else:
            chain = torch.randint(0, 256, (nbytes,), dtype=torch.uint8, device="cuda", generator=gen)
            if content == "smooth":
                low = torch.rand(1, bpp, 64, 64, device="cuda", generator=gen)
                img = torch.nn.functional.interpolate(low, size=(n, n), mode="bilinear", align_corners=False)[0].permute(1, 2, 0) * 250.0
                img = img + torch.randint(0, 5, (n, n, bpp), device="cuda", generator=gen)
                chain[: n * n * bpp] = img.clamp(0, 255).to(torch.uint8).reshape(-1)
        ts = []
        for i in range(6):
            flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            hkbcn.check(L.hkbcn_generate_mips_dev(0, fmt, chain.data_ptr(), n, n, mips, C.byref(cfg) if cfg else None, 0, C.c_void_p(st.cuda_stream)))
            e1.record(st)
            torch.cuda.synchronize()
            if i >= 2:
                ts.append(e0.elapsed_time(e1))
        ms = float(np.mean(ts))
        alg = n * n * bpp * (1.25 if mips == 2 else 1.6667)  # level i read + level i+1 written, summed over the levels built
        print(f"{content:6s} {name:45s} {n}^2 {'first level' if mips == 2 else 'whole chain'}: {ms * 1e3:8.1f} us  {alg / ms / 1e6:7.1f} GB/s algorithmic")
