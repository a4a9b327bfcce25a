#!/usr/bin/env python3
"""A/B timing of the BC7 encoder builds (HKBCN_LIB=build/variants/...): a 4096^2 surface tiled from a 1024^2 synthetic
albedo, opaque and with the soft alpha mask, device-resident, CUDA events, L2 flushed between iterations.
usage: [HKBCN_LIB=...] python tools/bc7_ab.py [size] [iters]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
import torch  # noqa: E402
import hkbcn  # noqa: E402
import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
L = hkbcn.lib()
hkbcn.init(1)
st = torch.cuda.current_stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name, kind in (("opaque", "opaque"), ("mask", "mask")):
    t = synth.albedo_rgba8(1024, 1024, 0xB2000004 if kind == "opaque" else 0xB2000005, kind).reshape(1024, 1024, 4)
    img = np.ascontiguousarray(np.tile(t, (n // 1024, n // 1024, 1)))
    src = torch.from_numpy(img.ravel()).cuda()
    dst = torch.empty((n // 4) ** 2 * 16, dtype=torch.uint8, device="cuda")
    ts = []
    for i in range(iters + 2):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        hkbcn.check(L.hkbcn_compress_dev(0, 66, src.data_ptr(), dst.data_ptr(), n, n, 0, C.c_void_p(st.cuda_stream)))
        e1.record(st)
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    ms = float(np.mean(ts))
    print(f"{os.path.basename(hkbcn.LIB_PATH):28s} {name:7s} {n}^2  {ms:8.3f} ms  {n * n / ms / 1e3:8.1f} Mpix/s  (min {min(ts):.3f})", flush=True)
