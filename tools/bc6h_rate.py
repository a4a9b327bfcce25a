#!/usr/bin/env python3
"""BC6H encode time per surface size (sky faces, device-resident, CUDA events, L2 flushed).  usage: [HKBCN_LIB=...] python tools/bc6h_rate.py [sizes]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import hkbcn  # noqa: E402
import synth  # noqa: E402

sizes = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "128,256,512,1024").split(",")]
L = hkbcn.lib()
hkbcn.init(1)
st = torch.cuda.current_stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for fmt, name in ((64, "unsigned"), (65, "signed")):
    for n in sizes:
        t = min(n, 256)
        img = np.tile(synth.sky_rgba32f(t, t, 0xB2000006, 0).reshape(t, t, 4), (n // t, n // t, 1))
        if fmt == 65:
            img = img.copy()
            img[..., :3] -= 1.0
        src = torch.from_numpy(np.ascontiguousarray(img).view(np.uint8).ravel()).cuda()
        dst = torch.empty((n // 4) ** 2 * 16, dtype=torch.uint8, device="cuda")
        ts = []
        for i in range(5):
            flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            hkbcn.check(L.hkbcn_compress_dev(0, fmt, src.data_ptr(), dst.data_ptr(), n, n, 0, C.c_void_p(st.cuda_stream)))
            e1.record(st)
            torch.cuda.synchronize()
            if i >= 1:
                ts.append(e0.elapsed_time(e1))
        ms = float(np.mean(ts))
        print(f"{os.path.basename(hkbcn.LIB_PATH):22s} {name:8s} {n:5d}^2 {ms:8.2f} ms {n * n / ms / 1e3:7.2f} Mpix/s  crc {int(dst.to(torch.int64).sum())}", flush=True)
