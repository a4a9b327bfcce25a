#!/usr/bin/env python3
"""Encode throughput per format (device-resident input, CUDA events on the launching stream, L2 flushed between iterations)
through hkbcn_compress_dev.  usage: python tools/encode_sweep.py [--formats BC1,BC7,...] [--size 2048] [--iters 5]"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))

FORMATS = {"BC1": 54, "BC2": 56, "BC3": 58, "BC4": 60, "BC5": 62, "BC6H": 64, "BC6H_S": 65, "BC7": 66, "BC7_alpha": 66}


def source(name, n):
    import synth
    if name in ("BC1",):
        return synth.albedo_rgba8(n, n, 0xB2000001)
    if name in ("BC2", "BC3", "BC7_alpha"):
        return synth.albedo_rgba8(n, n, 0xB2000005, "mask")
    if name == "BC7":
        return synth.albedo_rgba8(n, n, 0xB2000004)
    if name == "BC4":
        return synth.roughness_r8(n, n, 0xB2000002)
    if name == "BC5":
        return synth.normals_rg8(n, n, 0xB2000003)
    img = synth.sky_rgba32f(n, n, 0xB2000006, 0)
    if name == "BC6H_S":
        img[..., :3] -= 1.0
    return img


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--formats", default="BC1,BC3,BC4,BC5,BC7,BC7_alpha,BC6H")
    ap.add_argument("--size", type=int, default=2048)
    ap.add_argument("--hdr-size", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    import torch
    import hkbcn
    L = hkbcn.lib()
    hkbcn.init(1)
    stream = torch.cuda.current_stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for name in args.formats.split(","):
        fmt = FORMATS[name]
        n = args.hdr_size if name.startswith("BC6H") else args.size
        img = np.ascontiguousarray(source(name, n))
        d_src = torch.from_numpy(img.view(np.uint8).ravel()).cuda()
        d_dst = torch.empty((n // 4) * (n // 4) * hkbcn.block_bytes(fmt), dtype=torch.uint8, device="cuda")
        times = []
        for i in range(args.iters + 2):
            flush.fill_(i & 255)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            hkbcn.check(L.hkbcn_compress_dev(0, fmt, d_src.data_ptr(), d_dst.data_ptr(), n, n, 0, C.c_void_p(stream.cuda_stream)))
            e1.record(stream)
            torch.cuda.synchronize()
            if i >= 2:
                times.append(e0.elapsed_time(e1))
        ms = float(np.mean(times))
        rows.append({"format": name, "size": n, "ms": ms, "Mpix/s": n * n / ms / 1e3})
        print(f"{name:10s} {n:5d}^2  {ms:10.3f} ms  {n * n / ms / 1e3:12.1f} Mpix/s", flush=True)
    if args.json:
        with open(args.json, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
