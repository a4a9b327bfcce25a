#!/usr/bin/env python3
"""BCn decode bandwidth sweep (BASELINE.json configs[4]): whole-surface decode per format and size through
hkbcn_decompress_dev (device pointers), CUDA-event timed on the launching stream, L2 flushed between iterations.
Algorithmic bytes per pixel are SURVEY.md 8d's (block bytes read + decoded texel bytes written, no credit for re-reads).

usage: python tools/decode_sweep.py [--sizes 1024,4096,8192] [--formats BC1,BC7] [--iters 10] [--real]
--real: BC7 / BC6H inputs are blocks produced by this library's encoders from synthetic images (mode mix of real content)
        instead of uniform random bits (all modes, maximally divergent)."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))

# name: (format id, block bytes, decoded texel bytes, flags)
FORMATS = {"BC1": (54, 8, 4, 0), "BC2": (56, 16, 4, 0), "BC3": (58, 16, 4, 0), "BC4": (60, 8, 1, 0), "BC5": (62, 16, 2, 0),
           "BC7": (66, 16, 4, 0), "BC6H_f16": (64, 16, 6, 0), "BC6H_S_f16": (65, 16, 6, 0), "BC6H_f32": (64, 16, 12, 1 << 8)}


def peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except (OSError, KeyError, ValueError):
        return 6650.0


OPAQUE = False


def real_blocks(torch, hkbcn, L, name, n, stream):
    """encode a synthetic n x n image once (tiled from a 512^2 source) so the decoder sees a realistic mode mix"""
    import synth
    t = 512
    if name == "BC7":
        a = (synth.albedo_rgba8(t, t, 0xB2000004) if OPAQUE else synth.albedo_rgba8(t, t, 0xB2000005, "mask")).reshape(t, t, 4)
        img = np.tile(a, (n // t, n // t, 1))
        fmt = 66
    else:
        # BC6H encode is slow: encode 512^2 once and tile the BLOCKS up to n x n
        img = synth.sky_rgba32f(t, t, 0xB2000006, 0)
        if name == "BC6H_S":
            img[..., :3] -= 1.0  # mixed signs
        d_src = torch.from_numpy(np.ascontiguousarray(img).view(np.uint8).ravel()).cuda()
        d_dst = torch.empty((t // 4) * (t // 4) * 16, dtype=torch.uint8, device="cuda")
        hkbcn.check(L.hkbcn_compress_dev(0, 65 if name == "BC6H_S" else 64, d_src.data_ptr(), d_dst.data_ptr(), t, t, 0, C.c_void_p(stream.cuda_stream)))
        torch.cuda.synchronize()
        return d_dst.view(t // 4, t // 4, 16).repeat(n // t, n // t, 1).contiguous().view(-1)
    d_src = torch.from_numpy(np.ascontiguousarray(img).view(np.uint8).ravel()).cuda()
    d_dst = torch.empty((n // 4) * (n // 4) * 16, dtype=torch.uint8, device="cuda")
    hkbcn.check(L.hkbcn_compress_dev(0, fmt, d_src.data_ptr(), d_dst.data_ptr(), n, n, 0, C.c_void_p(stream.cuda_stream)))
    torch.cuda.synchronize()
    return d_dst


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1024,2048,4096,8192")
    ap.add_argument("--formats", default=",".join(FORMATS))
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--real", action="store_true")
    ap.add_argument("--opaque", action="store_true", help="--real BC7 input: opaque albedo (modes 1/6) instead of the alpha-mask variant")
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    global OPAQUE
    OPAQUE = args.opaque
    import torch
    import hkbcn
    L = hkbcn.lib()
    hkbcn.init(1)
    stream = torch.cuda.current_stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    gen = torch.Generator(device="cuda").manual_seed(7)
    pk = peak()
    rows = []
    for name in args.formats.split(","):
        fmt, blk, tex, flags = FORMATS[name]
        for n in [int(s) for s in args.sizes.split(",")]:
            nblk = (n // 4) * (n // 4)
            if args.real and name in ("BC7", "BC6H_f16", "BC6H_S_f16", "BC6H_f32") and n >= 512:
                src = real_blocks(torch, hkbcn, L, "BC7" if name == "BC7" else "BC6H_S" if name == "BC6H_S_f16" else "BC6H", n, stream)
                kind = "encoded"
            else:
                src = torch.randint(0, 256, (nblk * blk,), dtype=torch.uint8, device="cuda", generator=gen)
                kind = "random"
            dst = torch.empty(n * n * tex, dtype=torch.uint8, device="cuda")
            times = []
            for i in range(args.iters + 3):
                flush.fill_(i & 255)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                hkbcn.check(L.hkbcn_decompress_dev(0, fmt, src.data_ptr(), dst.data_ptr(), n, n, n * tex, flags, C.c_void_p(stream.cuda_stream)))
                e1.record(stream)
                torch.cuda.synchronize()
                if i >= 3:
                    times.append(e0.elapsed_time(e1))
            ms = float(np.mean(times))
            gbs = (nblk * blk + n * n * tex) / (ms * 1e-3) / 1e9
            rows.append({"format": name, "size": n, "input": kind, "ms": ms, "ms_min": float(np.min(times)), "GB/s": gbs, "frac": gbs / pk,
                         "Gpix/s": n * n / (ms * 1e-3) / 1e9})
            print(f"{name:9s} {n:6d}^2 {kind:8s} {ms * 1e3:9.1f} us  {gbs:8.1f} GB/s  {100 * gbs / pk:5.1f}% of {pk:.0f}", flush=True)
            del src, dst
    if args.json:
        with open(args.json, "w") as f:
            json.dump({"peak_gbs": pk, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
