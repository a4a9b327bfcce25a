#!/usr/bin/env python3
"""Randomised parity soak against the reference (oracle/_ref): random formats, sizes (ragged ones included), content kinds
and source address offsets through hkbcn_compress / hkbcn_compress_dev / hkbcn_decompress_dev, every result compared byte
for byte.  TEST INFRASTRUCTURE (uses oracle/).  usage: python tools/parity_soak.py [seconds] [seed]"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import hkbcn  # noqa: E402
import refshim  # noqa: E402
import synth  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
L = hkbcn.lib()
hkbcn.init(1)
FMT = {"BC1": 54, "BC2": 56, "BC3": 58, "BC4": 60, "BC5": 62, "BC6U": 64, "BC6S": 65, "BC7": 66}
BPP = {"BC4": 1, "BC5": 2, "BC6U": 16, "BC6S": 16}


def content(name, w, h):
    kind = rng.integers(0, 5)
    seed = int(rng.integers(1, 1 << 30))
    if name in ("BC6U", "BC6S"):
        img = synth.sky_rgba32f(w, h, seed, 0).reshape(h, w, 4)
        if kind == 1:
            img = (rng.random((h, w, 4)) * 10.0 ** rng.uniform(-3, 4.5)).astype(np.float32)
        if name == "BC6S":
            img[::2, ::3, :3] *= -1
        return np.ascontiguousarray(img)
    bpp = BPP.get(name, 4)
    if kind == 0:
        img = rng.integers(0, 256, (h, w, bpp), dtype=np.uint8)                                  # noise
    elif kind == 1:
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([(xx * (3 + c) + yy * (7 - c) + seed) & 255 for c in range(bpp)], -1).astype(np.uint8)  # ramps
    elif kind == 2:
        img = np.repeat(np.repeat(rng.integers(0, 256, ((h + 7) // 8, (w + 7) // 8, bpp), dtype=np.uint8), 8, 0), 8, 1)[:h, :w]  # flats
    elif kind == 3 and bpp == 4:
        img = synth.albedo_rgba8(w, h, seed, "mask").reshape(h, w, 4)
    else:
        base = rng.integers(0, 256, (1, 1, bpp))
        img = np.clip(base + rng.integers(-6, 7, (h, w, bpp)), 0, 255).astype(np.uint8)          # low contrast
    return np.ascontiguousarray(img)


t_end = time.time() + budget
n_cases = bad = 0
px_total = 0
while time.time() < t_end:
    name = list(FMT)[rng.integers(0, len(FMT))]
    fmt = FMT[name]
    big = name not in ("BC6U", "BC6S")
    w = int(rng.integers(1, 40 if not big else 160)) * (4 if rng.random() < 0.7 else 1) + (0 if rng.random() < 0.7 else int(rng.integers(0, 4)))
    h = int(rng.integers(1, 40 if not big else 160)) * 4 + (0 if rng.random() < 0.8 else int(rng.integers(0, 4)))
    w, h = max(w, 4), max(h, 4)
    img = content(name, w, h)
    want = refshim.compress(fmt, img, w, h, lib=refshim.ref0() if not big else None)
    got = hkbcn.compress(fmt, img, w, h)
    ok = np.array_equal(got, want)
    # device-pointer form with the source at an odd address (byte offsets for 8-bit texels, multiples of 4 for floats)
    off = int(rng.integers(0, 16)) * (4 if not big else 1)
    raw = np.ascontiguousarray(img).view(np.uint8).ravel()
    dbuf = torch.zeros(raw.size + 64, dtype=torch.uint8, device="cuda")
    dbuf[off: off + raw.size] = torch.from_numpy(raw.copy()).cuda()
    dout = torch.zeros(max(want.size, 16), dtype=torch.uint8, device="cuda")
    rc = L.hkbcn_compress_dev(0, fmt, dbuf.data_ptr() + off, dout.data_ptr(), w, h, 0, C.c_void_p(0))
    torch.cuda.synchronize()
    ok = ok and rc == 0 and np.array_equal(dout.cpu().numpy()[: want.size], want)
    # and the decoder on what was just encoded
    if want.size:
        tex = {54: 4, 56: 4, 58: 4, 60: 1, 62: 2, 64: 6, 65: 6, 66: 4}[fmt]
        wd, hd = (w // 4) * 4, (h // 4) * 4
        back = hkbcn.decompress(fmt, want, wd, hd)
        ok = ok and np.array_equal(np.asarray(back).ravel(), np.asarray(refshim.decompress(fmt, want, wd, hd)).ravel())
        del tex
    n_cases += 1
    px_total += w * h
    if not ok:
        bad += 1
        print(f"MISMATCH {name} {w}x{h} offset {off}", flush=True)
print(f"parity soak: {n_cases} cases, {px_total / 1e6:.2f} Mpix, {bad} mismatches")
sys.exit(1 if bad else 0)
