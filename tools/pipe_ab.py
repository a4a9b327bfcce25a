#!/usr/bin/env python3
"""The host-pointer piece pipeline (compress_host_pipelined) on the BC7 4096^2 chain: pinned and pageable buffers,
hkbcn_compress_chain (the chain crosses PCIe) and hkbcn_import_chain (level 0 only, mips on the device), for different band
counts (HKBCN_PIPE_BANDS), one or two encode streams (HKBCN_PIPE_ONE_STREAM) and the blocking form (HKBCN_PIPE_PINNED=0 /
HKBCN_NO_PIPELINE=1); all read per call.  usage: python tools/pipe_ab.py"""
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
import synth  # noqa: E402

n = 4096
L = hkbcn.lib()
hkbcn.init(1)
mips = L.hkbcn_calc_num_mips(n, n)
t = synth.albedo_rgba8(1024, 1024, 0xB2000004).reshape(1024, 1024, 4)
img = np.ascontiguousarray(np.tile(t, (4, 4, 1)))
chain = hkbcn.generate_mips(66, img, n, n, mips)
nbytes = L.hkbcn_chain_bytes(66, n, n, 1, mips)
px = nbytes
pin_src = torch.from_numpy(chain.copy()).pin_memory()
pin_dst = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
pg_dst = np.zeros(nbytes, np.uint8)
want = None


def run(label, fn, out, env):
    global want
    for k in ("HKBCN_PIPE_BANDS", "HKBCN_PIPE_ONE_STREAM", "HKBCN_PIPE_PINNED", "HKBCN_NO_PIPELINE", "HKBCN_PIPE_COARSE"):
        os.environ.pop(k, None)
    os.environ.update(env)
    ts = []
    for i in range(10):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    got = out() if callable(out) else out
    if want is None:
        want = got.copy()
    same = bool(np.array_equal(got, want))
    ms = float(np.mean(ts[2:])) * 1e3
    print(f"{label:58s} {ms:7.2f} ms {px / ms / 1e3:7.1f} Mpix/s  {'same bytes' if same else 'DIFFERENT BYTES'}", flush=True)


def pinned_chain():
    hkbcn.check(L.hkbcn_compress_chain(66, C.c_void_p(pin_src.data_ptr()), C.c_void_p(pin_dst.data_ptr()), n, n, mips, 0))


def pageable_chain():
    hkbcn.check(L.hkbcn_compress_chain(66, chain.ctypes.data_as(C.c_void_p), pg_dst.ctypes.data_as(C.c_void_p), n, n, mips, 0))


def pinned_import():
    hkbcn.check(L.hkbcn_import_chain(66, C.c_void_p(pin_src.data_ptr()), C.c_void_p(pin_dst.data_ptr()), n, n, mips, None, 0))


pd = lambda: pin_dst.numpy()
run("compress_chain pinned, blocking", pinned_chain, pd, {"HKBCN_PIPE_PINNED": "0"})
run("compress_chain pinned, pieces, one stream", pinned_chain, pd, {"HKBCN_PIPE_ONE_STREAM": "1"})
run("compress_chain pinned, pieces, two streams", pinned_chain, pd, {})
for b in ("2", "3", "4", "7"):
    run(f"compress_chain pinned, two streams, {b} bands", pinned_chain, pd, {"HKBCN_PIPE_BANDS": b})
run("compress_chain pageable, one stream", pageable_chain, pg_dst, {"HKBCN_PIPE_ONE_STREAM": "1"})
run("compress_chain pageable, two streams, coarse pieces", pageable_chain, pg_dst, {"HKBCN_PIPE_COARSE": "1"})
run("compress_chain pageable, two streams", pageable_chain, pg_dst, {})
run("compress_chain pinned, two streams, coarse pieces", pinned_chain, pd, {"HKBCN_PIPE_COARSE": "1"})
run("import_chain pinned, blocking", pinned_import, pd, {"HKBCN_NO_PIPELINE": "1"})
run("import_chain pinned, pieces, one stream", pinned_import, pd, {"HKBCN_PIPE_ONE_STREAM": "1"})
run("import_chain pinned, pieces, two streams, coarse", pinned_import, pd, {"HKBCN_PIPE_COARSE": "1"})
run("import_chain pinned, pieces, two streams", pinned_import, pd, {})
