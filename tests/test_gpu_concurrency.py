"""Concurrent use of the C ABI (include/hkbcn.h: "Work on one stream runs in order; different streams may run concurrently"; the
host-pointer calls are thread-safe).  The persistent encoders and decoders keep per-STREAM state on the device -- the tile
counters of every persistent kernel, the BC7 encoder's work list -- and the host-pointer calls share per-device staging
behind a mutex: launches queued on several streams at once, with growing and shrinking surfaces, and host threads calling in
parallel must all produce the one-at-a-time result."""
import ctypes as C
import threading

import numpy as np
import pytest

import hkbcn
import refshim
import synth

pytestmark = pytest.mark.gpu
BC1, BC4, BC6U, BC7 = 54, 60, 64, 66


def test_many_streams_in_flight():
    import torch
    L = hkbcn.lib()
    hkbcn.init()
    streams = [torch.cuda.Stream() for _ in range(4)]
    sizes = [(256, 256), (64, 128), (512, 256), (128, 64), (256, 512), (64, 64)]
    work = []
    for k, (w, h) in enumerate(sizes * 2):
        kind = "mask" if k % 2 else "opaque"
        img = synth.albedo_rgba8(w, h, 0x900 + k, kind)
        src = torch.from_numpy(np.ascontiguousarray(img).ravel().copy()).cuda()
        dst = torch.zeros((w // 4) * (h // 4) * 16, dtype=torch.uint8, device="cuda")
        back = torch.zeros(w * h * 4, dtype=torch.uint8, device="cuda")
        work.append((img, w, h, src, dst, back))
    torch.cuda.synchronize()
    # queue everything without waiting: BC7 encode, then the decode of its own output, on four streams round robin, twice
    for rep in range(2):
        for k, (img, w, h, src, dst, back) in enumerate(work):
            st = streams[(k + rep) % 4]
            cs = C.c_void_p(st.cuda_stream)
            hkbcn.check(L.hkbcn_compress_dev(0, BC7, src.data_ptr(), dst.data_ptr(), w, h, 0, cs))
            hkbcn.check(L.hkbcn_decompress_dev(0, BC7, dst.data_ptr(), back.data_ptr(), w, h, w * 4, 0, cs))
    torch.cuda.synchronize()
    for img, w, h, src, dst, back in work:
        want = refshim.compress(BC7, img, w, h)
        assert np.array_equal(dst.cpu().numpy(), want)
        assert np.array_equal(back.cpu().numpy(), refshim.decompress(BC7, want, w, h).ravel())


def test_host_threads_in_parallel():
    hkbcn.init()
    jobs = [(BC7, synth.albedo_rgba8(128, 128, 0xA00 + i, "mask"), 128, 128) for i in range(4)]
    jobs += [(BC1, synth.albedo_rgba8(256, 64, 0xA10 + i), 256, 64) for i in range(4)]
    jobs += [(BC4, synth.roughness_r8(64, 256, 0xA20 + i), 64, 256) for i in range(4)]
    jobs += [(BC6U, synth.sky_rgba32f(32, 32, 0xA30 + i, 0), 32, 32) for i in range(2)]
    out = [None] * len(jobs)
    err = []

    def run(i):
        try:
            fmt, src, w, h = jobs[i]
            for _ in range(3):
                out[i] = hkbcn.compress(fmt, src, w, h)
        except Exception as e:  # noqa: BLE001
            err.append(e)
    th = [threading.Thread(target=run, args=(i,)) for i in range(len(jobs))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not err, err
    for (fmt, src, w, h), o in zip(jobs, out):
        want = refshim.compress(fmt, src, w, h, lib=refshim.ref0() if fmt == BC6U else None)
        assert np.array_equal(o, want)
