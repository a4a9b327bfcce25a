"""SURVEY.md 8e on REAL devices: skipped unless the box shows at least two GPUs.  One process drives all of them through the
library's own sharding -- hkbcn_compress_batch (whole surfaces per GPU, longest first), hkbcn_import_batch (two pipeline
slots per GPU), hkbcn_compress_chain_multi (bands of block rows of one chain) -- and every result must be byte-identical to
the one-device result and to the reference.  No collective runs anywhere on this path."""
import ctypes as C

import numpy as np
import pytest

import hkbcn
import refshim
import synth

pytestmark = pytest.mark.gpu
BC1, BC4, BC5, BC7 = 54, 60, 62, 66


def n_devices():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs_two = pytest.mark.skipif(n_devices() < 2, reason="needs at least two GPUs")


@needs_two
def test_library_sees_every_device():
    hkbcn.init()
    assert hkbcn.lib().hkbcn_device_count() == n_devices()


@needs_two
def test_compress_batch_across_devices():
    hkbcn.init()
    nd = hkbcn.lib().hkbcn_device_count()
    jobs = []
    for i in range(4 * nd):
        w, h = 128 << (i % 2), 128
        fmt, src = ((BC1, synth.albedo_rgba8(w, h, 300 + i)), (BC5, synth.normals_rg8(w, h, 300 + i)), (BC7, synth.albedo_rgba8(w, h, 300 + i, "mask")),
                    (BC4, synth.roughness_r8(w, h, 300 + i)))[i % 4]
        jobs.append((fmt, src, w, h, 0))
    plan = hkbcn.plan_batch([(f, w_, h_) for f, _, w_, h_, _ in jobs], nd)
    assert set(plan) == set(range(nd)), f"plan leaves a GPU without work: {plan}"
    outs = hkbcn.compress_batch(jobs)
    for (fmt, src, w, h, _), o in zip(jobs, outs):
        assert np.array_equal(o, refshim.compress(fmt, src, w, h))


@needs_two
@pytest.mark.parametrize("fmt", [BC1, BC7])
def test_banded_chain_across_devices(fmt):
    hkbcn.init()
    nd = hkbcn.lib().hkbcn_device_count()
    w = h = 512
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    img = synth.albedo_rgba8(w, h, 0x410 + fmt, "mask").reshape(h, w, 4)
    chain = hkbcn.generate_mips(fmt, img, w, h, mips)
    one = hkbcn.compress_chain_multi(fmt, chain, w, h, mips, 1)
    many = hkbcn.compress_chain_multi(fmt, chain, w, h, mips, nd)
    assert np.array_equal(one, many)
    assert np.array_equal(one, refshim.compress_chain(fmt, chain, w, h, mips))
    # every device gets rows of level 0
    rows = [hkbcn.plan_bands(h, mips, nd, d) for d in range(nd)]
    assert all(r1[0] > r0[0] for r0, r1 in rows)


@needs_two
def test_import_batch_across_devices(tmp_path):
    hkbcn.init()
    L = hkbcn.lib()
    nd = L.hkbcn_device_count()
    n = 3 * nd
    w = h = 256
    mips = L.hkbcn_calc_num_mips(w, h)
    srcs = [np.ascontiguousarray(synth.albedo_rgba8(w, h, 0x500 + i, "mask" if i % 2 else "opaque")) for i in range(n)]
    dsts = [np.zeros(L.hkbcn_chain_bytes(BC7 if i % 2 else BC1, w, h, 1, mips), np.uint8) for i in range(n)]
    jobs = (hkbcn.ImportJob * n)()
    for i in range(n):
        jobs[i].fmt, jobs[i].width, jobs[i].height, jobs[i].generate_mips = (BC7 if i % 2 else BC1), w, h, 1
        jobs[i].src_level0 = srcs[i].ctypes.data_as(C.c_void_p)
        jobs[i].out_path = None
        jobs[i].dst = dsts[i].ctypes.data_as(C.c_void_p)
    where = (C.c_int * n)()
    hkbcn.check(L.hkbcn_plan_import_batch(jobs, n, nd, where))
    assert set(where) == set(range(nd))
    hkbcn.check(L.hkbcn_import_batch(jobs, n))
    busy = (C.c_double * nd)()
    assert L.hkbcn_batch_stats(busy, nd) == nd and all(b > 0 for b in busy)
    for i in range(n):
        assert jobs[i].status == 0
        assert np.array_equal(dsts[i], hkbcn.import_chain(BC7 if i % 2 else BC1, srcs[i], w, h, mips))
