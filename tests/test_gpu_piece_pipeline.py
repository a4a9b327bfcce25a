"""The host-pointer piece pipeline (surfaces from 16 MB up are cut into bands of level 0 plus the mip levels,
uploaded under the kernels before them and encoded on two alternating streams; pageable buffers are staged through pinned
ones; hkbcn_import_chain builds the mips on the device after the last band).  Whatever the format, the shape, the kind of
host memory and the band count, the bytes out are those of the blocking form of the same call (HKBCN_NO_PIPELINE=1, read per
call), which the other tests compare with the reference."""
import ctypes as C
import os

import numpy as np
import pytest

import hkbcn
import synth

pytestmark = pytest.mark.gpu
BC1, BC3, BC4, BC5, BC7 = 54, 58, 60, 62, 66
ENV = ("HKBCN_NO_PIPELINE", "HKBCN_PIPE_BANDS", "HKBCN_PIPE_ONE_STREAM", "HKBCN_PIPE_COARSE", "HKBCN_PIPE_PINNED")


@pytest.fixture(autouse=True)
def clean_env():
    saved = {k: os.environ.pop(k, None) for k in ENV}
    yield
    for k, v in saved.items():
        os.environ.pop(k, None)
        if v is not None:
            os.environ[k] = v


class Pinned:
    def __init__(self, nbytes):
        self.L = hkbcn.lib()
        self.p = self.L.hkbcn_alloc_pinned(nbytes)
        assert self.p
        self.a = np.ctypeslib.as_array(C.cast(self.p, C.POINTER(C.c_uint8)), shape=(nbytes,))

    def free(self):
        self.L.hkbcn_free_pinned(self.p)


def surface(fmt, w, h, seed):
    base = 256
    if fmt == BC4:
        t = synth.roughness_r8(base, base, seed).reshape(base, base, 1)
    elif fmt == BC5:
        t = synth.normals_rg8(base, base, seed).reshape(base, base, 2)
    else:
        t = synth.albedo_rgba8(base, base, seed, "mask").reshape(base, base, 4)
    img = np.tile(t, (h // base, w // base, 1))
    img[: h // 2] = img[: h // 2, ::-1]  # not periodic in the band direction
    return np.ascontiguousarray(img)


CASES = [  # fmt, width, height, mips?
    (BC7, 4096, 1024, True),    # non-square chain: 16 MB of level 0, 13 levels
    (BC1, 4096, 2048, False),   # tiles of 128 blocks, one level
    (BC3, 2048, 2048, True),
    (BC4, 8192, 4096, False),   # R8: 32 MB
    (BC5, 4096, 4096, True),    # RG8
]


@pytest.mark.parametrize("fmt,w,h,mipmapped", CASES)
def test_pieces_equal_the_blocking_call(fmt, w, h, mipmapped):
    hkbcn.init()
    L = hkbcn.lib()
    mips = L.hkbcn_calc_num_mips(w, h) if mipmapped else 1
    img = surface(fmt, w, h, 0x71 + fmt)
    cfg = hkbcn.MipConfig(3, 2, 0, 0) if fmt in (BC4, BC5) else None
    chain = hkbcn.generate_mips(fmt, img, w, h, mips, cfg) if mips > 1 else img.ravel()
    nbytes = L.hkbcn_chain_bytes(fmt, w, h, 1, mips)
    os.environ["HKBCN_NO_PIPELINE"] = "1"
    want = hkbcn.compress_chain(fmt, chain, w, h, mips)
    del os.environ["HKBCN_NO_PIPELINE"]
    psrc, pdst = Pinned(chain.nbytes), Pinned(nbytes)
    psrc.a[:] = chain.view(np.uint8).ravel()
    try:
        for env in ({}, {"HKBCN_PIPE_BANDS": "2"}, {"HKBCN_PIPE_BANDS": "7"}, {"HKBCN_PIPE_ONE_STREAM": "1"}, {"HKBCN_PIPE_COARSE": "1"}):
            os.environ.update(env)
            # pageable in, pageable out
            got = hkbcn.compress_chain(fmt, chain, w, h, mips)
            assert np.array_equal(got, want), f"pageable, {env}"
            # pinned in, pinned out: the copy engine works on the caller's buffers
            pdst.a[:] = 0
            hkbcn.check(L.hkbcn_compress_chain(fmt, C.c_void_p(psrc.p), C.c_void_p(pdst.p), w, h, mips, 0))
            assert np.array_equal(pdst.a, want), f"pinned, {env}"
            # pinned in, pageable out
            out = np.zeros(nbytes, np.uint8)
            hkbcn.check(L.hkbcn_compress_chain(fmt, C.c_void_p(psrc.p), out.ctypes.data_as(C.c_void_p), w, h, mips, 0))
            assert np.array_equal(out, want), f"pinned in / pageable out, {env}"
            for k in env:
                del os.environ[k]
        if mips > 1:
            # the import form: level 0 only goes up, the cascade runs after the last band
            pdst.a[:] = 0
            hkbcn.check(L.hkbcn_import_chain(fmt, C.c_void_p(psrc.p), C.c_void_p(pdst.p), w, h, mips, C.byref(cfg) if cfg else None, 0))
            assert np.array_equal(pdst.a, want), "import_chain, pinned"
            out = np.zeros(nbytes, np.uint8)
            hkbcn.check(L.hkbcn_import_chain(fmt, img.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), w, h, mips, C.byref(cfg) if cfg else None, 0))
            assert np.array_equal(out, want), "import_chain, pageable"
    finally:
        psrc.free()
        pdst.free()


def test_import_chain_argument_errors_come_before_any_launch():
    hkbcn.init()
    L = hkbcn.lib()
    n = 2048
    img = surface(BC7, n, n, 5)
    mips = L.hkbcn_calc_num_mips(n, n)
    out = np.zeros(L.hkbcn_chain_bytes(BC7, n, n, 1, mips), np.uint8)
    bad = hkbcn.MipConfig(9, 5, 0, 0)
    assert L.hkbcn_import_chain(BC7, img.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), n, n, mips, C.byref(bad), 0) == -2
    assert b"edge mode" in L.hkbcn_last_error()
    assert (out == 0).all()
    # and the device is fine afterwards
    small = surface(BC7, 256, 256, 6)
    assert hkbcn.compress(BC7, small, 256, 256).any()
