"""hkbcn_create_normal_map / hkbcn_create_normal_and_roughness: Hk::CreateNormalMap (Image.cpp:2262-2352) and
Hk::CreateNormalAndRoughness (Image.cpp:2459-2638) as single calls.  Every stage has its own oracle test (packers, the
normalisation and the bake against the reference's object code in test_gpu_normals.py; the cascade and the resampler against
the reference's stbir in test_gpu_mips.py / test_gpu_resample.py; BC1/3/4/5 against rgbcx in test_gpu_parity.py): here the
composition must equal the same stages called one by one, and the sizes must follow the reference's rules."""
import ctypes as C

import numpy as np
import pytest

import hkbcn

pytestmark = pytest.mark.gpu
BC1, BC3, BC4, BC5 = 54, 58, 60, 62
RGB32F, R8 = 9, 1
FMT_OF_PACK = {0: BC1, 1: BC5, 2: BC5, 3: BC5, 4: BC5, 5: BC3}


def normals(w, h, seed):
    rng = np.random.default_rng(seed)
    n = rng.normal(size=(h, w, 3)).astype(np.float32)
    n[..., 2] = np.abs(n[..., 2]) + 0.5
    return np.ascontiguousarray(n * rng.uniform(0.5, 1.5, (h, w, 1)).astype(np.float32))  # not normalised, as loaded


def normalize(n):
    v = np.ascontiguousarray(n.reshape(-1, 3)).copy()
    hkbcn.check(hkbcn.lib().hkbcn_normalize_vectors(v.ctypes.data_as(C.c_void_p), v.shape[0]))
    return v.reshape(n.shape)


def resample(raw, img, ow, oh, edge, filt):
    h, w, _ = img.shape
    p = hkbcn.ResampleParams(0, edge, edge, filt, filt, ow, oh)
    out = np.zeros((oh, ow, img.shape[2]), img.dtype)
    hkbcn.check(hkbcn.lib().hkbcn_resample_raw_image(raw, np.ascontiguousarray(img).ctypes.data_as(C.c_void_p), w, h, C.byref(p), out.ctypes.data_as(C.c_void_p)))
    return out


def packed_chain(n_unit, pack, edge, filt, mipmapped=True):
    h, w, _ = n_unit.shape
    L = hkbcn.lib()
    tb = L.hkbcn_pack_texel_bytes(pack)
    packed = np.zeros(w * h * tb, np.uint8)
    hkbcn.check(L.hkbcn_pack_normals(pack, np.ascontiguousarray(n_unit).ctypes.data_as(C.c_void_p), packed.ctypes.data_as(C.c_void_p), w * h))
    mips = L.hkbcn_calc_num_mips(w, h) if mipmapped else 1
    return hkbcn.import_chain(FMT_OF_PACK[pack], packed, w, h, mips, cfg=hkbcn.MipConfig(edge, filt, 1, 0)), mips


@pytest.mark.parametrize("pack", [0, 1, 3, 5])
def test_create_normal_map_equals_its_stages(pack):
    L = hkbcn.lib()
    w, h = 128, 64
    n = normalize(normals(w, h, 40 + pack))
    desc = hkbcn.TextureDesc()
    assert L.hkbcn_create_normal_map(n.ctypes.data_as(C.c_void_p), w, h, pack, 1, 1, 2, C.byref(desc), None, 0) == 0  # size query
    need = L.hkbcn_chain_bytes(desc.fmt, desc.width, desc.height, 1, desc.num_mips)
    dst = np.zeros(need, np.uint8)
    hkbcn.check(L.hkbcn_create_normal_map(n.ctypes.data_as(C.c_void_p), w, h, pack, 1, 1, 2, C.byref(desc), dst.ctypes.data_as(C.c_void_p), need))
    want, mips = packed_chain(n, pack, 1, 2)
    assert (desc.fmt, desc.num_mips, desc.storage_flags) == (FMT_OF_PACK[pack], mips, 1)
    assert np.array_equal(dst, want)
    # the reference's refusal
    assert L.hkbcn_create_normal_map(n.ctypes.data_as(C.c_void_p), 100, 64, pack, 1, 0, 0, C.byref(desc), None, 0) == -2
    assert b"power of two" in L.hkbcn_last_error()


@pytest.mark.parametrize("nsize,rsize,pack,bake,dx", [((128, 128), (128, 128), 1, 1, 0), ((100, 90), (128, 128), 0, 2, 1), ((256, 128), (60, 70), 5, 1, 0),
                                                     ((64, 64), (128, 128), 2, 2, 0)])
def test_create_normal_and_roughness_equals_its_stages(nsize, rsize, pack, bake, dx):
    L = hkbcn.lib()
    (nw, nh), (rw0, rh0) = nsize, rsize
    edge, tri = 1, 2
    n_src = normals(nw, nh, 7 + nw)
    rough = np.random.default_rng(rw0).integers(0, 256, (rh0, rw0, 1), dtype=np.uint8)
    s = hkbcn.NormalRoughnessSettings(n_src.ctypes.data_as(C.c_void_p), nw, nh, rough.ctypes.data_as(C.c_void_p), rw0, rh0, pack, dx, edge, bake)
    nd, rd = hkbcn.TextureDesc(), hkbcn.TextureDesc()
    hkbcn.check(L.hkbcn_create_normal_and_roughness(C.byref(s), C.byref(nd), None, 0, C.byref(rd), None, 0))
    nout = np.zeros(L.hkbcn_chain_bytes(nd.fmt, nd.width, nd.height, 1, nd.num_mips), np.uint8)
    rout = np.zeros(L.hkbcn_chain_bytes(rd.fmt, rd.width, rd.height, 1, rd.num_mips), np.uint8)
    hkbcn.check(L.hkbcn_create_normal_and_roughness(C.byref(s), C.byref(nd), nout.ctypes.data_as(C.c_void_p), nout.size, C.byref(rd),
                                                    rout.ctypes.data_as(C.c_void_p), rout.size))

    def pow2(v):
        hi = 1
        while hi < v:
            hi <<= 1
        lo = hi >> 1
        return max(hi if (hi - v) < (v - lo) else lo, 4)
    rw, rh = pow2(rw0), pow2(rh0)
    assert (rd.width, rd.height, rd.fmt) == (rw, rh, BC4)
    # the stages, one by one
    r0 = rough if (rw, rh) == (rw0, rh0) else resample(R8, rough, rw, rh, edge, tri)
    src = n_src.copy()
    if dx:
        src[..., 1] = np.float32(1.0) - src[..., 1]
    nb = src if (nw, nh) == (rw, rh) else resample(RGB32F, src, rw, rh, edge, tri)
    nb = normalize(nb)
    rmips = L.hkbcn_calc_num_mips(rw, rh)
    rchain = hkbcn.generate_mips(BC4, r0, rw, rh, rmips, cfg=hkbcn.MipConfig(edge, tri, 1, 0))
    hkbcn.check(L.hkbcn_bake_roughness(np.ascontiguousarray(nb).ctypes.data_as(C.c_void_p), rchain.ctypes.data_as(C.c_void_p), rw, rh, rmips, bake, tri, edge))
    assert np.array_equal(rout, hkbcn.compress_chain(BC4, rchain, rw, rh, rmips))
    pw, ph = pow2(nw), pow2(nh)
    if (pw, ph) == (rw, rh) or (nw, nh) == (pw, ph):
        nfinal = nb
    else:
        nfinal = normalize(resample(RGB32F, src, pw, ph, edge, tri))
    want, mips = packed_chain(nfinal, pack, edge, tri)
    assert (nd.width, nd.height, nd.num_mips, nd.fmt) == (nfinal.shape[1], nfinal.shape[0], mips, FMT_OF_PACK[pack])
    assert np.array_equal(nout, want)
