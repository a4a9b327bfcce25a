"""hkbcn_resample_raw_image = Hk::ResampleRawImage (Image.cpp:2188-2243): stbir_resize of a raw image, up or down, per-axis edge
mode and filter, optional alpha weighting, on the device.  Checked against the reference's own stb_image_resize.h
(oracle/_ref): linear data must match bit for bit (uint8 and float), sRGB uint8 within +-1 LSB."""
import ctypes as C

import numpy as np
import pytest

import hkbcn
import refshim

R8, R8A, RGB8, BGR8, RGBA8, BGRA8, R32F, R32AF, RGB32F, BGR32F, RGBA32F, BGRA32F = range(1, 13)
HAS_ALPHA, PREMULT, SRGB = 1, 2, 4
CH = {R8: 1, R8A: 2, RGB8: 3, RGBA8: 4, R32F: 1, R32AF: 2, RGB32F: 3, RGBA32F: 4}


def resample(raw, img, ow, oh, flags=0, edge=(0, 0), filt=(0, 0)):
    h, w, _ = img.shape
    p = hkbcn.ResampleParams(flags, edge[0], edge[1], filt[0], filt[1], ow, oh)
    out = np.zeros((oh, ow, img.shape[2]), img.dtype)
    src = np.ascontiguousarray(img)
    hkbcn.check(hkbcn.lib().hkbcn_resample_raw_image(raw, src.ctypes.data_as(C.c_void_p), w, h, C.byref(p), out.ctypes.data_as(C.c_void_p)))
    return out


def test_upsampling_taps_reproduce_stbir_on_the_host():
    """hkbcn_mip_taps (host only) for out > in: stbir's upsampling contributors (stb_image_resize.h:1009-1090)"""
    if not refshim.have_ref():
        pytest.skip("oracle/_ref not built")
    for iw, ih, ow, oh in ((32, 32, 64, 64), (20, 12, 37, 29), (100, 60, 128, 64), (5, 7, 16, 8), (3, 3, 4, 4)):
        rng = np.random.default_rng(iw * 7 + ow)
        for filt in (1, 2, 3, 4, 5):
            for edge in (1, 2, 3, 4):
                img = rng.random((ih, iw, 4)).astype(np.float32)
                ix, wx = hkbcn.mip_taps(iw, ow, filt, edge)
                iy, wy = hkbcn.mip_taps(ih, oh, filt, edge)
                hbuf = np.zeros((ih, ow, 4), np.float32)
                for k in range(ix.shape[1]):
                    hbuf += img[:, ix[:, k], :] * wx[:, k][None, :, None]
                o = np.zeros((oh, ow, 4), np.float32)
                for k in range(iy.shape[1]):
                    o += hbuf[iy[:, k], :, :] * wy[:, k][:, None, None]
                ref = refshim.stbir_resize(img, ow, oh, alpha_channel=-1, flags=0, edge=edge, filt=filt, space=0)
                assert np.array_equal(o, ref), (iw, ih, ow, oh, filt, edge)


@pytest.mark.gpu
@pytest.mark.parametrize("raw", [R8, R8A, RGB8, RGBA8, R32F, R32AF, RGB32F, RGBA32F])
@pytest.mark.parametrize("size", [(48, 40, 64, 64), (100, 60, 128, 64), (64, 64, 24, 36), (30, 50, 47, 20)])
def test_resample_matches_stbir(raw, size):
    iw, ih, ow, oh = size
    ch = CH[raw]
    rng = np.random.default_rng(raw * 100 + iw)
    f32 = raw >= R32F
    img = (rng.random((ih, iw, ch)) * (4.0 if f32 else 255.0)).astype(np.float32 if f32 else np.uint8)
    cases = [(0, -1, 0)]
    if ch in (2, 4):
        cases += [(HAS_ALPHA, ch - 1, 0), (HAS_ALPHA | PREMULT, ch - 1, 1)]
    for k, (flags, alpha, sflags) in enumerate(cases):
        # the reference shim's stbir entry takes one edge mode and one filter for both axes
        edge, filt = 1 + (k + iw) % 4, 1 + (k + ih) % 5
        got = resample(raw, img, ow, oh, flags, (edge, edge), (filt, filt))
        want = refshim.stbir_resize(img, ow, oh, alpha_channel=alpha, flags=sflags, edge=edge, filt=filt, space=0)
        assert np.array_equal(got, want), f"raw {raw} {size} flags {flags} edge {edge} filter {filt}: {np.abs(got.astype(np.float64) - want).max()}"


@pytest.mark.gpu
def test_resample_srgb_and_refusals():
    rng = np.random.default_rng(5)
    img = (rng.random((40, 56, 4)) * 255).astype(np.uint8)
    got = resample(RGBA8, img, 64, 32, HAS_ALPHA | SRGB, (1, 1), (4, 4))
    want = refshim.stbir_resize(img, 64, 32, alpha_channel=3, flags=0, edge=1, filt=4, space=1)
    d = np.abs(got.astype(int) - want.astype(int))
    if hkbcn.lib().hkbcn_build_has(hkbcn.BUILD_STBIR_SRGB):
        assert np.array_equal(got, want), f"{int((d != 0).sum())} values differ from stbir"  # stbir's own decode table and encode segments
    else:
        assert d.max() <= 1 and (d > 0).mean() < 0.1
    L = hkbcn.lib()
    p = hkbcn.ResampleParams(SRGB, 0, 0, 0, 0, 8, 8)
    f = np.zeros((4, 4, 4), np.float32)
    o = np.zeros((8, 8, 4), np.float32)
    assert L.hkbcn_resample_raw_image(RGBA32F, f.ctypes.data_as(C.c_void_p), 4, 4, C.byref(p), o.ctypes.data_as(C.c_void_p)) == -6
    p = hkbcn.ResampleParams(0, 0, 0, 0, 0, 0, 8)
    assert L.hkbcn_resample_raw_image(RGBA32F, f.ctypes.data_as(C.c_void_p), 4, 4, C.byref(p), o.ctypes.data_as(C.c_void_p)) == -2
    assert L.hkbcn_resample_raw_image(0, f.ctypes.data_as(C.c_void_p), 4, 4, C.byref(p), o.ctypes.data_as(C.c_void_p)) == -1


@pytest.mark.gpu
def test_create_image_resamples_to_the_required_size_first():
    """Image.cpp:1064-1096: 100 x 60 RGBA8, mipmapped -> resampled to 128 x 64 with the mip config's edge mode / filter and alpha
    weighting, then the cascade and BC3; equal to the reference composition (stbir + stbir cascade + rgbcx)"""
    from test_gpu_mips import ref_cascade
    L = hkbcn.lib()
    rng = np.random.default_rng(11)
    img = (rng.random((60, 100, 4)) * 255).astype(np.uint8)
    cfg = hkbcn.MipConfig(1, 2, 0, 0)  # clamp, triangle
    mips = L.hkbcn_calc_num_mips(128, 64)
    dst = np.zeros(L.hkbcn_chain_bytes(58, 128, 64, 1, mips), np.uint8)
    desc = hkbcn.TextureDesc()
    rc = L.hkbcn_create_image(RGBA8, img.ctypes.data_as(C.c_void_p), 100, 60, C.byref(cfg), 0, 2, C.byref(desc), dst.ctypes.data_as(C.c_void_p), dst.size)
    assert rc == 0, L.hkbcn_last_error()
    assert (desc.width, desc.height, desc.num_mips, desc.fmt) == (128, 64, mips, 58)
    big = refshim.stbir_resize(img, 128, 64, alpha_channel=3, flags=0, edge=1, filt=2, space=0)
    levels = ref_cascade(big, mips, 3, False, edge=1, filt=2)
    want = refshim.compress_chain(58, np.concatenate([l.ravel() for l in levels]), 128, 64, mips)
    assert np.array_equal(dst, want)
    # without mips: aligned up to the block size (22 x 10 -> 24 x 12), wrap / mitchell
    small = (rng.random((10, 22, 3)) * 255).astype(np.uint8)
    dst = np.zeros(L.hkbcn_chain_bytes(54, 24, 12, 1, 1), np.uint8)
    rc = L.hkbcn_create_image(RGB8, small.ctypes.data_as(C.c_void_p), 22, 10, None, 0, 2, C.byref(desc), dst.ctypes.data_as(C.c_void_p), dst.size)
    assert rc == 0 and (desc.width, desc.height, desc.num_mips, desc.fmt) == (24, 12, 1, 54)
    up = refshim.stbir_resize(small, 24, 12, alpha_channel=-1, flags=0, edge=3, filt=5, space=0)
    rgba = np.concatenate([up, np.full((12, 24, 1), 255, np.uint8)], -1)
    assert np.array_equal(dst, refshim.compress(54, rgba, 24, 12))
