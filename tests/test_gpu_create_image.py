"""hkbcn_plan_image / hkbcn_create_image / hkbcn_create_roughness_map: the compressed branch of Hk::CreateImage(RawImage...)
(Image.cpp:866-1289) and Hk::CreateRoughnessMap (Image.cpp:2399-2456) as single calls.

The reference composes: format selection by raw format and flags -> channel pass (RGB -> RGBA with alpha 255 / 1.0, BGR ->
RGB) -> ImageStorage::GenerateMipmaps (stbir cascade, alpha handling by the storage flags) -> Compress* per level.  The
oracle does the same with the reference's own stbir and encoders (oracle/_ref); linear formats must match byte for byte."""
import ctypes as C

import numpy as np
import pytest

import hkbcn
import refshim
import synth
from test_gpu_mips import ref_cascade

BC1, BC1S, BC3, BC3S, BC4, BC5, BC6U = 54, 55, 58, 59, 60, 62, 64
R8, R8A, RGB8, BGR8, RGBA8, BGRA8, R32F, R32AF, RGB32F, BGR32F, RGBA32F, BGRA32F = range(1, 13)
NO_ALPHA, PREMULT = 1, 2
HALF, COMPRESS, HDR, SRGB = 1, 2, 4, 8


def plan(raw, w, h, mipmapped, sflags, iflags):
    p = hkbcn.ImagePlan()
    rc = hkbcn.lib().hkbcn_plan_image(raw, w, h, int(mipmapped), sflags, iflags, C.byref(p))
    assert rc == 0
    return p


def test_plan_image_follows_the_reference_table():
    """Image.cpp:878-1056, case by case (host only: no device needed)"""
    for raw, sf, imp, fmt, add, swap, sf_out in (
            (R8, 0, COMPRESS | SRGB, BC4, 0, 0, 0), (R8A, 0, COMPRESS, BC5, 0, 0, 0),
            (RGB8, 0, COMPRESS | SRGB, BC1S, 1, 0, NO_ALPHA), (RGB8, 0, COMPRESS, BC1, 1, 0, NO_ALPHA),
            (BGR8, 0, COMPRESS | SRGB, BC1S, 1, 1, NO_ALPHA),
            (RGBA8, 0, COMPRESS | SRGB, BC3S, 0, 0, 0), (RGBA8, NO_ALPHA, COMPRESS | SRGB, BC1S, 0, 0, NO_ALPHA),
            (RGBA8, 0, COMPRESS, BC3, 0, 0, 0), (BGRA8, PREMULT, COMPRESS, BC3, 0, 1, PREMULT), (BGRA8, NO_ALPHA, COMPRESS, BC1, 0, 1, NO_ALPHA),
            (R32F, 0, COMPRESS | HDR, 0, 0, 0, 0), (R32AF, 0, COMPRESS | HDR, 0, 0, 0, 0),
            (RGB32F, 0, COMPRESS | HDR | HALF, BC6U, 1, 0, NO_ALPHA), (RGB32F, 0, COMPRESS | HALF, 0, 0, 0, 0),
            (BGR32F, 0, COMPRESS | HDR, BC6U, 1, 1, NO_ALPHA), (RGBA32F, 0, COMPRESS | HDR, BC6U, 0, 0, 0), (BGRA32F, 0, COMPRESS | HDR, BC6U, 0, 1, 0),
            (RGBA8, 0, SRGB | HALF, 0, 0, 0, 0)):  # IMAGE_IMPORT_FLAGS_DEFAULT: no compression
        p = plan(raw, 64, 64, True, sf, imp)
        assert (p.format, p.add_alpha, p.swap_rb, p.storage_flags) == (fmt, add, swap, sf_out), (raw, sf, imp)
    assert hkbcn.lib().hkbcn_plan_image(0, 4, 4, 0, 0, COMPRESS, C.byref(hkbcn.ImagePlan())) == -1
    # required size (Image.cpp:1064-1068): closest power of two (ties to the smaller, BaseMath.h:417-423), at least one block
    for w, mip, want in ((600, True, 512), (768, True, 512), (769, True, 1024), (1024, True, 1024), (3, True, 4), (1, True, 4), (22, False, 24),
                         (24, False, 24), (1, False, 4)):
        assert plan(RGBA8, w, w, mip, 0, COMPRESS).required_width == want, (w, mip)


def create_image(raw, img, w, h, cfg, sflags, iflags):
    L = hkbcn.lib()
    p = plan(raw, w, h, cfg is not None, sflags, iflags)
    mips = L.hkbcn_calc_num_mips(w, h) if cfg is not None else 1
    dst = np.zeros(L.hkbcn_chain_bytes(p.format, w, h, 1, mips) if p.format else 16, np.uint8)
    desc = hkbcn.TextureDesc()
    src = np.ascontiguousarray(img)
    rc = L.hkbcn_create_image(raw, src.ctypes.data_as(C.c_void_p), w, h, C.byref(cfg) if cfg is not None else None, sflags, iflags,
                              C.byref(desc), dst.ctypes.data_as(C.c_void_p), dst.size)
    return rc, desc, dst, p


@pytest.mark.gpu
@pytest.mark.parametrize("raw,sflags,fmt", [(RGB8, 0, BC1), (BGR8, 0, BC1), (RGBA8, 0, BC3), (BGRA8, NO_ALPHA, BC1), (RGBA8, PREMULT, BC3),
                                            (R8, 0, BC4), (R8A, 0, BC5)])
def test_create_image_8bit_linear_equals_the_reference_composition(raw, sflags, fmt):
    w, h = 128, 128
    rgba = synth.albedo_rgba8(w, h, 0x51 + raw, "mask").reshape(h, w, 4)
    if raw in (RGB8, BGR8):
        src = np.ascontiguousarray(rgba[..., :3])
    elif raw == R8:
        src = np.ascontiguousarray(rgba[..., :1])
    elif raw == R8A:
        src = np.ascontiguousarray(rgba[..., :2])
    else:
        src = rgba
    cfg = hkbcn.MipConfig(1, 2, 0, 0)  # clamp, triangle
    rc, desc, got, p = create_image(raw, src, w, h, cfg, sflags, COMPRESS)
    assert rc == 0, hkbcn.lib().hkbcn_last_error()
    assert (desc.fmt, desc.num_mips, desc.width, desc.storage_flags) == (fmt, 6, w, p.storage_flags)
    # the reference's channel pass, then its cascade with the alpha handling of ImageStorage::GenerateMipmaps (Image.cpp:767-774)
    img = src
    if raw in (RGB8, BGR8):
        img = np.concatenate([src, np.full((h, w, 1), 255, np.uint8)], -1)
    if raw in (BGR8, BGRA8):
        img = np.ascontiguousarray(img[..., [2, 1, 0, 3]])
    ch = img.shape[-1]
    alpha = ch - 1 if (ch == 4 and not (p.storage_flags & NO_ALPHA)) else -1
    levels = [np.ascontiguousarray(img)]
    for i in range(1, desc.num_mips):
        lh, lw, _ = levels[-1].shape
        levels.append(refshim.stbir_resize(levels[-1], max(lw >> 1, 4), max(lh >> 1, 4), alpha_channel=alpha,
                                           flags=1 if (alpha >= 0 and (p.storage_flags & PREMULT)) else 0, edge=1, filt=2, space=0))
    chain = np.concatenate([l.ravel() for l in levels])
    want = refshim.compress_chain(fmt, chain, w, h, desc.num_mips)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_create_image_hdr_and_srgb_and_refusals():
    L = hkbcn.lib()
    w = h = 64
    sky = synth.sky_rgba32f(w, h, 0x77, 0).reshape(h, w, 4)
    rgb = np.ascontiguousarray(sky[..., :3])
    cfg = hkbcn.MipConfig(0, 0, 0, 0)
    rc, desc, got, p = create_image(RGB32F, rgb, w, h, cfg, 0, COMPRESS | HDR)
    assert rc == 0 and desc.fmt == BC6U and desc.storage_flags == NO_ALPHA
    img = np.concatenate([rgb, np.ones((h, w, 1), np.float32)], -1)
    assert np.array_equal(got, hkbcn.import_chain(BC6U, img, w, h, desc.num_mips, cfg=hkbcn.MipConfig(0, 0, 1, 0)))
    # level 0 against the reference encoder directly
    assert np.array_equal(got[: (w // 4) * (h // 4) * 16], refshim.compress(BC6U, img, w, h, lib=refshim.ref0()))
    # sRGB 8-bit: against the library's own import of the expanded image, and against the reference's cascade + encoder
    rgba = synth.albedo_rgba8(w, h, 0x78, "mask").reshape(h, w, 4)
    rc, desc, got, p = create_image(RGBA8, rgba, w, h, cfg, 0, COMPRESS | SRGB)
    assert rc == 0 and desc.fmt == BC3S
    assert np.array_equal(got, hkbcn.import_chain(BC3S, rgba, w, h, desc.num_mips))
    if hkbcn.lib().hkbcn_build_has(hkbcn.BUILD_STBIR_SRGB):
        # with stbir's own sRGB tables in the build the cascade is stbir's bit for bit, and so is the whole compressed chain
        from test_gpu_mips import ref_cascade
        levels = ref_cascade(rgba, desc.num_mips, 3, True)
        chain = np.concatenate([l.ravel() for l in levels])
        assert np.array_equal(got, refshim.compress_chain(BC3, chain, w, h, desc.num_mips))
    assert np.array_equal(got[: 256 * 16], refshim.compress(BC3, rgba, w, h))
    # no mip config: one level
    rc, desc, got, p = create_image(RGBA8, rgba, w, h, None, 0, COMPRESS)
    assert rc == 0 and desc.num_mips == 1 and np.array_equal(got, refshim.compress(BC3, rgba, w, h))
    # what the reference does not compress, or resamples first, is refused with a reason
    rc, _, _, _ = create_image(RGBA8, rgba, w, h, cfg, 0, SRGB | HALF)
    assert rc == -6 and b"uncompressed" in L.hkbcn_last_error()
    # a size the reference resamples first is resampled here too (tests/test_gpu_resample.py)
    odd = np.zeros((60, 100, 4), np.uint8)
    rc, desc, _, p = create_image(RGBA8, odd, 100, 60, cfg, 0, COMPRESS)
    assert rc == -2 and b"dst holds" in L.hkbcn_last_error()  # the helper sized dst for 100 x 60; the storage is 128 x 64
    assert (p.required_width, p.required_height) == (128, 64)


@pytest.mark.gpu
def test_create_roughness_map_equals_the_reference_composition():
    L = hkbcn.lib()
    w, h = 256, 256
    r = synth.roughness_r8(w, h, 0x99).reshape(h, w, 1)
    mips = L.hkbcn_calc_num_mips(w, h)
    dst = np.zeros(L.hkbcn_chain_bytes(BC4, w, h, 1, mips), np.uint8)
    desc = hkbcn.TextureDesc()
    rc = L.hkbcn_create_roughness_map(r.ctypes.data_as(C.c_void_p), w, h, 1, 3, 2, C.byref(desc), dst.ctypes.data_as(C.c_void_p), dst.size)
    assert rc == 0, L.hkbcn_last_error()
    assert (desc.fmt, desc.num_mips, desc.storage_flags) == (BC4, mips, NO_ALPHA)
    levels = ref_cascade(r, mips, -1, False, edge=3, filt=2)  # wrap, triangle
    want = refshim.compress_chain(BC4, np.concatenate([l.ravel() for l in levels]), w, h, mips)
    assert np.array_equal(dst, want)
    # unmipmapped, and the reference's refusal with its message
    one = np.zeros(L.hkbcn_chain_bytes(BC4, w, h, 1, 1), np.uint8)
    assert L.hkbcn_create_roughness_map(r.ctypes.data_as(C.c_void_p), w, h, 0, 0, 0, C.byref(desc), one.ctypes.data_as(C.c_void_p), one.size) == 0
    assert np.array_equal(one, refshim.compress(BC4, r, w, h)) and desc.num_mips == 1
    assert L.hkbcn_create_roughness_map(r.ctypes.data_as(C.c_void_p), 100, 60, 1, 0, 0, C.byref(desc), one.ctypes.data_as(C.c_void_p), one.size) == -2
    assert b"power of two" in L.hkbcn_last_error()
