"""GPU parity at the sizes BASELINE.json's configs name, against the reference itself (oracle/_ref = the unmodified
rgbcx / bc7enc / bc6h_enc / stb_image_resize sources compiled by oracle/Makefile), through the C ABI.

  config 1  BC1 (and BC1 sRGB) 2048^2 with its full mip chain                      byte-identical
  config 2  BC4 4096^2 roughness, BC5 4096^2 tangent-space normals                 byte-identical
  config 3  BC7 4096^2 albedo: level 0 and the whole 11-level chain built by the
            reference's own stbir cascade (Image.cpp:1755-1803)                    byte-identical
  config 4  one 1024^2 RGBA32F cube face -> BC6H                                   byte-identical to the zero-init build,
                                                                                    within 0.05 dB of the as-shipped build
The reference runs on all host threads in contiguous block ranges (ref_shim.cpp: ref_compress)."""
import numpy as np
import pytest

import hkbcn
import refshim
import synth
from test_gpu_bc6h import quality
from test_gpu_mips import ref_cascade

pytestmark = pytest.mark.gpu
BC1, BC1S, BC4, BC5, BC6U, BC7 = 54, 55, 60, 62, 64, 66


def tiled(gen, n, base):
    """an n x n image made of a base x base synthetic tile (numpy synthesis of 4096^2 noise takes minutes; tiling with a
    mirrored band keeps the content, every block position of the large surface is still encoded and compared)"""
    t = gen(base)
    reps = n // base
    img = np.tile(t, (reps, reps, 1))
    img[base:2 * base] = img[base:2 * base, ::-1]
    return np.ascontiguousarray(img)


def same_blocks(got, want, bb):
    g, w = got.reshape(-1, bb), want.reshape(-1, bb)
    assert g.shape == w.shape
    diff = np.flatnonzero((g != w).any(axis=1))
    assert diff.size == 0, f"{diff.size} of {g.shape[0]} blocks differ from the reference, first at {diff[:8]}"


def test_config1_bc1_2048_chain_linear_and_srgb():
    n = 2048
    mips = hkbcn.lib().hkbcn_calc_num_mips(n, n)
    assert mips == 10
    img = tiled(lambda b: synth.albedo_rgba8(b, b, 0xB2000001).reshape(b, b, 4), n, 512)
    for fmt, srgb in ((BC1, False), (BC1S, True)):
        levels = ref_cascade(img, mips, 3, srgb)  # the importers' cascade with the real stbir: Mitchell, wrap
        chain = np.concatenate([l.ravel() for l in levels])
        got = hkbcn.compress_chain(fmt, chain, n, n, mips)
        want = refshim.compress_chain(fmt, chain, n, n, mips)
        assert got.size == 349525 * 8
        same_blocks(got, want, 8)
        # level 0 in, the device cascade, chain out: the same bytes for the linear format always, for the sRGB one when the build
        # carries stbir's sRGB segment table (hkbcn_build_has)
        if not srgb or hkbcn.lib().hkbcn_build_has(hkbcn.BUILD_STBIR_SRGB):
            same_blocks(hkbcn.import_chain(fmt, img, n, n, mips), want, 8)


def test_config2_bc4_bc5_4096():
    n = 4096
    rough = tiled(lambda b: synth.roughness_r8(b, b, 0xB2000002).reshape(b, b, 1), n, 1024)
    same_blocks(hkbcn.compress(BC4, rough, n, n), refshim.compress(BC4, rough, n, n), 8)
    normals = tiled(lambda b: synth.normals_rg8(b, b, 0xB2000003).reshape(b, b, 2), n, 1024)
    same_blocks(hkbcn.compress(BC5, normals, n, n), refshim.compress(BC5, normals, n, n), 16)


def test_config3_bc7_4096_level0_and_whole_chain():
    n = 4096
    mips = hkbcn.lib().hkbcn_calc_num_mips(n, n)
    assert mips == 11
    img = tiled(lambda b: synth.albedo_rgba8(b, b, 0xB2000004).reshape(b, b, 4), n, 1024)
    levels = ref_cascade(img, mips, 3, False)
    chain = np.concatenate([l.ravel() for l in levels])
    got = hkbcn.compress_chain(BC7, chain, n, n, mips)
    want = refshim.compress_chain(BC7, chain, n, n, mips)
    assert got.size == 1398101 * 16
    same_blocks(got[: n * n], want[: n * n], 16)  # level 0
    same_blocks(got, want, 16)                      # every level at its ImageStorage offset
    # the device cascade builds the same chain for this linear format, so hkbcn_import_chain emits the same bytes from level 0 alone
    mine = hkbcn.generate_mips(BC7, img, n, n, mips)
    assert np.array_equal(mine, chain), "device mip cascade differs from the stbir cascade on a linear format"
    same_blocks(hkbcn.import_chain(BC7, img, n, n, mips), want, 16)


def test_config3_bc7_alpha_2048():
    """variant B' of config 3 (smooth alpha mask: modes 5 / 6 / 7), one 2048^2 level"""
    n = 2048
    img = tiled(lambda b: synth.albedo_rgba8(b, b, 0xB2000005, "mask").reshape(b, b, 4), n, 512)
    same_blocks(hkbcn.compress(BC7, img, n, n), refshim.compress(BC7, img, n, n), 16)


def test_config4_bc6h_1024_face():
    n = 1024
    img = tiled(lambda b: synth.sky_rgba32f(b, b, 0xB2000006, 0).reshape(b, b, 4), n, 256)
    got = hkbcn.compress(BC6U, img, n, n)
    same_blocks(got, refshim.compress(BC6U, img, n, n, lib=refshim.ref0()), 16)
    # the as-shipped build reads uninitialised stack (bc6h_enc.h:3134-3150): PSNR bar, on a 512^2 quarter of the face
    q = np.ascontiguousarray(img[:512, :512])
    gq = np.ascontiguousarray(got.reshape(n // 4, n // 4, 16)[:128, :128]).ravel()
    pg, eg = quality(gq, q, 512, 512, BC6U)
    pr, er = quality(refshim.compress(BC6U, q, 512, 512), q, 512, 512, BC6U)
    print(f"BC6H 512^2 quarter: psnr gpu {pg:.3f} dB, as-shipped reference {pr:.3f} dB; err/block {eg:.1f} vs {er:.1f}")
    assert pg >= pr - 0.05
    assert eg <= er * 10 ** (0.05 / 10)
