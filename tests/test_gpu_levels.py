"""GPU parity of the quality levels of the per-block entry points (row a15: Encode_BC1/2/3(Level), Encode_BC3(bMaxQuality =
false), Encode_BC7(Level), ImageEncoders.cpp:138-169, 207-212) against the reference's rgbcx / bc7enc at the same level:
byte-identical, like the presets.  Levels that are not built must be refused, not encoded as something else."""
import numpy as np
import pytest

import hkbcn
import refshim
import synth
from test_gpu_parity import adversarial_rgba8

pytestmark = pytest.mark.gpu
BC1, BC2, BC3, BC4, BC7 = 54, 56, 58, 60, 66


def LEVEL(n):
    return (n + 1) << 4


def corpus():
    img = synth.albedo_rgba8(64, 64, 0xB2000001, "mask").reshape(16, 4, 16, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64)
    gray = np.repeat(synth.roughness_r8(32, 32, 5).reshape(8, 4, 8, 4).transpose(0, 2, 1, 3).reshape(-1, 16, 1), 4, axis=2).reshape(-1, 64)
    gray[:, 3::4] = 255
    return np.concatenate([adversarial_rgba8()[: 64 * 600].reshape(-1, 64), img, gray, synth.random_blocks(300, 4, 17).reshape(-1, 64)]).astype(np.uint8)


@pytest.mark.parametrize("level", [0, 1, 2, 3, 4, 5])
def test_bc1_levels(level):
    blocks = corpus()
    n = blocks.shape[0]
    got = hkbcn.encode_blocks(BC1, blocks, n, LEVEL(level))
    want = refshim.encode_bc1_level_blocks(blocks, n, level)
    bad = np.flatnonzero((got.reshape(n, 8) != want.reshape(n, 8)).any(axis=1))
    assert bad.size == 0, f"level {level}: {bad.size} of {n} blocks differ, first {bad[:5]}"
    if level in (0, 1, 2, 4):
        assert not np.array_equal(got, refshim.encode_bc1_level_blocks(blocks, n, 5)), "the level must actually change the encoding"


@pytest.mark.parametrize("level", [0, 1, 3, 4])
def test_bc2_bc3_levels_compose_alpha_and_colour(level):
    blocks = corpus()
    n = blocks.shape[0]
    colour = refshim.encode_bc1_level_blocks(blocks, n, level).reshape(n, 8)
    ref5_bc2 = refshim.compress(BC2, blocks, 4, 4 * n, 1).reshape(n, 16)
    ref5_bc3 = refshim.compress(BC3, blocks, 4, 4 * n, 1).reshape(n, 16)
    got2 = hkbcn.encode_blocks(BC2, blocks, n, LEVEL(level)).reshape(n, 16)
    got3 = hkbcn.encode_blocks(BC3, blocks, n, LEVEL(level)).reshape(n, 16)
    assert np.array_equal(got2[:, :8], ref5_bc2[:, :8]) and np.array_equal(got2[:, 8:], colour)   # Encode_BC2: 4-bit alpha + encode_bc1(Level)
    assert np.array_equal(got3[:, :8], ref5_bc3[:, :8]) and np.array_equal(got3[:, 8:], colour)   # encode_bc3_hq: bc4_hq alpha + encode_bc1(Level)


def test_bc3_without_max_quality():
    """Encode_BC3(.., bMaxQuality = false) = rgbcx::encode_bc3: the non-hq encode_bc4 on alpha (rgbcx.cpp:2886-2903)"""
    blocks = corpus()
    n = blocks.shape[0]
    got = hkbcn.encode_blocks(BC3, blocks, n, hkbcn.FLAG_FAST | LEVEL(5)).reshape(n, 16)
    alpha = np.ascontiguousarray(blocks.reshape(n, 16, 4)[:, :, 3])
    want_a = refshim.encode_bc45_fast(BC4, alpha, n).reshape(n, 8)
    want_c = refshim.encode_bc1_level_blocks(blocks, n, 5).reshape(n, 8)
    assert np.array_equal(got[:, :8], want_a)
    assert np.array_equal(got[:, 8:], want_c)


@pytest.mark.parametrize("level", [0, 1, 2, 3, 4])
def test_bc7_levels(level):
    img = np.concatenate([synth.albedo_rgba8(64, 64, 0xB2000004 + k, a).reshape(16, 4, 16, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64)
                          for k, a in enumerate(("opaque", "mask", "noise"))])
    blocks = np.concatenate([img, synth.random_blocks(200, 4, 3).reshape(-1, 64)]).astype(np.uint8)
    n = blocks.shape[0]
    got = hkbcn.encode_blocks(BC7, blocks, n, LEVEL(level))
    want = refshim.encode_bc7_level_blocks(blocks, n, level)
    bad = np.flatnonzero((got.reshape(n, 16) != want.reshape(n, 16)).any(axis=1))
    assert bad.size == 0, f"uber level {level}: {bad.size} of {n} blocks differ, first {bad[:5]}"
    if level < 4:
        assert not np.array_equal(got, hkbcn.encode_blocks(BC7, blocks, n)), "the level must actually change the encoding"


def test_levels_that_are_not_built_are_refused():
    blocks = corpus()[:4]
    dst = np.zeros(64, np.uint8)
    L = hkbcn.lib()
    for fmt, level in ((BC1, 6), (BC1, 18), (BC3, 10), (BC7, 5), (BC4, 2)):
        rc = L.hkbcn_encode_blocks(fmt, blocks.ctypes.data, dst.ctypes.data, 4, LEVEL(level))
        assert rc == -6, (fmt, level, rc)  # HKBCN_ERR_UNSUPPORTED
        assert b"not built" in L.hkbcn_last_error()
