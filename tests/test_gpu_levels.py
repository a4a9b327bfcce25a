"""GPU parity of the quality levels of the per-block entry points (row a15: Encode_BC1/2/3(Level), Encode_BC3(bMaxQuality =
false), Encode_BC7(Level), ImageEncoders.cpp:138-169, 207-212) against the reference's rgbcx / bc7enc at the same level:
byte-identical, like the presets -- every level Hork's *_ENCODE_MAX_LEVEL admits (0..18) and both 3-colour options of
Encode_BC1.  Levels that do not exist must be refused, not encoded as something else."""
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


def dark_corpus():
    """blocks with near-black texels (r | g | b < 4: what bTransparentPixelsForBlack looks for) next to bright ones, dark
    gradients, and blocks where the midpoint of two colours is a better third colour than the 1/3 - 2/3 pair"""
    rng = np.random.default_rng(0xB1AC)
    out = []
    for k in range(500):
        px = np.zeros((16, 4), np.uint8)
        px[:, 3] = 255
        kind = k % 5
        if kind == 0:    # a few black texels in a coloured block
            px[:, :3] = rng.integers(40, 256, (16, 3))
            px[rng.permutation(16)[: rng.integers(1, 9)], :3] = rng.integers(0, 4, 3)
        elif kind == 1:  # dark gradient
            base = rng.integers(0, 24, 3)
            px[:, :3] = np.clip(base + rng.integers(0, 12, (16, 1)), 0, 255)
        elif kind == 2:  # two colours and their midpoint
            a, b = rng.integers(0, 256, 3), rng.integers(0, 256, 3)
            pick = rng.integers(0, 3, 16)
            px[:, :3] = np.where(pick[:, None] == 0, a, np.where(pick[:, None] == 1, b, (a + b) // 2))
        elif kind == 3:  # two colours, midpoint, black
            a, b = rng.integers(30, 256, 3), rng.integers(30, 256, 3)
            pick = rng.integers(0, 4, 16)
            px[:, :3] = np.where(pick[:, None] == 0, a, np.where(pick[:, None] == 1, b, np.where(pick[:, None] == 2, (a + b) // 2, 0)))
        else:            # solid and near-solid colours (the 3-colour single-colour tables)
            px[:, :3] = rng.integers(0, 256, 3)
            if k % 10 == 4:
                px[rng.integers(0, 16), :3] = np.clip(px[0, :3].astype(int) + rng.integers(-2, 3, 3), 0, 255)
        out.append(px.ravel())
    return np.stack(out).astype(np.uint8)


def differing(got, want, n, bb=8):
    return np.flatnonzero((got.reshape(n, bb) != want.reshape(n, bb)).any(axis=1))


@pytest.mark.parametrize("level", list(range(6, 19)))
def test_bc1_levels_with_the_orderings_search(level):
    """levels 6..12: likely total orderings (1 .. 32 of them, three selector evaluators); 13..18: both initial endpoint sets,
    six power iterations, exhaustive selectors, 20 .. 256 endpoint search rounds, up to 128 orderings, two iterations at 18"""
    blocks = corpus()
    if level >= 15:
        blocks = blocks[::3]  # the reference takes seconds per thousand blocks up here
    n = blocks.shape[0]
    got = hkbcn.encode_blocks(BC1, blocks, n, LEVEL(level))
    want = refshim.encode_bc1_level_blocks(blocks, n, level)
    bad = differing(got, want, n)
    assert bad.size == 0, f"level {level}: {bad.size} of {n} blocks differ, first {bad[:5]}"
    if level in (6, 9, 13):
        assert not np.array_equal(got, refshim.encode_bc1_level_blocks(blocks, n, level - 1)), "the level must actually change the encoding"


@pytest.mark.parametrize("level", [5, 6, 9, 12, 13, 18])
@pytest.mark.parametrize("allow3,black", [(True, False), (False, True), (True, True)])
def test_bc1_three_colour_options(level, allow3, black):
    """Encode_BC1(.., Level, b3ColorMode, bTransparentPixelsForBlack): 3-colour blocks with their own orderings search, transparent
    black for near-black texels, the 3-colour solid block, search rounds continuing on a 3-colour result"""
    blocks = np.concatenate([dark_corpus(), corpus()[:400]])
    if level >= 13:
        blocks = blocks[::2]
    n = blocks.shape[0]
    flags = LEVEL(level) | (hkbcn.FLAG_BC1_3COLOR if allow3 else 0) | (hkbcn.FLAG_BC1_TRANSPARENT_BLACK if black else 0)
    got = hkbcn.encode_blocks(BC1, blocks, n, flags)
    want = refshim.encode_bc1_level_blocks(blocks, n, level, allow3, black)
    bad = differing(got, want, n)
    assert bad.size == 0, f"level {level} 3colour {allow3} black {black}: {bad.size} of {n} blocks differ, first {bad[:5]}"
    three = (got.reshape(n, 8)[:, 0:2].view(np.uint16) <= got.reshape(n, 8)[:, 2:4].view(np.uint16)).ravel()
    assert three.any(), "the corpus must actually produce 3-colour blocks"
    assert not np.array_equal(got, refshim.encode_bc1_level_blocks(blocks, n, level)), "the options must actually change the encoding"


def test_three_colour_options_are_not_read_below_level_5_or_by_bc2_bc3():
    blocks = np.concatenate([dark_corpus()[:200], corpus()[:200]])
    n = blocks.shape[0]
    opts = hkbcn.FLAG_BC1_3COLOR | hkbcn.FLAG_BC1_TRANSPARENT_BLACK
    for level in (0, 3, 4):
        assert np.array_equal(hkbcn.encode_blocks(BC1, blocks, n, LEVEL(level) | opts), refshim.encode_bc1_level_blocks(blocks, n, level, True, True))
        assert np.array_equal(hkbcn.encode_blocks(BC1, blocks, n, LEVEL(level) | opts), hkbcn.encode_blocks(BC1, blocks, n, LEVEL(level)))
    for fmt in (BC2, BC3):
        assert np.array_equal(hkbcn.encode_blocks(fmt, blocks, n, LEVEL(7) | opts), hkbcn.encode_blocks(fmt, blocks, n, LEVEL(7)))


@pytest.mark.parametrize("level", [6, 10, 14, 18])
def test_bc2_bc3_high_levels_compose_alpha_and_colour(level):
    blocks = corpus()[::4]
    n = blocks.shape[0]
    colour = refshim.encode_bc1_level_blocks(blocks, n, level).reshape(n, 8)
    ref5_bc2 = refshim.compress(BC2, blocks, 4, 4 * n, 1).reshape(n, 16)
    ref5_bc3 = refshim.compress(BC3, blocks, 4, 4 * n, 1).reshape(n, 16)
    got2 = hkbcn.encode_blocks(BC2, blocks, n, LEVEL(level)).reshape(n, 16)
    got3 = hkbcn.encode_blocks(BC3, blocks, n, LEVEL(level)).reshape(n, 16)
    assert np.array_equal(got2[:, :8], ref5_bc2[:, :8]) and np.array_equal(got2[:, 8:], colour)
    assert np.array_equal(got3[:, :8], ref5_bc3[:, :8]) and np.array_equal(got3[:, 8:], colour)
    # never a 3-colour block under BC2 / BC3 (punch-through is not defined there): colour0 > colour1 or the solid-block forms
    c0, c1 = colour[:, 0:2].copy().view(np.uint16).ravel(), colour[:, 2:4].copy().view(np.uint16).ravel()
    assert (c0 >= c1).all()


def test_surface_at_a_high_level():
    """hkbcn_compress (the Compress* shape: a surface, not packed blocks) at level 10 with both options"""
    w, h = 96, 64
    img = synth.albedo_rgba8(w, h, 0xB2000009, "mask").reshape(h, w, 4).copy()
    img[8:24, 8:40, :3] //= 64  # a dark patch
    flags = LEVEL(10) | hkbcn.FLAG_BC1_3COLOR | hkbcn.FLAG_BC1_TRANSPARENT_BLACK
    got = hkbcn.compress(BC1, img, w, h, flags)
    blocks = img.reshape(h // 4, 4, w // 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64)
    want = refshim.encode_bc1_level_blocks(blocks, blocks.shape[0], 10, True, True)
    assert np.array_equal(got, want)


def test_levels_that_do_not_exist_are_refused():
    blocks = corpus()[:4]
    dst = np.zeros(64, np.uint8)
    L = hkbcn.lib()
    for fmt, level in ((BC1, 19), (BC3, 25), (BC7, 5), (BC4, 2)):
        rc = L.hkbcn_encode_blocks(fmt, blocks.ctypes.data, dst.ctypes.data, 4, LEVEL(level))
        assert rc == -6, (fmt, level, rc)  # HKBCN_ERR_UNSUPPORTED
        assert b"does not exist" in L.hkbcn_last_error()
