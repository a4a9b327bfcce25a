"""GPU parity tests of the device-side mip cascade against the reference's own stbir_resize
(ThirdParty/stb/stb_image_resize.h v0.97, built into oracle/_ref) driven as the importers drive it: level i from level
i-1, channels / alpha channel / colour space by format (Hork/Image/Image.cpp:1755-1803, :1852-1899).

Bars: LINEAR uint8 formats (BC1 / BC3 / BC4 / BC5 / BC7 UNORM) are bit-identical to stbir on every level -- the float sums
follow stbir's order without contraction -- which is what lets bench.py and hkbcn_import_chain claim the reference's
blocks byte for byte.  sRGB formats: +-1 LSB per channel and >= 50 dB (the contract SURVEY.md 8f states: the sRGB decode
table is computed instead of typed in and the encode is the exact transfer function instead of stbir's piecewise-linear
table).  Float chains within 1e-6 relative."""
import numpy as np
import pytest

import hkbcn
import refshim
import synth

pytestmark = pytest.mark.gpu
BC1, BC1S, BC3, BC4, BC5, BC6U, BC7, BC7S = 54, 55, 58, 60, 62, 64, 66, 67
PSNR_TOL_DB = 0.05


def ref_cascade(img, mips, alpha_channel, srgb, edge=3, filt=5):
    """the importers' loop with the real stbir_resize: list of levels (h, w, c)"""
    levels = [img]
    for i in range(1, mips):
        h, w, _ = levels[-1].shape
        levels.append(refshim.stbir_resize(levels[-1], max(w >> 1, 4), max(h >> 1, 4), alpha_channel=alpha_channel, flags=0,
                                           edge=edge, filt=filt, space=1 if srgb else 0))
    return levels


def split_chain(buf, w, h, mips, ch, dtype=np.uint8):
    out, off = [], 0
    a = buf.view(dtype)
    for i in range(mips):
        lw, lh = max(w >> i, 4), max(h >> i, 4)
        out.append(a[off:off + lw * lh * ch].reshape(lh, lw, ch))
        off += lw * lh * ch
    return out


@pytest.mark.parametrize("fmt,ch,alpha,srgb,kind", [(BC1, 4, 3, False, "opaque"), (BC7S, 4, 3, True, "mask"), (BC7, 4, 3, False, "noise"),
                                                    (BC1S, 4, 3, True, "opaque"), (BC4, 1, -1, False, ""), (BC5, 2, -1, False, "")])
def test_uint8_cascade_matches_stbir(fmt, ch, alpha, srgb, kind):
    w, h = 256, 128
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    if ch == 4:
        img = synth.albedo_rgba8(w, h, 0xB2000005, kind).reshape(h, w, 4)
    elif ch == 1:
        img = synth.roughness_r8(w, h, 2).reshape(h, w, 1)
    else:
        img = synth.normals_rg8(w, h, 3).reshape(h, w, 2)
    got = split_chain(hkbcn.generate_mips(fmt, img, w, h, mips), w, h, mips, ch)
    want = ref_cascade(img, mips, alpha, srgb)
    assert np.array_equal(got[0], img)
    for i in range(1, mips):
        d = np.abs(got[i].astype(int) - want[i].astype(int))
        if srgb and not hkbcn.lib().hkbcn_build_has(hkbcn.BUILD_STBIR_SRGB):
            # a library built without the engine tree encodes sRGB with the exact transfer function, not stbir's segment table
            assert d.max() <= 1, f"level {i}: max diff {d.max()} LSB"
            assert refshim.psnr(got[i], want[i]) >= 50.0, f"level {i}: {refshim.psnr(got[i], want[i]):.2f} dB"
        else:
            assert np.array_equal(got[i], want[i]), f"level {i}: {int((d != 0).sum())} values differ from stbir (max {d.max()} LSB)"
    frac = np.mean([np.mean(got[i] != want[i]) for i in range(1, mips)])
    print(f"fmt {fmt}: mean fraction of channel values off by one LSB over the chain: {frac:.5f}")


@pytest.mark.parametrize("edge,filt", [(1, 1), (2, 2), (3, 3), (1, 4), (4, 5)])
def test_uint8_edge_modes_and_filters(edge, filt):
    w = h = 64
    img = synth.albedo_rgba8(w, h, 0xB2000001, "mask").reshape(h, w, 4)
    cfg = hkbcn.MipConfig(edge, filt, 0, 0)
    got = split_chain(hkbcn.generate_mips(BC7, img, w, h, 5, cfg), w, h, 5, 4)
    want = ref_cascade(img, 5, 3, False, edge, filt)
    for i in range(1, 5):
        assert np.array_equal(got[i], want[i]), f"edge {edge} filter {filt} level {i}"


@pytest.mark.parametrize("w,h", [(512, 512), (1024, 256), (64, 512)])
def test_uint8_cascade_larger_and_non_square(w, h):
    """tile edges of the separable kernel, non-square chains (the short axis clamps at 4 while the long one still halves)"""
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    img = synth.albedo_rgba8(w, h, 0xB2000009, "mask").reshape(h, w, 4)
    for fmt, alpha, cfg in ((BC7, 3, None), (BC7, -1, hkbcn.MipConfig(3, 5, 1, 0)), (BC3, 3, hkbcn.MipConfig(1, 2, 0, 0))):
        edge, filt = (cfg.edge_mode or 3, cfg.filter or 5) if cfg else (3, 5)
        got = split_chain(hkbcn.generate_mips(fmt, img, w, h, mips, cfg), w, h, mips, 4)
        want = ref_cascade(img, mips, alpha, False, edge, filt)
        for i in range(1, mips):
            assert np.array_equal(got[i], want[i]), f"fmt {fmt} alpha {alpha} level {i}"
    r = synth.roughness_r8(w, h, 4).reshape(h, w, 1)
    got = split_chain(hkbcn.generate_mips(BC4, r, w, h, mips, hkbcn.MipConfig(3, 2, 0, 0)), w, h, mips, 1)
    want = ref_cascade(r, mips, -1, False, 3, 2)
    for i in range(1, mips):
        assert np.array_equal(got[i], want[i]), f"BC4 level {i}"


def test_float_cascade_matches_stbir():
    w = h = 128
    img = synth.sky_rgba32f(w, h, 0xB2000006, 0)
    mips = 6
    got = split_chain(hkbcn.generate_mips(BC6U, img, w, h, mips), w, h, mips, 4, np.float32)
    want = ref_cascade(img, mips, -1, False)
    for i in range(1, mips):
        assert np.array_equal(got[i].view(np.uint32), want[i].view(np.uint32)), f"level {i}: {int((got[i] != want[i]).sum())} floats differ from stbir's bit patterns"


def test_import_chain_equals_reference_cascade_and_encoder():
    """level 0 in, compressed chain out: the bytes of the stbir cascade + the reference encoder, level by level"""
    w = h = 256
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    img = synth.albedo_rgba8(w, h, 0xB2000004, "opaque").reshape(h, w, 4)
    got = hkbcn.import_chain(BC7, img, w, h, mips)
    levels = ref_cascade(img, mips, 3, False)
    off = 0
    for i, lv in enumerate(levels):
        lh, lw, _ = lv.shape
        nb = (lw // 4) * (lh // 4) * 16
        assert np.array_equal(got[off:off + nb], refshim.compress(BC7, lv, lw, lh)), f"level {i}"
        off += nb
    assert off == got.size
    # level 0 is untouched by the cascade: its blocks are the reference's, byte for byte
    assert np.array_equal(got[: (w // 4) * (h // 4) * 16], refshim.compress(BC7, img, w, h))


@pytest.mark.parametrize("fmt,ch", [(BC4, 1), (BC5, 2)])
@pytest.mark.parametrize("edge", [1, 2, 3, 4])
def test_r8_rg8_triangle_chain_at_size(fmt, ch, edge):
    """the four-tap R8 / RG8 kernel (levels of at least 64 x 16 outputs) and the kernels below it, every edge mode: interior
    tiles (aligned word staging) and edge tiles (through the maps) against stbir, bit for bit"""
    w, h = 1024, 512
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    img = (synth.roughness_r8(w, h, 21).reshape(h, w, 1) if ch == 1 else synth.normals_rg8(w, h, 22).reshape(h, w, 2))
    cfg = hkbcn.MipConfig(edge, 2, 0, 0)
    got = split_chain(hkbcn.generate_mips(fmt, img, w, h, mips, cfg), w, h, mips, ch)
    want = ref_cascade(img, mips, -1, False, edge, 2)
    for i in range(1, mips):
        assert np.array_equal(got[i], want[i]), f"level {i}: {int((got[i] != want[i]).sum())} values differ"
