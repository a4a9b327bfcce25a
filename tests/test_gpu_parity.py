"""GPU parity tests: the CUDA path (through the C ABI, libhkbcn.so) against the reference codecs
(oracle/_ref/libhkref.so = the unmodified rgbcx / bcdec / bc7decomp sources behind oracle/ref_shim.cpp).

Bar (BASELINE.json north_star): bit-exact for BC1..BC5 encode and every decode."""
import numpy as np
import pytest

import hkbcn
import refshim
import synth

pytestmark = pytest.mark.gpu

BC1, BC2, BC3, BC4, BC5, BC6U, BC6S, BC7 = 54, 56, 58, 60, 62, 64, 65, 66


def blocks_to_surface(blocks, bpp):
    """n packed 4x4 blocks (n*16*bpp bytes) == a 4-wide, 4n-tall surface: same bytes."""
    return blocks


def assert_same(a, b, what, blk):
    a = np.asarray(a).ravel()
    b = np.asarray(b).ravel()
    assert a.shape == b.shape
    if not np.array_equal(a, b):
        bad = np.nonzero((a.reshape(-1, blk) != b.reshape(-1, blk)).any(axis=1))[0]
        raise AssertionError(f"{what}: {len(bad)} of {len(a) // blk} blocks differ, first at {bad[0]}: "
                             f"got {a.reshape(-1, blk)[bad[0]].tolist()} want {b.reshape(-1, blk)[bad[0]].tolist()}")


# ----------------------------------------------------------------------------- BC4 / BC5
def adversarial_r8():
    rng = np.random.default_rng(1234)
    blocks = []
    for v in range(256):                     # all solid values (rgbcx.cpp:2744-2754)
        blocks.append(np.full(16, v, np.uint8))
    for lo in (0, 1, 2, 3, 4, 250, 252, 255):  # clamp region of the +-3 search (rgbcx.cpp:2806-2807)
        for hi in (0, 1, 2, 3, 5, 251, 253, 254, 255):
            if lo != hi:
                blocks.append(np.where(rng.random(16) < 0.5, lo, hi).astype(np.uint8))
    for _ in range(300):                     # narrow ranges: many tied trials
        base = rng.integers(0, 250)
        blocks.append((base + rng.integers(0, rng.integers(2, 7), 16)).astype(np.uint8))
    for _ in range(300):                     # ramps with 0 / 255 outliers (6-value mode wins)
        b = np.linspace(rng.integers(20, 100), rng.integers(120, 230), 16).astype(np.uint8)
        b[rng.integers(0, 16, 3)] = rng.choice([0, 255])
        blocks.append(b)
    blocks.append(rng.integers(0, 256, 16 * 1500).astype(np.uint8))
    return np.concatenate(blocks)


def test_bc4_adversarial_blocks():
    src = adversarial_r8()
    n = src.size // 16
    got = hkbcn.encode_blocks(BC4, src, n)
    want = refshim.compress(BC4, src, 4, 4 * n)
    assert_same(got, want, "BC4 adversarial", 8)


def test_bc5_random_blocks():
    n = 2000
    src = synth.random_blocks(n, 2, 77)
    src[: 16 * 2 * 256] = np.repeat(np.arange(256, dtype=np.uint8), 32)  # solid in both channels
    got = hkbcn.encode_blocks(BC5, src, n)
    want = refshim.compress(BC5, src, 4, 4 * n)
    assert_same(got, want, "BC5 random", 16)


@pytest.mark.parametrize("w,h", [(256, 256), (64, 32), (20, 12), (8, 8), (4, 4), (512, 4), (4, 512), (36, 100)])
def test_bc4_bc5_surfaces(w, h):
    r = synth.roughness_r8(w, h, 0xB2000002)
    assert_same(hkbcn.compress(BC4, r, w, h), refshim.compress(BC4, r, w, h), f"BC4 {w}x{h}", 8)
    n = synth.normals_rg8(w, h, 0xB2000003)
    assert_same(hkbcn.compress(BC5, n, w, h), refshim.compress(BC5, n, w, h), f"BC5 {w}x{h}", 16)


def test_bc4_bc5_fast_mode():
    n = 1000
    src = adversarial_r8()[: n * 16]
    assert_same(hkbcn.encode_blocks(BC4, src, n, hkbcn.FLAG_FAST), refshim.encode_bc45_fast(BC4, src, n), "BC4 non-hq", 8)
    src2 = synth.random_blocks(n, 2, 5)
    assert_same(hkbcn.encode_blocks(BC5, src2, n, hkbcn.FLAG_FAST), refshim.encode_bc45_fast(BC5, src2, n), "BC5 non-hq", 16)


# ----------------------------------------------------------------------------- BC1 / BC2 / BC3
def adversarial_rgba8():
    rng = np.random.default_rng(99)
    blocks = []
    lat = np.arange(0, 256, 17)
    for r in lat:                             # solid colours on a 16^3 lattice (rgbcx.cpp:629-691)
        for g in lat:
            for b in lat:
                blocks.append(np.tile(np.array([r, g, b, 255], np.uint8), 16))
    for v in range(256):                      # solid grays incl. black / white
        blocks.append(np.tile(np.array([v, v, v, v], np.uint8), 16))
    for _ in range(400):                      # grayscale ramps / near-solid grays (rgbcx.cpp:1818-1836)
        a, b = rng.integers(0, 256, 2)
        if rng.random() < 0.3:
            b = min(255, a + 1)
        g = np.linspace(a, b, 16).astype(np.uint8)[rng.permutation(16)]
        blocks.append(np.stack([g, g, g, rng.integers(0, 256, 16).astype(np.uint8)], 1).ravel())
    for _ in range(600):                      # two-colour blocks
        c0, c1 = rng.integers(0, 256, 4), rng.integers(0, 256, 4)
        m = rng.random(16) < rng.random()
        blocks.append(np.where(m[:, None], c0, c1).astype(np.uint8).ravel())
    for _ in range(600):                      # colour ramps + noise
        c0, c1 = rng.integers(0, 256, 4), rng.integers(0, 256, 4)
        t = rng.random(16)[:, None]
        blocks.append(np.clip(c0 * (1 - t) + c1 * t + rng.normal(0, 4, (16, 4)), 0, 255).astype(np.uint8).ravel())
    for a in (255, 254, 0):                   # alpha cases for BC2/BC3
        px = rng.integers(0, 256, (16, 4)).astype(np.uint8)
        px[:, 3] = a
        blocks.append(px.ravel())
    blocks.append(rng.integers(0, 256, 64 * 3000).astype(np.uint8))
    return np.concatenate(blocks)


@pytest.mark.parametrize("fmt,blk", [(BC1, 8), (BC2, 16), (BC3, 16)])
def test_bc123_adversarial_blocks(fmt, blk):
    src = adversarial_rgba8()
    n = src.size // 64
    got = hkbcn.encode_blocks(fmt, src, n)
    want = refshim.compress(fmt, src, 4, 4 * n)
    assert_same(got, want, f"fmt {fmt} adversarial", blk)


@pytest.mark.parametrize("w,h", [(256, 256), (128, 64), (20, 12), (8, 8), (4, 4), (4, 64)])
@pytest.mark.parametrize("fmt,blk", [(BC1, 8), (BC2, 16), (BC3, 16)])
def test_bc123_surfaces(fmt, blk, w, h):
    img = synth.albedo_rgba8(w, h, 0xB2000001, alpha="mask" if fmt != BC1 else "opaque")
    assert_same(hkbcn.compress(fmt, img, w, h), refshim.compress(fmt, img, w, h), f"fmt {fmt} {w}x{h}", blk)


def test_bc1_chain_matches_per_level():
    """hkbcn_compress_chain: ImageStorage layout (Image.cpp:498-574), one launch for all levels."""
    w = h = 128
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    assert mips == 6
    levels = [synth.albedo_rgba8(lw, lh, 0xB2000001 + i) for i, (lw, lh) in enumerate(synth.mip_dims(w, h, mips))]
    src = np.concatenate([l.ravel() for l in levels])
    got = hkbcn.compress_chain(BC1, src, w, h, mips)
    want = refshim.compress_chain(BC1, src, w, h, mips)
    assert_same(got, want, "BC1 chain", 8)
    got4 = hkbcn.compress_chain(BC4, src[: sum(lw * lh for lw, lh in synth.mip_dims(w, h, mips))], w, h, mips)
    want4 = refshim.compress_chain(BC4, src, w, h, mips)
    assert_same(got4, want4, "BC4 chain", 8)


# ----------------------------------------------------------------------------- decode
def force_modes_bc7(blocks):
    b = blocks.reshape(-1, 16).copy()
    for i in range(len(b)):
        m = i % 9
        v = int(b[i, 0])
        b[i, 0] = 0 if m == 8 else ((v | (1 << m)) & ~((1 << m) - 1)) & 0xFF   # mode m; 8 = reserved (byte0 == 0)
    return b.ravel()


def force_modes_bc6h(blocks):
    codes = [0, 1, 2, 6, 10, 14, 18, 22, 26, 30, 3, 7, 11, 15, 19, 23, 27, 31]  # 14 modes + 4 reserved
    b = blocks.reshape(-1, 16).copy()
    for i in range(len(b)):
        c = codes[i % len(codes)]
        v = int(b[i, 0])
        b[i, 0] = (v & ~3 | c) & 0xFF if c < 2 else (v & ~31 | c) & 0xFF
    return b.ravel()


@pytest.mark.parametrize("fmt,blk", [(BC1, 8), (BC2, 16), (BC3, 16), (BC4, 8), (BC5, 16), (BC7, 16)])
def test_decode_random_blocks(fmt, blk):
    w, h = 256, 128
    n = (w // 4) * (h // 4)
    blocks = synth.random_blocks(n, 1, 4242 + fmt)[: n * blk] if blk == 16 else synth.random_blocks(n, 1, 4242 + fmt)[: n * blk]
    if fmt == BC7:
        blocks = force_modes_bc7(blocks)
    got = hkbcn.decompress(fmt, blocks, w, h)
    want = refshim.decompress(fmt, blocks, w, h)
    tb = got.size // (w * h)
    assert_same(got, want, f"decode fmt {fmt}", tb * 4)


@pytest.mark.parametrize("fmt", [BC6U, BC6S])
@pytest.mark.parametrize("f32", [False, True])
def test_decode_bc6h_random_blocks(fmt, f32):
    w, h = 256, 144
    n = (w // 4) * (h // 4)
    blocks = force_modes_bc6h(synth.random_blocks(n, 1, 777 + fmt))
    got = hkbcn.decompress(fmt, blocks, w, h, hkbcn.FLAG_DST_F32 if f32 else 0)
    want = refshim.decompress(fmt, blocks, w, h, f32out=f32)
    assert_same(got, want, f"decode BC6H fmt {fmt} f32={f32}", 12 if f32 else 6)


def test_decode_bc6h_each_mode_reported():
    """per-mode breakdown so a wrong header layout names its mode"""
    codes = [0, 1, 2, 6, 10, 14, 18, 22, 26, 30, 3, 7, 11, 15]
    n = 512
    for mi, c in enumerate(codes):
        b = synth.random_blocks(n, 1, 9000 + mi).reshape(-1, 16).copy()
        b[:, 0] = ((b[:, 0].astype(np.int32) & (~3 if c < 2 else ~31)) | c).astype(np.uint8)
        for fmt in (BC6U, BC6S):
            got = hkbcn.decompress(fmt, b.ravel(), 4, 4 * n)
            want = refshim.decompress(fmt, b.ravel(), 4, 4 * n)
            assert np.array_equal(got, want), f"BC6H mode #{mi + 1} (code {c:05b}) signed={fmt == BC6S} differs"


@pytest.mark.parametrize("fmt,blk,tb", [(BC1, 8, 4), (BC3, 16, 4), (BC4, 8, 1), (BC5, 16, 2), (BC7, 16, 4), (BC6U, 16, 6)])
def test_decode_odd_strides_and_sizes(fmt, blk, tb):
    """unaligned RowStride (Decode_BC*(.., RowStride)) and narrow surfaces"""
    for (w, h) in [(4, 4), (8, 4), (12, 20), (36, 8)]:
        n = (w // 4) * (h // 4)
        blocks = synth.random_blocks(n, 1, 31 + w + fmt)[: n * blk]
        if fmt == BC7:
            blocks = force_modes_bc7(blocks)
        if fmt == BC6U:
            blocks = force_modes_bc6h(blocks)
        stride = w * tb + 6   # not a multiple of 4/8/16
        got = hkbcn.decompress(fmt, blocks, w, h, row_stride=stride)
        want = refshim.decompress(fmt, blocks, w, h, row_stride=stride)
        got = got.reshape(h, stride)[:, : w * tb]
        want = want.reshape(h, stride)[:, : w * tb]
        assert np.array_equal(got, want), f"fmt {fmt} {w}x{h} stride {stride}"


def test_decode_blocks_api():
    n = 100
    blocks = force_modes_bc7(synth.random_blocks(n, 1, 5)[: n * 16])
    got = hkbcn.decode_blocks(BC7, blocks, n, 16)
    want = refshim.decompress(BC7, blocks, 4, 4 * n)
    assert np.array_equal(got, want)


def test_full_size_bit_exact_and_roundtrip():
    """BASELINE config sizes: 2048^2 BC1 / BC4 / BC5 encodes are bit-exact against the reference run over all host
    threads, and the decoded surfaces stay close to the source (encode -> decode round trip)."""
    w = h = 2048
    r = synth.roughness_r8(w, h, 0xB2000002)
    b4 = hkbcn.compress(BC4, r, w, h)
    assert_same(b4, refshim.compress(BC4, r, w, h), "BC4 2048^2", 8)
    d4 = hkbcn.decompress(BC4, b4, w, h).reshape(h, w)
    assert refshim.psnr(d4, r) > 45.0
    assert np.array_equal(d4.ravel(), refshim.decompress(BC4, b4, w, h))
    n = synth.normals_rg8(w, h, 0xB2000003)
    assert_same(hkbcn.compress(BC5, n, w, h), refshim.compress(BC5, n, w, h), "BC5 2048^2", 16)
    a = synth.albedo_rgba8(w, h, 0xB2000001)
    b1 = hkbcn.compress(BC1, a, w, h)
    assert_same(b1, refshim.compress(BC1, a, w, h), "BC1 2048^2", 8)
    d1 = hkbcn.decompress(BC1, b1, w, h)
    assert np.array_equal(d1, refshim.decompress(BC1, b1, w, h))
    assert refshim.psnr(d1.reshape(h, w, 4)[..., :3], a.reshape(h, w, 4)[..., :3]) > 30.0


def test_batch_api_matches_single():
    jobs = []
    for i in range(6):
        w, h = 64 << (i % 2), 64
        if i % 3 == 0:
            jobs.append((BC1, synth.albedo_rgba8(w, h, 100 + i), w, h, 0))
        elif i % 3 == 1:
            jobs.append((BC5, synth.normals_rg8(w, h, 100 + i), w, h, 0))
        else:
            jobs.append((BC4, synth.roughness_r8(w, h, 100 + i), w, h, 0))
    outs = hkbcn.compress_batch(jobs)
    for (fmt, src, w, h, _), o in zip(jobs, outs):
        assert np.array_equal(o, refshim.compress(fmt, src, w, h))


@pytest.mark.parametrize("n_gpus", [0, -3, -7])
def test_banded_chain_equals_single_device(n_gpus):
    """one surface split over GPUs in bands of whole block rows.  n_gpus = 0 uses every GPU on the box; negative
    counts deal that many bands over the GPUs there are, so the banded path also runs on a one-GPU box.  Bands of a level that
    has fewer block rows than bands are empty for some devices (the 4^2 and 8^2 levels here)."""
    w, h = 128, 64
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    for fmt, gen in ((BC1, lambda lw, lh: synth.albedo_rgba8(lw, lh, 21)), (BC5, lambda lw, lh: synth.normals_rg8(lw, lh, 22)),
                     (BC7, lambda lw, lh: synth.albedo_rgba8(lw, lh, 23, "mask"))):
        src = np.concatenate([np.asarray(gen(lw, lh)).ravel() for lw, lh in synth.mip_dims(w, h, mips)])
        want = hkbcn.compress_chain(fmt, src, w, h, mips)
        got = hkbcn.compress_chain_multi(fmt, src, w, h, mips, n_gpus)
        assert np.array_equal(got, want), f"fmt {fmt}, n_gpus {n_gpus}"
