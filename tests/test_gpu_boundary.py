"""GPU tests of the drop-in boundary itself: the C++ shim executed (not just nm-checked), widths that are
not multiples of 4, pointer-alignment and chain-dimension checks of the *_dev entry points, the committed golden encodes,
and the NaN / Inf corner of the BC6H corpus."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import hkbcn
import refshim
import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "hork-source_b200")
BC1, BC2, BC3, BC4, BC5, BC6U, BC6S, BC7 = 54, 56, 58, 60, 62, 64, 65, 66
GOLD = os.path.join(ROOT, "tests", "golden", "bcn_golden.npz")


def test_cpp_shim_runs_and_matches_reference(tmp_path):
    """links libhkbcn_hork.so, stores &Hk::TextureBlockCompression::CompressBC7 in the void(*)(void const*, void*, uint32_t,
    uint32_t) that Image.cpp:1718 uses, calls all seven Compress* and the sixteen block functions on malloc'd buffers"""
    exe = str(tmp_path / "shim_check")
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", exe, os.path.join(ROOT, "tests", "shim_check.cpp"),
                    "-L" + PKG, "-lhkbcn_hork", "-lhkbcn", "-ldl", "-Wl,-rpath," + PKG], check=True)
    r = subprocess.run([exe, refshim.REF_PATH, refshim.REF0_PATH], capture_output=True, text=True)
    last = (r.stdout.strip().splitlines() or ["<no output>"])[-1]
    assert r.returncode == 0 and last.endswith(" 0 differ"), f"{last}\n{r.stderr[-2000:]}"
    assert int(last.split()[1]) > 600, last


def _src(fmt, w, h, seed):
    if fmt == BC4:
        return synth.random_blocks((w * h + 15) // 16, 1, seed)[: w * h]
    if fmt == BC5:
        return synth.random_blocks((w * h * 2 + 15) // 16 + 1, 1, seed)[: w * h * 2]
    if fmt in (BC6U, BC6S):
        rng = np.random.default_rng(seed)
        img = (rng.random((h, w, 4)) * 10 ** rng.uniform(-2, 3, (h, w, 1))).astype(np.float32)
        if fmt == BC6S:
            img[::2, ::3, :3] *= -1
        return img
    # smooth colour + mixed alpha: a translucent, an opaque and a ramp region
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([(xx * 7 + yy * 3) & 255, (xx * 2 + yy * 9) & 255, (xx * 5 + yy) & 255, np.where(xx < w // 2, 255, (yy * 13) & 255)], -1)
    return np.ascontiguousarray(img.astype(np.uint8))


@pytest.mark.parametrize("fmt", [BC1, BC2, BC3, BC4, BC5, BC6U, BC7])
@pytest.mark.parametrize("w,h", [(22, 8), (37, 19), (5, 4), (131, 9), (4, 7)])
def test_width_not_multiple_of_four(fmt, w, h):
    """numBlocksX = Width / 4 truncates but rowStride stays Width * bpp (ImageEncoders.cpp:224-226): the last Width % 4 texels
    of each row and the last Height % 4 rows are skipped, nothing is sheared"""
    src = _src(fmt, w, h, 1000 + w)
    ref = refshim.ref0() if fmt in (BC6U, BC6S) else None
    want = refshim.compress(fmt, src, w, h, lib=ref)
    got = hkbcn.compress(fmt, src, w, h)
    assert np.array_equal(got, want)
    # the device-pointer form takes the same ragged surface
    import torch
    d_src = torch.from_numpy(np.ascontiguousarray(src).view(np.uint8).ravel().copy()).cuda()
    d_dst = torch.zeros(max(got.size, 16), dtype=torch.uint8, device="cuda")
    hkbcn.check(hkbcn.lib().hkbcn_compress_dev(0, fmt, d_src.data_ptr(), d_dst.data_ptr(), w, h, 0, C.c_void_p(0)))
    torch.cuda.synchronize()
    assert np.array_equal(d_dst.cpu().numpy()[: got.size], want)


def test_batch_takes_ragged_widths():
    jobs = [(fmt, _src(fmt, w, h, 7 + w), w, h, 0) for fmt, w, h in ((BC1, 22, 8), (BC7, 37, 19), (BC4, 131, 9), (BC5, 64, 64), (BC3, 18, 18))]
    outs = hkbcn.compress_batch(jobs)
    for (fmt, src, w, h, _), got in zip(jobs, outs):
        assert np.array_equal(got, refshim.compress(fmt, src, w, h))


def test_dev_entry_points_check_alignment_and_chain_dims():
    import torch
    L = hkbcn.lib()
    hkbcn.init()
    buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
    base = buf.data_ptr()
    # a block stream at an odd offset (a .texture payload uploaded with its 30-byte header): refused, not a sticky fault
    assert L.hkbcn_decompress_dev(0, BC7, base + 30, base + 32768, 16, 16, 64, 0, C.c_void_p(0)) == -2
    assert L.hkbcn_decompress_dev(0, BC1, base + 4, base + 32768, 16, 16, 64, 0, C.c_void_p(0)) == -2
    assert b"aligned" in L.hkbcn_last_error()
    # encoders store whole blocks: the destination must be block-aligned
    assert L.hkbcn_compress_dev(0, BC7, base, base + 32768 + 8, 16, 16, 0, C.c_void_p(0)) == -2
    assert L.hkbcn_compress_dev(0, BC1, base, base + 32768 + 4, 16, 16, 0, C.c_void_p(0)) == -2
    # mipmapped compressed storage: powers of two only, at most CalcNumMips levels (Image.cpp:415-431)
    assert L.hkbcn_compress_chain_dev(0, BC1, base, base + 32768, 12, 12, 2, 0, C.c_void_p(0)) == -2
    assert b"power-of-two" in L.hkbcn_last_error()
    assert L.hkbcn_compress_chain_dev(0, BC1, base, base + 32768, 16, 16, 4, 0, C.c_void_p(0)) == -2
    assert L.hkbcn_compress_chain_dev(0, BC1, base, base + 32768, 12, 12, 1, 0, C.c_void_p(0)) == 0
    torch.cuda.synchronize()
    # the context is still healthy
    img = synth.albedo_rgba8(16, 16, 3)
    assert np.array_equal(hkbcn.compress(BC1, img, 16, 16), refshim.compress(BC1, img, 16, 16))


@pytest.mark.parametrize("fmt,tb", [(BC1, 4), (BC3, 4), (BC4, 1), (BC5, 2), (BC7, 4), (BC6U, 6)])
def test_decompress_dev_unaligned_destination_and_odd_stride(fmt, tb):
    """hkbcn_decompress_dev called directly with a destination at an odd address and a row stride that is not a multiple of
    16: the non-vector store kernels (the host-pointer entry rounds its device stride up and never reaches them)"""
    import torch
    w, h = 40, 24
    nb = (w // 4) * (h // 4)
    bb = hkbcn.block_bytes(fmt)
    blocks = synth.random_blocks(nb, 1, 5 + fmt)[: nb * bb] if bb == 8 else synth.random_blocks(nb, 1, 5 + fmt)
    if fmt == BC7:
        from test_gpu_parity import force_modes_bc7
        blocks = force_modes_bc7(blocks.copy())
    if fmt == BC6U:
        from test_gpu_parity import force_modes_bc6h
        blocks = force_modes_bc6h(blocks.copy())
    align = 2 if fmt == BC6U else 1  # halfs must stay 2-byte aligned
    stride = w * tb + 6 * align + (0 if fmt == BC6U else 1)
    off = 2 if fmt == BC6U else 3
    d_src = torch.from_numpy(blocks.copy()).cuda()
    d_dst = torch.full((stride * h + 64,), 0xAB, dtype=torch.uint8, device="cuda")
    hkbcn.check(hkbcn.lib().hkbcn_decompress_dev(0, fmt, d_src.data_ptr(), d_dst.data_ptr() + off, w, h, stride, 0, C.c_void_p(0)))
    torch.cuda.synchronize()
    got = d_dst.cpu().numpy()
    want = refshim.decompress(fmt, blocks, w, h, row_stride=stride)
    rows = got[off: off + stride * h].reshape(h, stride)
    assert np.array_equal(rows[:, : w * tb], want.reshape(h, stride)[:, : w * tb])
    assert (rows[:-1, w * tb:] == 0xAB).all() and (got[:off] == 0xAB).all(), "bytes between rows were overwritten"


def test_golden_encodes():
    """tests/golden/bcn_golden.npz (tools/gen_golden.py: the reference's own output on seeded inputs, committed): the
    BC7 and BC6H encoders reproduce the stored blocks byte for byte (BC6H: the zero-init build's)"""
    g = np.load(GOLD)
    assert np.array_equal(hkbcn.compress(BC7, g["bc7_src"], 64, 64), g["bc7_ref_enc"])
    got = hkbcn.compress(BC6U, g["bc6h_src"], 32, 32)
    assert np.array_equal(got, g["bc6h_ref0_enc"])
    # as-shipped build (stack-dependent, bc6h_enc.h:3134-3150): same decoded quality within the north_star band
    from test_gpu_bc6h import quality
    pg, _ = quality(got, g["bc6h_src"], 32, 32, BC6U)
    pr, _ = quality(g["bc6h_ref_enc"], g["bc6h_src"], 32, 32, BC6U)
    assert pg >= pr - 0.05
    for key, fmt, src in (("bc1", BC1, "rgba"), ("bc2", BC2, "rgba"), ("bc3", BC3, "rgba"), ("bc4", BC4, "r8"), ("bc5", BC5, "rg8")):
        s = g[src]
        bpp = {"rgba": 4, "r8": 1, "rg8": 2}[src]
        n = s.size // (16 * bpp)
        assert np.array_equal(hkbcn.encode_blocks(fmt, s, n), g[key + "_enc"]), key


def nonfinite_hdr_blocks():
    """NaN / +-Inf / -0 texels (SURVEY.md 8c corpus: "document whatever the oracle emits")"""
    rng = np.random.default_rng(5)
    out = []
    specials = [np.nan, np.inf, -np.inf, -0.0, np.float32(3.4e38), -np.float32(3.4e38)]
    for s in specials:
        for k in (1, 4, 16):  # one texel, a few, the whole block
            px = (rng.random((16, 4)) * 8).astype(np.float32)
            px[:, 3] = 1.0
            idx = rng.permutation(16)[:k]
            px[idx, rng.integers(0, 3)] = s
            out.append(px.ravel())
        px = np.full((16, 4), s, np.float32)
        out.append(px.ravel())
    return np.concatenate(out)


@pytest.mark.parametrize("fmt", [BC6U, BC6S])
def test_bc6h_nonfinite_inputs_match_the_defined_reference(fmt):
    """half_from_float (Hork/Core/half.c:802-864) maps NaN to a NaN half and +-Inf to +-Inf; F16ToINT / the quantiser then
    treat the bit patterns as numbers (bc6h_enc.h:1015-1040).  Whatever that yields, the zero-init reference build yields it
    deterministically and the kernel must emit the same 16 bytes."""
    src = nonfinite_hdr_blocks()
    n = src.size // 64
    got = hkbcn.encode_blocks(fmt, src, n)
    want = refshim.compress(fmt, src, 4, 4 * n, lib=refshim.ref0())
    same = (got.reshape(n, 16) == want.reshape(n, 16)).all(axis=1)
    assert same.all(), f"blocks {np.flatnonzero(~same)} differ on non-finite input"
