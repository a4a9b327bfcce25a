"""CPU tests of the host side: the C++ shim exports the reference's TextureBlockCompression names, and the
BC7 kernel's per-block function -- compiled for the host by tools/bc7_host_check.cpp, test infrastructure only --
reproduces the reference's bc7enc byte for byte on the committed golden vectors and on fresh adversarial blocks."""
import os
import subprocess

import numpy as np
import pytest

import refshim
import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "hork-source_b200", "libhkbcn_hork.so")

# Hork/Image/ImageEncoders.h:48-85
REFERENCE_NAMES = ["Decode_BC1", "Decode_BC2", "Decode_BC3", "Decode_BC4", "Decode_BC5", "Decode_BC6h_f16", "Decode_BC6h_f32", "Decode_BC7",
                   "Encode_BC1", "Encode_BC2", "Encode_BC3", "Encode_BC4", "Encode_BC5", "Encode_BC6h_f16", "Encode_BC6h_f32", "Encode_BC7",
                   "CompressBC1", "CompressBC2", "CompressBC3", "CompressBC4", "CompressBC5", "CompressBC6h", "CompressBC7"]


def test_shim_exports_reference_names():
    assert os.path.exists(SHIM), "run `python __graft_entry__.py build`"
    out = subprocess.run(["nm", "-DC", "--defined-only", SHIM], capture_output=True, text=True, check=True).stdout
    for n in REFERENCE_NAMES:
        assert f"Hk::TextureBlockCompression::{n}(" in out, f"shim lacks Hk::TextureBlockCompression::{n}"
    # the load-bearing 4-argument signature (Image.cpp:1718: void (*)(void const*, void*, uint32_t, uint32_t))
    assert "Hk::TextureBlockCompression::CompressBC7(void const*, void*, unsigned int, unsigned int)" in out


@pytest.fixture(scope="module")
def host_check(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bc7") / "bc7_host_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-Wno-unused", "-o", exe,
                    os.path.join(ROOT, "tools", "bc7_host_check.cpp"), "-ldl"], check=True)
    return exe


def run_check(exe, blocks, tmp_path):
    f = tmp_path / "blocks.bin"
    np.ascontiguousarray(blocks, dtype=np.uint8).tofile(f)
    out = subprocess.run([exe, refshim.REF_PATH, str(f)], capture_output=True, text=True, check=True).stdout
    return out.strip().splitlines()[-1]


@pytest.mark.skipif(not refshim.have_ref(), reason="oracle/_ref not built")
def test_bc7_block_function_matches_reference_on_host(host_check, tmp_path):
    rng = np.random.default_rng(11)
    parts = []
    for seed, alpha in ((0xB2000004, "opaque"), (0xB2000005, "mask"), (0xB2000007, "noise")):
        img = synth.albedo_rgba8(128, 128, seed, alpha)
        parts.append(img.reshape(32, 4, 32, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64))
    adv = []
    for v in range(0, 256, 5):
        adv.append(np.tile(np.array([v, v, v, 255], np.uint8), 16))
        adv.append(np.tile(np.array([v, 255 - v, v // 2, v], np.uint8), 16))
    for _ in range(400):  # two-colour and ramp blocks, opaque and translucent
        c0, c1 = rng.integers(0, 256, 4), rng.integers(0, 256, 4)
        if rng.random() < 0.5:
            c0[3] = c1[3] = 255
        m = rng.random(16) < rng.random()
        adv.append(np.where(m[:, None], c0, c1).astype(np.uint8).ravel())
        t = (np.arange(16) / 15.0)[rng.permutation(16)][:, None]
        adv.append(np.clip(c0 * (1 - t) + c1 * t + rng.normal(0, 3, (16, 4)), 0, 255).astype(np.uint8).ravel())
    for _ in range(200):  # constant colour, varying alpha (mode 7 single-colour path, bc7enc.cpp:587-643)
        c = rng.integers(0, 256, 3)
        a = rng.integers(200, 256, 16)
        adv.append(np.concatenate([np.tile(c, (16, 1)), a[:, None]], 1).astype(np.uint8).ravel())
    parts.append(np.stack(adv))
    parts.append(synth.random_blocks(1500, 4, 99).reshape(-1, 64))
    blocks = np.concatenate(parts)
    last = run_check(host_check, blocks, tmp_path)
    assert last == f"{len(blocks)} blocks, 0 differ", last


@pytest.mark.skipif(not refshim.have_ref(), reason="oracle/_ref not built")
def test_lean_decoders_match_reference_on_host(tmp_path):
    """the register-level BC1-5 / BC7 / BC6H block decoders of the CUDA kernels (bcn_decode_lean.cuh, bc7_decode_lean.cuh,
    bc6h_decode_lean.cuh), compiled for the host with PRMT emulated, against bcdec / bc7decomp: per-channel exhaustive endpoint
    pairs, all alpha endpoint pairs, random blocks with every BC7 / BC6H mode forced in turn, signed and unsigned."""
    exe = str(tmp_path / "decode_host_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-DWITH_BC7", "-o", exe, os.path.join(ROOT, "tools", "decode_host_check.cpp"), "-ldl"], check=True)
    out = subprocess.run([exe, refshim.REF_PATH, "60000"], capture_output=True, text=True, check=True).stdout.strip().splitlines()[-1]
    assert out.endswith(", 0 differ") and int(out.split()[0]) > 500000, out


def test_plan_bands_tiles_every_level():
    """hkbcn_plan_bands (host only): the bands of n GPUs partition the block rows of every level, in order, without overlap"""
    import hkbcn
    for height, mips in ((4096, 11), (64, 5), (4, 1), (1000, 6)):
        for n in (1, 2, 3, 8):
            edges = [hkbcn.plan_bands(height, mips, n, d) for d in range(n)]
            for lvl in range(mips):
                nby = max(height >> lvl, 4) // 4
                assert edges[0][0][lvl] == 0 and edges[-1][1][lvl] == nby
                for d in range(1, n):
                    assert edges[d][0][lvl] == edges[d - 1][1][lvl]
                sizes = [e[1][lvl] - e[0][lvl] for e in edges]
                assert max(sizes) - min(sizes) <= 1
