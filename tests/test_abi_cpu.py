"""CPU tests of the C-ABI boundary: the library loads, exports every symbol include/hkbcn.h
declares, its host-only helpers agree with the reference's ImageStorage arithmetic, and without a
GPU every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import hkbcn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "hkbcn.h")).read()
    return sorted(set(re.findall(r"HKBCN_API\s+[\w\s\*]+?\b(hkbcn_\w+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    L = hkbcn.lib()
    syms = declared_symbols()
    assert len(syms) >= 20
    assert set(syms) == set(hkbcn.EXPORTS)
    for s in syms:
        assert hasattr(L, s), f"libhkbcn.so does not export {s}"


def test_format_table_matches_reference_enum():
    # Hk::TEXTURE_FORMAT values, Image.h:122-148; BytesPerBlock from TexFormat[] (Image.cpp:49-119)
    L = hkbcn.lib()
    assert [L.hkbcn_block_bytes(f) for f in range(54, 68)] == [8, 8, 16, 16, 16, 16, 8, 8, 16, 16, 16, 16, 16, 16]
    assert L.hkbcn_block_bytes(53) < 0 and L.hkbcn_block_bytes(68) < 0
    assert [L.hkbcn_src_texel_bytes(f, 0) for f in (54, 56, 58, 60, 62, 64, 66)] == [4, 4, 4, 1, 2, 16, 4]
    assert L.hkbcn_src_texel_bytes(64, hkbcn.FLAG_SRC_F16) == 8
    assert L.hkbcn_dst_texel_bytes(64, 0) == 6 and L.hkbcn_dst_texel_bytes(65, hkbcn.FLAG_DST_F32) == 12


def test_calc_num_mips_and_chain_bytes():
    # CalcNumMips (Image.cpp:188-216): log2(max(w,h)/4)+1; sizes from SURVEY.md 8 conventions
    L = hkbcn.lib()
    assert L.hkbcn_calc_num_mips(2048, 2048) == 10
    assert L.hkbcn_calc_num_mips(4096, 4096) == 11
    assert L.hkbcn_calc_num_mips(1024, 1024) == 9
    assert L.hkbcn_calc_num_mips(4, 4) == 1 and L.hkbcn_calc_num_mips(64, 8) == 5
    assert L.hkbcn_chain_bytes(54, 2048, 2048, 1, 10) == 349525 * 8
    assert L.hkbcn_chain_bytes(66, 4096, 4096, 1, 11) == 1398101 * 16
    assert L.hkbcn_chain_bytes(64, 1024, 1024, 6, 1) == 393216 * 16


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="GPU present")
def test_no_cpu_fallback_without_gpu():
    src = np.zeros(64, np.uint8)
    with pytest.raises(hkbcn.HkbcnError) as e:
        hkbcn.compress(hkbcn.BC1_UNORM, src, 4, 4)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.decompress(hkbcn.BC1_UNORM, np.zeros(8, np.uint8), 4, 4)
