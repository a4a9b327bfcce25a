"""The two sRGB tables of stb_image_resize: the decode table is recomputed by the library (float arithmetic, six
decimals) and must equal the header's 256 literals; the encode segments come from the engine tree through
tools/gen_stbir_tables.py, which writes a stub without it."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import hkbcn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "tools", "gen_stbir_tables.py")
STB = "/root/reference/ThirdParty/stb"
HAVE = os.path.exists(os.path.join(STB, "stb_image_resize.h"))


def test_stub_without_the_engine_tree(tmp_path):
    out = tmp_path / "s.inc"
    subprocess.run([sys.executable, GEN, str(tmp_path / "nowhere"), str(out)], check=True)
    assert "#define HK_HAVE_STBIR_SRGB 0" in out.read_text()


def test_decode_table_is_monotone_and_pinned_at_the_ends():
    L = hkbcn.lib()
    t = np.array([L.hkbcn_srgb_decode_entry(v) for v in range(256)], np.float32)
    assert t[0] == 0.0 and t[255] == 1.0 and (np.diff(t) > 0).all()
    assert L.hkbcn_srgb_decode_entry(-1) == 0.0 and L.hkbcn_srgb_decode_entry(256) == 0.0


@pytest.mark.skipif(not HAVE, reason="reference tree not present")
def test_decode_table_equals_the_headers_literals():
    src = open(os.path.join(STB, "stb_image_resize.h")).read()
    body = src[src.index("float stbir__srgb_uchar_to_linear_float[256] = {"):]
    body = body[: body.index("};")]
    want = np.array([np.float32(v) for v in re.findall(r"(\d+\.\d+)f", body)], np.float32)
    assert want.size == 256
    L = hkbcn.lib()
    got = np.array([L.hkbcn_srgb_decode_entry(v) for v in range(256)], np.float32)
    assert np.array_equal(got, want), np.flatnonzero(got != want)


@pytest.mark.skipif(not HAVE, reason="reference tree not present")
def test_encode_segments_from_the_engine_tree(tmp_path):
    out = tmp_path / "s.inc"
    subprocess.run([sys.executable, GEN, STB, str(out)], check=True)
    text = out.read_text()
    assert "#define HK_HAVE_STBIR_SRGB 1" in text
    vals = [int(v, 16) for v in re.findall(r"0x[0-9a-f]{8}", text)]
    assert len(vals) == 104 and vals[0] == 0x0073000D and vals[-1] == 0x7C330723
    assert hkbcn.lib().hkbcn_build_has(hkbcn.BUILD_STBIR_SRGB) == 1 and hkbcn.lib().hkbcn_build_has(hkbcn.BUILD_BC1_ORDERINGS) == 1
    assert hkbcn.lib().hkbcn_build_has(77) == 0
