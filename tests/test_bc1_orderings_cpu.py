"""tools/gen_bc1_orderings.py, the build step that reads rgbcx's total-ordering tables out of the engine tree:
without the tree it writes a stub that switches levels 6..18 off; with it, one blob holding four tables of the right shapes whose histograms
sum to 16 and whose special entries sit where rgbcx.cpp:259-268 says they do."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "tools", "gen_bc1_orderings.py")
REF = "/root/reference/ThirdParty/bc7enc_rdo"


def test_stub_without_the_engine_tree(tmp_path):
    out = tmp_path / "o.inc"
    subprocess.run([sys.executable, GEN, str(tmp_path / "nowhere"), str(out)], check=True)
    text = out.read_text()
    assert "#define HK_HAVE_BC1_ORDERINGS 0" in text and not (tmp_path / "o.bin").exists()


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "rgbcx.cpp")), reason="reference tree not present")
def test_tables_from_the_engine_tree(tmp_path):
    out = tmp_path / "o.inc"
    subprocess.run([sys.executable, GEN, REF, str(out)], check=True)
    text = out.read_text()
    assert "#define HK_HAVE_BC1_ORDERINGS 1" in text

    import struct
    assert "HK_BC1_ORDERINGS_BYTES 257300" in text
    blob = (tmp_path / "o.bin").read_bytes()
    assert blob[:4] == b"HKO1" and len(blob) == 257300
    o4 = [list(blob[4 + 4 * i: 8 + 4 * i]) for i in range(969)]
    at = 4 + 969 * 4
    o3 = [list(blob[at + 3 * i: at + 3 * i + 3]) for i in range(153)]
    at += 153 * 3 + 1
    b4 = struct.unpack("<%dH" % (969 * 128), blob[at: at + 969 * 128 * 2])
    b3 = blob[at + 969 * 128 * 2:]
    assert len(b3) == 153 * 32
    assert all(sum(r) == 16 for r in o4) and all(sum(r) == 16 for r in o3)
    assert len({tuple(r) for r in o4}) == 969 and len({tuple(r) for r in o3}) == 153  # every histogram once
    # TOTAL_ORDER_4_0_16 .. _3_16 and TOTAL_ORDER_3_0_16 .. _2_16
    assert o4[15] == [16, 0, 0, 0] and o4[700] == [0, 16, 0, 0] and o4[753] == [0, 0, 16, 0] and o4[515] == [0, 0, 0, 16]
    # (the reference's constants for the two single-selector 3-colour histograms point at each other's entries -- hist3 says
    # m_hist[1] == 16 -> 15, the table's entry 15 is {0, 0, 16}; both only ever select the single-colour fallback, and the
    # kernel uses the same constants, rgbcx.cpp:266-268, :1634)
    assert o3[12] == [16, 0, 0] and o3[15] == [0, 0, 16] and o3[89] == [0, 16, 0]
    assert max(b4) < 969 and max(b3) < 153
    assert list(b4[:4]) == [202, 403, 51, 450]  # the head of rgbcx_table4.h
