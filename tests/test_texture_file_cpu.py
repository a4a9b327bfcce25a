"""Row f2: the .texture container.  The oracle here is a restatement of the reference's writer in Python
(AssetUtils::CreateTexture, Hork/Resources/Resource_Texture.cpp:104-109; MakeResourceMagic, ResourceBase.h:72-75;
ImageStorage::Write / Read, Hork/Image/Image.cpp:833-864; little-endian WriteUInt32, Core/BinaryStream.h:408-412);
the library's writer must produce the same bytes and its reader must take them back.  Host-only entry points: no GPU."""
import os
import struct
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "hork-source_b200", "python"))
import hkbcn  # noqa: E402

RESOURCE_TEXTURE, VERSION = 4, 1  # ResourceBase.h:40-45 (UNDEFINED, MESH, ANIMATION, NODE_MOTION, TEXTURE), Resource_Texture.h:45-46


def oracle_texture_file(type_, width, height, depth, mips, fmt, flags, payload):
    magic = ord("H") | (ord("k") << 8) | (RESOURCE_TEXTURE << 16) | (VERSION << 24)
    out = struct.pack("<I", magic)                        # stream.WriteUInt32(MakeResourceMagic(Type, Version))
    out += struct.pack("<B", type_)                       # WriteUInt8(m_Desc.Type)
    out += struct.pack("<III", width, height, depth)      # WriteUInt32 x 3
    out += struct.pack("<I", mips)
    out += struct.pack("<B", fmt)
    out += struct.pack("<I", flags)
    out += struct.pack("<I", len(payload))                # WriteUInt32(m_Data.Size())
    return out + bytes(payload)


CASES = [(2, 2048, 2048, 1, 12, 54, 0), (2, 4096, 4096, 1, 1, 66, 1), (5, 1024, 1024, 6, 9, 64, 0), (2, 4, 4, 1, 1, 60, 3),
         (2, 65536, 4, 1, 1, 67, 2)]


@pytest.mark.parametrize("case", CASES)
def test_header_bytes_match_the_reference_writer(case):
    n = 12345
    want = oracle_texture_file(*case, b"\0" * n)[:30]
    got = hkbcn.texture_header(*case, n)
    assert len(got) == 30 and got == want
    assert got[:4] == b"Hk\x04\x01"


def test_write_read_round_trip(tmp_path):
    rng = np.random.default_rng(5)
    for i, case in enumerate(CASES):
        payload = rng.integers(0, 256, 1000 + 17 * i, dtype=np.uint8)
        path = str(tmp_path / f"t{i}.texture")
        hkbcn.write_texture(path, *case, payload)
        raw = open(path, "rb").read()
        assert raw == oracle_texture_file(*case, payload.tobytes())
        d, back = hkbcn.read_texture(path)
        assert (d.type, d.width, d.height, d.depth, d.num_mips, d.fmt, d.storage_flags) == case
        assert np.array_equal(back, payload)


def test_reader_rejects_other_files(tmp_path):
    p = tmp_path / "x.texture"
    p.write_bytes(b"Hk\x01\x01" + b"\0" * 40)  # a mesh resource's magic
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.read_texture(str(p))
    p.write_bytes(b"Hk\x04\x01\x02")           # truncated header
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.read_texture(str(p))
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.read_texture(str(tmp_path / "missing.texture"))
