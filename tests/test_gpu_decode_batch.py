"""GPU parity of the storage / batch decode entry points (hkbcn_decompress_slices, _chain, _batch): every (level, slice)
subresource and every job bit-exact against the reference decoders (bcdec / bc7decomp through oracle/_ref), all seven
formats including BC2 and signed BC6H, random 128-bit blocks with every BC7 / BC6H mode forced in turn."""
import numpy as np
import pytest

import hkbcn
import refshim
import synth
from test_gpu_parity import force_modes_bc6h, force_modes_bc7

pytestmark = pytest.mark.gpu
FORMATS = [(54, 8, 4, 0), (56, 16, 4, 0), (58, 16, 4, 0), (60, 8, 1, 0), (62, 16, 2, 0), (66, 16, 4, 0), (64, 16, 6, 0), (65, 16, 6, 0),
           (64, 16, 12, hkbcn.FLAG_DST_F32), (65, 16, 12, hkbcn.FLAG_DST_F32)]


def blocks_for(fmt, n, seed):
    bb = hkbcn.block_bytes(fmt)
    b = synth.random_blocks(n, 1, seed)[: n * bb].copy()
    if fmt == 66:
        b = force_modes_bc7(b)
    if fmt in (64, 65):
        b = force_modes_bc6h(b)
    return b


@pytest.mark.parametrize("fmt,bb,tb,flags", FORMATS)
def test_decompress_slices_matches_reference_per_subresource(fmt, bb, tb, flags):
    w, h, slices = 64, 32, 6
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    total = hkbcn.lib().hkbcn_chain_bytes(fmt, w, h, slices, mips) // bb
    blocks = blocks_for(fmt, total, 100 + fmt)
    got = hkbcn.decompress_slices(fmt, blocks, w, h, slices, mips, flags)
    assert got.size == hkbcn.lib().hkbcn_decoded_chain_bytes(fmt, w, h, slices, mips, flags)
    so = do = 0
    for lw, lh in synth.mip_dims(w, h, mips):
        for s in range(slices):
            nb = (lw // 4) * (lh // 4) * bb
            want = refshim.decompress(fmt, blocks[so:so + nb], lw, lh, f32out=bool(flags))
            assert np.array_equal(got[do:do + want.size], want), f"level {lw}x{lh} slice {s}"
            so += nb
            do += want.size
    assert so == blocks.size and do == got.size


def test_decompress_batch_mixed_jobs():
    jobs = []
    for i, (fmt, bb, tb, flags) in enumerate(FORMATS):
        w, h = 16 * (1 + i % 5), 8 * (1 + i % 3)
        jobs.append((fmt, blocks_for(fmt, (w // 4) * (h // 4), 7 + i), w, h, flags))
    outs = hkbcn.decompress_batch(jobs)
    for (fmt, blocks, w, h, flags), got in zip(jobs, outs):
        assert np.array_equal(got, refshim.decompress(fmt, blocks, w, h, f32out=bool(flags))), f"fmt {fmt} {w}x{h}"


def test_batch_of_equal_textures_is_the_slices_form():
    """16 BC7 512^2 textures as one storage with 16 slices: one launch, byte-identical to 16 single decodes"""
    n, cnt = 512, 16
    nb = (n // 4) ** 2
    blocks = blocks_for(66, nb * cnt, 5)
    l0 = hkbcn.lib().hkbcn_launch_count()
    got = hkbcn.decompress_slices(66, blocks, n, n, cnt, 1)
    assert hkbcn.lib().hkbcn_launch_count() == l0 + 1
    for i in range(cnt):
        one = hkbcn.decompress(66, blocks[i * nb * 16:(i + 1) * nb * 16], n, n)
        assert np.array_equal(got[i * n * n * 4:(i + 1) * n * n * 4], one)
