"""CPU test of the mip cascade's host logic: the resampling taps hkbcn_mip_taps builds (stbir's downsampling contributors,
normalisation and edge mapping restated, csrc/mip_resize.cu) applied with numpy reproduce the reference's own stbir_resize
on float images exactly, for every filter and edge mode, 2:1 and odd ratios, and the 4 -> 4 step of clamped chain levels."""
import numpy as np
import pytest

import hkbcn
import refshim


@pytest.mark.skipif(not refshim.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("iw,ow", [(64, 32), (8, 4), (4, 4), (100, 50), (37, 18)])
def test_taps_reproduce_stbir(iw, ow):
    rng = np.random.default_rng(iw * 131 + ow)
    for filt in (1, 2, 3, 4, 5):
        for edge in (1, 2, 3, 4):
            img = rng.random((iw, iw, 4)).astype(np.float32)
            ix, wx = hkbcn.mip_taps(iw, ow, filt, edge)
            h = np.zeros((iw, ow, 4), np.float32)
            for k in range(ix.shape[1]):
                h += img[:, ix[:, k], :] * wx[:, k][None, :, None]
            o = np.zeros((ow, ow, 4), np.float32)
            for k in range(ix.shape[1]):
                o += h[ix[:, k], :, :] * wx[:, k][:, None, None]
            ref = refshim.stbir_resize(img, ow, ow, alpha_channel=-1, flags=0, edge=edge, filt=filt, space=0)
            assert np.array_equal(o, ref), f"filter {filt} edge {edge}: max diff {np.abs(o - ref).max()}"


def test_taps_shape_and_normalisation():
    ix, wx = hkbcn.mip_taps(4096, 2048)  # Mitchell, wrap: eight taps everywhere, the same kernel for every output
    assert ix.shape == (2048, 8)
    assert np.allclose(wx.sum(axis=1), 1.0, atol=1e-6)
    assert np.array_equal(wx[0], wx[1000])
    assert list(ix[0]) == [4093, 4094, 4095, 0, 1, 2, 3, 4]
    ix, wx = hkbcn.mip_taps(4, 8)  # upsampling: stbir's other branch (tests/test_gpu_resample.py)
    assert ix.shape[0] == 8 and np.allclose(wx.sum(axis=1), 1.0, atol=1e-6)
