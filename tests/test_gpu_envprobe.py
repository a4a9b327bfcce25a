"""Row f3 parity: the CUDA env-probe kernels against the CPU restatement of the reference's shaders
(oracle/oracle_envprobe.c; envprobegen.frag:44-111, irradiancegen.frag:46-76).  Tolerance: relative 1e-3 (SURVEY.md 8f f3:
GL leaves filtering precision to the implementation, so there is no bit contract; here the only differences are the order
of the float sums and sincosf vs sinf / cosf)."""
import numpy as np
import pytest

import hkbcn
import oracle_c
from cubesynth import sky_cube

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-3, 1e-4


@pytest.mark.parametrize("src,max_lod", [(32, 4), (16, 4), (64, 3)])
def test_prefilter_matches_the_shader_restatement(src, max_lod):
    cube = sky_cube(src, src + max_lod)
    got = hkbcn.envprobe_prefilter(cube, src, max_lod)
    want = oracle_c.envprobe_prefilter(cube, src, max_lod)
    assert len(got) == max_lod + 1
    for lod, (g, w) in enumerate(zip(got, want)):
        assert g.shape == w.shape == (6, (1 << max_lod) >> lod, (1 << max_lod) >> lod, 4)
        np.testing.assert_allclose(g[..., :3], w[..., :3], rtol=RTOL, atol=ATOL, err_msg=f"level {lod}")
        assert np.all(g[..., 3] == 1.0)


def test_prefilter_level0_reproduces_the_source():
    cube = sky_cube(32, 9)
    got = hkbcn.envprobe_prefilter(cube, 32, 5)
    np.testing.assert_allclose(got[0][..., :3], cube[..., :3], rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("src,out", [(32, 8), (16, 4)])
def test_irradiance_matches_the_shader_restatement(src, out):
    cube = sky_cube(src, 5)
    got = hkbcn.irradiance(cube, src, out)
    want = oracle_c.irradiance(cube, src, out)
    np.testing.assert_allclose(got[..., :3], want[..., :3], rtol=RTOL, atol=ATOL)
    assert np.all(got[..., 3] == 1.0)


def test_prefiltered_levels_feed_the_bc6h_encoder():
    """config 4's chain: every level at least one block wide goes through CompressBC6h face by face"""
    cube = sky_cube(32, 2)
    levels = hkbcn.envprobe_prefilter(cube, 32, 4)  # 16, 8, 4, 2, 1
    for lv in levels:
        s = lv.shape[1]
        if s < 4:
            continue
        for f in range(6):
            blocks = hkbcn.compress(64, lv[f], s, s)
            dec = hkbcn.decompress(64, blocks, s, s, flags=1 << 8).view(np.float32).reshape(s, s, 3)
            err = np.abs(dec - lv[f][..., :3]) / (np.abs(lv[f][..., :3]) + 0.05)
            assert np.median(err) < 0.08 and err.max() < 1.0  # BC6H quantisation on a noisy 4-16 px face; a sanity bound, not a quality claim


def test_bad_arguments_are_refused():
    cube = sky_cube(8, 0)
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.envprobe_prefilter(cube, 8, 0)   # roughness would be 0 / 0
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.irradiance(cube, 0, 4)
