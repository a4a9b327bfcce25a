"""Row f3, CPU side: the C restatement of the env-probe shaders (oracle/oracle_envprobe.c) against properties with known
answers -- it is the only checker the CUDA kernels have (the reference runs this math through OpenGL only), so it is
pinned here by what the shader must produce analytically."""
import numpy as np

import oracle_c
from cubesynth import sky_cube


def test_constant_environment_is_a_fixed_point():
    cube = np.zeros((6, 8, 8, 4), np.float32)
    cube[..., :3] = [0.25, 1.5, 3.0]
    cube[..., 3] = 1.0
    for lv in oracle_c.envprobe_prefilter(cube, 8, 3):
        np.testing.assert_allclose(lv[..., :3], np.broadcast_to(np.float32([0.25, 1.5, 3.0]), lv[..., :3].shape), rtol=2e-6)
        assert np.all(lv[..., 3] == 1.0)
    # irradiance of constant radiance L is pi * mean(cos * sin) * L = L up to the 0.025 rad Riemann sum
    irr = oracle_c.irradiance(cube, 8, 4)
    np.testing.assert_allclose(irr[..., :3], np.broadcast_to(np.float32([0.25, 1.5, 3.0]), irr[..., :3].shape), rtol=0.02)


def test_level0_is_the_source_when_sizes_match():
    """roughness 0: H = N, L = R, every sample fetches the texel under R; with a 2^max_lod source R is a texel centre"""
    cube = sky_cube(16, 3)
    lv = oracle_c.envprobe_prefilter(cube, 16, 4)
    np.testing.assert_allclose(lv[0][..., :3], cube[..., :3], rtol=1e-4, atol=1e-3)  # L = R up to float normalisation; the sun is steep
    assert [l.shape for l in lv] == [(6, 16, 16, 4), (6, 8, 8, 4), (6, 4, 4, 4), (6, 2, 2, 4), (6, 1, 1, 4)]


def test_rougher_levels_blur_and_conserve_energy_roughly():
    cube = sky_cube(32, 1)
    lv = oracle_c.envprobe_prefilter(cube, 32, 3)
    peaks = [float(l[..., :3].max()) for l in lv]
    assert peaks[0] > peaks[1] > peaks[2] > peaks[3]          # the sun spreads out
    lo = float(cube[..., :3].min())
    assert all(float(l[..., :3].min()) >= lo - 1e-4 for l in lv)  # convex combinations of source texels
    irr = oracle_c.irradiance(cube, 32, 4)
    # the +Y (sky) face is brighter than the -Y (ground) face
    assert irr[2, ..., :3].mean() > irr[3, ..., :3].mean()
