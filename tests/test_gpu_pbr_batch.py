"""hkbcn_import_pbr_batch: PBRImporter's ORMX texture (Tools/PBRImporter/Main.cpp:284-735) as a library job -- CreateORMX gather,
the cascade of the RGBA8 storage with NO_ALPHA, the roughness channel of levels >= 1 updated from the triangle-filtered
normals, the ORMX_BC1 / _BC3 / _BC7 presets -- against the steps composed by hand from the entry points that each have their
own oracle test (generate_mips: test_gpu_mips vs stbir; bake_roughness: test_gpu_normals vs the reference's object code;
compress_chain: the reference's encoders), and level 0 against the reference's CompressBC* on the gathered image directly."""
import ctypes as C

import numpy as np
import pytest

import hkbcn
import refshim
import synth

pytestmark = pytest.mark.gpu
BC1, BC3, BC7 = 54, 58, 66


def make_job(fmt, n, maps, channels, normals, bake, mips, cfg, dst=None, path=None):
    j = hkbcn.PbrJob()
    j.fmt, j.width, j.height = fmt, n, n
    for k in range(4):
        j.channel[k] = channels[k]
        j.maps[k] = maps[k].ctypes.data if maps[k] is not None else None
    j.normals = normals.ctypes.data if normals is not None else None
    j.roughness_bake, j.generate_mips, j.mip = bake, mips, cfg
    j.out_path = path.encode() if path else None
    j.dst = dst.ctypes.data if dst is not None else None
    return j


def by_hand(fmt, n, maps, channels, normals, bake, mips, cfg):
    ormx = np.empty((n, n, 4), np.uint8)
    ormx[...] = (255, 255, 0, 255)
    for k in range(4):
        if maps[k] is not None:
            ormx[:, :, k] = maps[k].reshape(n, n, 4)[:, :, channels[k]]
    nm = hkbcn.lib().hkbcn_calc_num_mips(n, n) if mips else 1
    chain = ormx.ravel().copy()
    if nm > 1:
        chain = hkbcn.generate_mips(BC7, ormx, n, n, nm, hkbcn.MipConfig(cfg.edge_mode, cfg.filter, 1, 0)).copy()
        if bake and normals is not None:
            chain[1::4] = hkbcn.bake_roughness(normals, chain[1::4].copy(), n, n, nm, bake, 2, cfg.edge_mode)
    return ormx, hkbcn.compress_chain(fmt, chain, n, n, nm), nm


def test_jobs_equal_the_steps_composed_by_hand(tmp_path):
    hkbcn.init()
    n = 128
    rgba = [np.ascontiguousarray(synth.albedo_rgba8(n, n, 0xB2000100 + k, "mask")) for k in range(4)]
    nrm8 = np.concatenate([synth.normals_rg8(n, n, 13).reshape(n, n, 2), np.full((n, n, 1), 255, np.uint8)], axis=2)
    normals = np.ascontiguousarray(hkbcn.normalize_vectors((nrm8.astype(np.float32) / np.float32(255.0)).reshape(-1, 3)))
    cases = [
        # fmt, maps present, channels, normals?, bake, mips, (edge, filter)
        (BC7, (1, 1, 1, 1), (0, 1, 2, 3), True, 1, 1, (3, 5)),
        (BC1, (1, 1, 0, 0), (1, 1, 0, 0), True, 2, 1, (1, 2)),   # the reference's default channels, Toksvig, clamp / triangle
        (BC3, (0, 1, 0, 1), (0, 2, 0, 0), False, 1, 1, (3, 5)),  # bake asked for without normals: nothing to anti-alias
        (BC7, (1, 0, 1, 0), (3, 0, 0, 0), True, 1, 0, (0, 0)),   # no mips: a single level, no bake
        (BC1, (1, 1, 1, 0), (2, 2, 2, 0), True, 0, 1, (2, 4)),   # normals given, no bake; reflect / Catmull-Rom
    ]
    jobs = (hkbcn.PbrJob * len(cases))()
    keep, want = [], []
    for i, (fmt, present, channels, with_n, bake, mips, (edge, filt)) in enumerate(cases):
        maps = [rgba[k] if present[k] else None for k in range(4)]
        cfg = hkbcn.MipConfig(edge, filt, 0, 0)
        nm = hkbcn.lib().hkbcn_calc_num_mips(n, n) if mips else 1
        dst = np.zeros(hkbcn.lib().hkbcn_chain_bytes(fmt, n, n, 1, nm), np.uint8)
        path = str(tmp_path / f"orm{i}.texture") if i != 2 else None
        jobs[i] = make_job(fmt, n, maps, channels, normals if with_n else None, bake, mips, cfg, dst if i != 1 else None, path)
        keep.append((maps, dst, path))
        cfg_eff = hkbcn.MipConfig(edge or 3, filt or 5, 0, 0)
        want.append(by_hand(fmt, n, maps, channels, normals if with_n else None, bake, mips, cfg_eff))
    hkbcn.check(hkbcn.lib().hkbcn_import_pbr_batch(jobs, len(cases)))
    for i, (fmt, *_rest) in enumerate(cases):
        maps, dst, path = keep[i]
        ormx, blocks, nm = want[i]
        assert jobs[i].status == 0
        if i != 1:
            assert np.array_equal(dst, blocks), f"job {i}: chain in memory differs from the composed steps"
        if path:
            d, payload = hkbcn.read_texture(path)
            assert (d.type, d.width, d.height, d.depth, d.num_mips, d.fmt, d.storage_flags) == (2, n, n, 1, nm, fmt, 1)
            assert np.array_equal(payload, blocks), f"job {i}: file payload differs"
        level0 = (n // 4) * (n // 4) * hkbcn.block_bytes(fmt)
        assert np.array_equal(blocks[:level0], refshim.compress(fmt, ormx, n, n)), f"job {i}: level 0 differs from the reference's encoder on the gathered image"
    # the bake really ran where it should have
    plain = by_hand(BC7, n, [rgba[k] for k in range(4)], (0, 1, 2, 3), None, 0, 1, hkbcn.MipConfig(3, 5, 0, 0))[1]
    assert not np.array_equal(plain, want[0][1]), "the roughness update changed nothing: test content too flat"


def test_bad_jobs_are_refused_before_anything_runs():
    hkbcn.init()
    L = hkbcn.lib()
    n = 64
    img = np.zeros((n, n, 4), np.uint8)
    dst = np.zeros(1 << 16, np.uint8)
    cfg = hkbcn.MipConfig(0, 0, 0, 0)

    def rc(job):
        arr = (hkbcn.PbrJob * 1)(job)
        return L.hkbcn_import_pbr_batch(arr, 1)
    assert rc(make_job(60, n, [img, None, None, None], (0, 0, 0, 0), None, 0, 0, cfg, dst)) == -1          # BC4 is not an ORMX preset
    assert rc(make_job(BC7, n, [None] * 4, (0, 0, 0, 0), None, 0, 0, cfg, dst)) == -2                       # no map at all
    assert rc(make_job(BC7, n, [img, None, None, None], (0, 4, 0, 0), None, 0, 0, cfg, dst)) == -2          # channel 4
    assert rc(make_job(BC7, n, [img, None, None, None], (0, 0, 0, 0), None, 7, 0, cfg, dst)) == -2          # unknown bake
    assert rc(make_job(BC7, n, [img, None, None, None], (0, 0, 0, 0), None, 0, 0, cfg, None)) == -2         # nowhere to put the result
    j = make_job(BC7, 48, [np.zeros((48, 48, 4), np.uint8), None, None, None], (0, 0, 0, 0), None, 0, 1, cfg, dst)
    assert rc(j) == -2 and b"power-of-two" in L.hkbcn_last_error()                                          # mips need powers of two
    assert (dst == 0).all()
