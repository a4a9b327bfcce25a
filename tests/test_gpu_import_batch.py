"""Row f2: hkbcn_import_batch = what one TextureImporter run per texture does (Tools/TextureImporter/Main.cpp:58-253 ->
CreateImage, Hork/Image/Image.cpp:1700-1803 -> AssetUtils::CreateTexture), as one sharded batch.  The files must be the
bytes the reference's writer would produce around the chain (oracle restatement in test_texture_file_cpu.py), the
chain must be hkbcn_import_chain's, and where the whole pipeline is deterministic integer work (BC4 / BC5: linear uint8
cascade, bit-exact encoders) the file must equal the one built from the reference's own stbir + rgbcx chain."""
import ctypes as C

import numpy as np
import pytest

import hkbcn
import refshim
import synth
from test_gpu_mips import ref_cascade
from test_texture_file_cpu import oracle_texture_file

pytestmark = pytest.mark.gpu
BC1S, BC3, BC4, BC5, BC7 = 55, 58, 60, 62, 66


def _jobs(tmp_path):
    tri_wrap = hkbcn.MipConfig(3, 2, 0, 0)  # normal / roughness maps are mipped with the triangle filter (Image.cpp:2359, :2463)
    return [
        dict(fmt=BC4, level0=synth.roughness_r8(512, 512, 2).reshape(512, 512, 1), width=512, height=512, mips=True, cfg=tri_wrap,
             out_path=str(tmp_path / "rough.texture")),
        dict(fmt=BC5, level0=synth.normals_rg8(256, 128, 3).reshape(128, 256, 2), width=256, height=128, mips=True, cfg=tri_wrap,
             out_path=str(tmp_path / "normal.texture")),
        dict(fmt=BC1S, level0=synth.albedo_rgba8(256, 256, 0xB2000001, "opaque").reshape(256, 256, 4), width=256, height=256, mips=True,
             cfg=None, out_path=str(tmp_path / "albedo.texture")),
        dict(fmt=BC7, level0=synth.albedo_rgba8(128, 128, 0xB2000004, "opaque").reshape(128, 128, 4), width=128, height=128, mips=False,
             cfg=None, out_path=str(tmp_path / "albedo7.texture")),
        dict(fmt=BC3, level0=synth.albedo_rgba8(64, 64, 0xB2000005, "mask").reshape(64, 64, 4), width=64, height=64, mips=False,
             cfg=hkbcn.MipConfig(0, 0, 0, 1), out_path=str(tmp_path / "decal.texture")),
        dict(fmt=BC4, level0=synth.roughness_r8(64, 64, 9).reshape(64, 64, 1), width=64, height=64, mips=False, cfg=None, out_path=None),
    ]


def test_batch_writes_the_importers_files(tmp_path):
    jobs = _jobs(tmp_path)
    chains = hkbcn.import_batch(jobs)
    assert len(chains) == len(jobs)
    for j, chain in zip(jobs, chains):
        w, h, fmt = j["width"], j["height"], j["fmt"]
        mips = hkbcn.lib().hkbcn_calc_num_mips(w, h) if j["mips"] else 1
        # the chain is the single-texture entry point's, byte for byte
        assert np.array_equal(chain, hkbcn.import_chain(fmt, j["level0"], w, h, mips, j["cfg"]))
        if j["out_path"] is None:
            continue
        cfg = j["cfg"] or hkbcn.MipConfig()
        flags = (1 if cfg.no_alpha else 0) | (2 if cfg.alpha_premultiplied else 0)
        raw = open(j["out_path"], "rb").read()
        assert raw == oracle_texture_file(2, w, h, 1, mips, fmt, flags, chain.tobytes())
        d, payload = hkbcn.read_texture(j["out_path"])
        assert (d.type, d.width, d.height, d.depth, d.num_mips, d.fmt, d.storage_flags) == (2, w, h, 1, mips, fmt, flags)
        assert payload.nbytes == hkbcn.lib().hkbcn_chain_bytes(fmt, w, h, 1, mips)


@pytest.mark.parametrize("which", [0, 1])
def test_bc4_bc5_files_equal_the_reference_pipeline(tmp_path, which):
    """stbir cascade (triangle, wrap) + rgbcx encode_bc4_hq / bc5_hq of every level, wrapped by the reference writer"""
    j = _jobs(tmp_path)[which]
    w, h, fmt = j["width"], j["height"], j["fmt"]
    mips = hkbcn.lib().hkbcn_calc_num_mips(w, h)
    hkbcn.import_batch([j])
    levels = ref_cascade(j["level0"], mips, -1, False, edge=3, filt=2)
    ref_chain = np.concatenate([refshim.compress(fmt, lv, lv.shape[1], lv.shape[0]) for lv in levels])
    raw = np.frombuffer(open(j["out_path"], "rb").read(), np.uint8)
    want = np.frombuffer(oracle_texture_file(2, w, h, 1, mips, fmt, 0, ref_chain.tobytes()), np.uint8)
    assert raw.size == want.size
    assert np.array_equal(raw[:30], want[:30])
    # linear uint8 sources: the device cascade is bit-identical to stbir and the encoders to rgbcx, so the whole file is the
    # reference pipeline's file -- zero differing blocks, not a bound
    bb = 8 if fmt == BC4 else 16
    n0 = (w // 4) * (h // 4) * bb
    assert np.array_equal(raw[30:30 + n0], want[30:30 + n0])
    diff_blocks = np.any((raw[30:] != want[30:]).reshape(-1, bb), axis=1)
    print(f"fmt {fmt}: {diff_blocks.sum()} of {diff_blocks.size} blocks differ from the reference pipeline's file")
    assert diff_blocks.sum() == 0 and np.array_equal(raw, want)


def _write_pnm(path, img):
    h, w, c = img.shape
    with open(path, "wb") as f:
        f.write((b"P5" if c == 1 else b"P6") + f"\n# test\n{w} {h}\n255\n".encode())
        f.write(np.ascontiguousarray(img).tobytes())


def test_importer_tool_batch_manifest(tmp_path):
    """hk_texture_import -batch: the reference tool's flags per line; files equal the library calls' results"""
    import os
    import subprocess
    tool = os.path.join(os.path.dirname(hkbcn.LIB_PATH), "hk_texture_import")
    assert os.path.exists(tool), "hk_texture_import not built (python __graft_entry__.py build)"
    rough = synth.roughness_r8(256, 256, 2).reshape(256, 256, 1)
    albedo = synth.albedo_rgba8(128, 64, 0xB2000001, "opaque").reshape(64, 128, 4)
    _write_pnm(tmp_path / "rough.pgm", rough)
    _write_pnm(tmp_path / "albedo.ppm", albedo[:, :, :3])
    manifest = tmp_path / "batch.txt"
    manifest.write_text(
        f"# roughness: mipped with the triangle filter\n"
        f"-s {tmp_path}/rough.pgm -o {tmp_path}/rough.texture -format BC4_UNORM -mip_filter triangle -mip_edge_mode wrap\n"
        f"-s {tmp_path}/albedo.ppm -o {tmp_path}/albedo.texture -format BC1_UNORM_SRGB -no_alpha -mip_filter mitchell\n"
        f"-s {tmp_path}/albedo.ppm -o {tmp_path}/albedo7.texture -format BC7_UNORM   # no mips\n"
        f"-s gen:sky:64:64:3 -o {tmp_path}/sky.texture -format BC6H_UFLOAT\n")
    r = subprocess.run([tool, "-batch", str(manifest)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    rgba = np.concatenate([albedo[:, :, :3], np.full((64, 128, 1), 255, np.uint8)], axis=2)
    d, p = hkbcn.read_texture(str(tmp_path / "rough.texture"))
    m4 = hkbcn.lib().hkbcn_calc_num_mips(256, 256)
    assert (d.fmt, d.width, d.height, d.num_mips, d.storage_flags) == (BC4, 256, 256, m4, 0)
    assert np.array_equal(p, hkbcn.import_chain(BC4, rough, 256, 256, m4, hkbcn.MipConfig(3, 2, 0, 0)))
    d, p = hkbcn.read_texture(str(tmp_path / "albedo.texture"))
    m1 = hkbcn.lib().hkbcn_calc_num_mips(128, 64)
    assert (d.fmt, d.width, d.height, d.num_mips, d.storage_flags) == (BC1S, 128, 64, m1, 1)
    assert np.array_equal(p, hkbcn.import_chain(BC1S, rgba, 128, 64, m1, hkbcn.MipConfig(0, 5, 1, 0)))
    d, p = hkbcn.read_texture(str(tmp_path / "albedo7.texture"))
    assert (d.fmt, d.num_mips) == (BC7, 1)
    assert np.array_equal(p, hkbcn.compress(BC7, rgba, 128, 64))
    d, p = hkbcn.read_texture(str(tmp_path / "sky.texture"))
    assert (d.fmt, d.width, d.num_mips, p.nbytes) == (64, 64, 1, 16 * 16 * 16)
    # the reference tool's error behaviour: a missing output name is an error before any work
    r = subprocess.run([tool, "-s", "gen:albedo:64:64:1", "-format", "BC7_UNORM"], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "Output file is not specified" in r.stderr


def _write_pfm(path, rgb):
    """little-endian colour PFM, bottom row first"""
    h, w, _ = rgb.shape
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(np.ascontiguousarray(rgb[::-1].astype("<f4")).tobytes())


def test_importer_tool_cubemap_mode(tmp_path):
    """hk_texture_import -cubemap = Tools/CubemapImporter (Main.cpp:58-145 -> LoadSkyboxImages, Image.cpp:1918-2040): six faces ->
    TEXTURE_CUBE storage, six slices, one mip, IMAGE_STORAGE_NO_ALPHA; each face's blocks are the reference's CompressBC6h /
    CompressBC1 of that face (BC6H against the zero-init build, see test_gpu_bc6h.py), -hdri_scale / -hdri_pow applied first."""
    import os
    import subprocess
    tool = os.path.join(os.path.dirname(hkbcn.LIB_PATH), "hk_texture_import")
    n, scale, pw = 32, 0.5, 1.25
    faces = [synth.sky_rgba32f(n, n, 0xB2000006, f).reshape(n, n, 4) for f in range(6)]
    names = []
    for f, img in enumerate(faces):
        names.append(str(tmp_path / f"face{f}.pfm"))
        _write_pfm(names[-1], img[:, :, :3])
    out = str(tmp_path / "sky.texture")
    r = subprocess.run([tool, "-cubemap", "-s", *names, "-o", out, "-format", "BC6H_UFLOAT", "-hdri_scale", str(scale), "-hdri_pow", str(pw)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    d, payload = hkbcn.read_texture(out)
    assert (d.type, d.width, d.height, d.depth, d.num_mips, d.fmt, d.storage_flags) == (5, n, n, 6, 1, 64, 1)
    want = []
    for img in faces:
        adj = img.copy()
        adj[:, :, :3] = np.power(adj[:, :, :3] * np.float32(scale), np.float32(pw)).astype(np.float32)
        want.append(refshim.compress(64, adj, n, n, lib=refshim.ref0()))
    want = np.concatenate(want)
    same = (payload.reshape(-1, 16) == want.reshape(-1, 16)).all(axis=1)
    # powf on the host vs numpy's float32 power may differ in the last bit of a few texels; the blocks must still agree
    # wherever the inputs do -- demand at least 98 % identical blocks and equal size
    assert payload.nbytes == want.nbytes and same.mean() >= 0.98, f"{100 * same.mean():.2f}% identical"
    # LDR skybox: BC1_UNORM_SRGB from 8-bit faces, bit-exact
    names8 = []
    faces8 = [synth.albedo_rgba8(n, n, 40 + f, "opaque").reshape(n, n, 4) for f in range(6)]
    for f, img in enumerate(faces8):
        names8.append(str(tmp_path / f"face{f}.ppm"))
        _write_pnm(names8[-1], img[:, :, :3])
    out8 = str(tmp_path / "sky8.texture")
    r = subprocess.run([tool, "-cubemap", "-s", *names8, "-o", out8, "-format", "BC1_UNORM_SRGB"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    d, payload = hkbcn.read_texture(out8)
    assert (d.type, d.depth, d.fmt, d.storage_flags) == (5, 6, BC1S, 1)
    want8 = np.concatenate([refshim.compress(BC1S, img, n, n) for img in faces8])
    raw = open(out8, "rb").read()
    assert raw == oracle_texture_file(5, n, n, 6, 1, BC1S, 1, want8.tobytes())
    # reference error behaviour
    r = subprocess.run([tool, "-cubemap", "-s", *names8[:3], "-o", out8], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0


def test_importer_tool_pbr_mode(tmp_path):
    """hk_texture_import -pbr_preset ... -normal ... = Tools/PBRImporter (Main.cpp:284-755): ORMX gather (defaults 255 255 0 255),
    RGBA8 mips with NO_ALPHA, vMF roughness update from the triangle-filtered normals, BC7 chain; normal map normalised, packed
    RG_BC5, the packed texels mipped with the triangle filter, BC5 chain.  Level 0 of both files is checked against the
    reference's own encoders on the gathered / packed image, the whole files against the library calls the tool composes
    (each of which has its own oracle test: test_gpu_mips, test_gpu_normals)."""
    import os
    import subprocess
    tool = os.path.join(os.path.dirname(hkbcn.LIB_PATH), "hk_texture_import")
    n = 64
    occl = synth.roughness_r8(n, n, 11).reshape(n, n, 1)
    rough_src = synth.albedo_rgba8(n, n, 12, "opaque").reshape(n, n, 4)[:, :, :3]  # roughness taken from channel 2 of an RGB file
    nrm8 = np.concatenate([synth.normals_rg8(n, n, 13).reshape(n, n, 2), np.full((n, n, 1), 255, np.uint8)], axis=2)
    _write_pnm(tmp_path / "occl.pgm", occl)
    _write_pnm(tmp_path / "rough.ppm", rough_src)
    _write_pnm(tmp_path / "normal.ppm", nrm8)
    out, nout = str(tmp_path / "orm.texture"), str(tmp_path / "normal.texture")
    r = subprocess.run([tool, "-occlusion", str(tmp_path / "occl.pgm"), "0", "-roughness", str(tmp_path / "rough.ppm"), "2",
                        "-pbr_preset", "ORMX_BC7", out, "-normal", str(tmp_path / "normal.ppm"), nout, "-a", "vMF",
                        "-normal_pack", "RG_BC5_COMPATIBLE", "-mip_edge_mode", "wrap", "-mip_filter", "mitchell"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    mips = hkbcn.lib().hkbcn_calc_num_mips(n, n)
    # what the tool composes
    ormx = np.empty((n, n, 4), np.uint8)
    ormx[:, :, 0] = occl[:, :, 0]
    ormx[:, :, 1] = rough_src[:, :, 2]
    ormx[:, :, 2] = 0
    ormx[:, :, 3] = 255
    normals = hkbcn.normalize_vectors((nrm8.astype(np.float32) / np.float32(255.0)).reshape(-1, 3))
    chain = hkbcn.generate_mips(BC7, ormx, n, n, mips, hkbcn.MipConfig(3, 5, 1, 0)).copy()
    rough = hkbcn.bake_roughness(normals, chain[1::4], n, n, mips, 1, 2, 3)
    assert not np.array_equal(rough, chain[1::4]), "the bake changed nothing: test content too flat"
    chain[1::4] = rough
    d, payload = hkbcn.read_texture(out)
    assert (d.type, d.width, d.height, d.depth, d.num_mips, d.fmt, d.storage_flags) == (2, n, n, 1, mips, BC7, 1)
    assert np.array_equal(payload, hkbcn.compress_chain(BC7, chain, n, n, mips))
    assert np.array_equal(payload[: n * n], refshim.compress(BC7, ormx, n, n)), "ORMX level 0 differs from the reference's CompressBC7"
    packed = hkbcn.pack_normals(1, normals)
    d, payload = hkbcn.read_texture(nout)
    assert (d.type, d.width, d.num_mips, d.fmt, d.storage_flags) == (2, n, mips, BC5, 1)
    assert np.array_equal(payload, hkbcn.import_chain(BC5, packed, n, n, mips, hkbcn.MipConfig(3, 2, 1, 0)))
    assert np.array_equal(payload[: n * n], refshim.compress(BC5, packed, n, n)), "normal level 0 differs from the reference's CompressBC5"
    # reference error behaviour: antialiasing needs equal sizes
    _write_pnm(tmp_path / "occl32.pgm", synth.roughness_r8(32, 32, 11).reshape(32, 32, 1))
    r = subprocess.run([tool, "-occlusion", str(tmp_path / "occl32.pgm"), "0", "-pbr_preset", "ORMX_BC1", out, "-normal", str(tmp_path / "normal.ppm"),
                        nout, "-a", "vMF", "-mip_filter", "mitchell"], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "must be the same size" in r.stderr


def test_import_batch_skip_existing_and_atomic_write(tmp_path):
    """SURVEY.md 5 (checkpoint / resume): files appear under their final name only when complete, and a re-run with
    HKBCN_FLAG_SKIP_EXISTING leaves complete outputs alone (status HKBCN_SKIPPED) while redoing a truncated one"""
    import ctypes as C
    import os
    L = hkbcn.lib()
    imgs = [synth.albedo_rgba8(64, 64, 40 + i, "opaque") for i in range(3)]
    paths = [str(tmp_path / f"t{i}.texture").encode() for i in range(3)]
    SKIP = 1 << 16

    def run(flags):
        arr = (hkbcn.ImportJob * 3)()
        for i in range(3):
            arr[i] = hkbcn.ImportJob(BC1S, 64, 64, 1, hkbcn.MipConfig(), flags, imgs[i].ctypes.data, paths[i], None, 0)
        hkbcn.check(L.hkbcn_import_batch(arr, C.c_size_t(3)))
        return [arr[i].status for i in range(3)]
    assert run(SKIP) == [0, 0, 0]
    assert sorted(os.listdir(tmp_path)) == ["t0.texture", "t1.texture", "t2.texture"]  # no temporaries left behind
    first = [open(p, "rb").read() for p in paths]
    with open(paths[1], "r+b") as f:  # a crash half way through a write would never leave this, but a truncated copy might
        f.truncate(100)
    mt = [os.stat(p).st_mtime_ns for p in paths]
    assert run(SKIP) == [1, 0, 1]
    assert [open(p, "rb").read() for p in paths] == first
    assert os.stat(paths[0]).st_mtime_ns == mt[0] and os.stat(paths[2]).st_mtime_ns == mt[2]
    assert run(0) == [0, 0, 0]


def test_import_batch_device_sources_and_plan():
    """HKBCN_FLAG_SRC_DEVICE: level 0 already on the GPU the plan assigns the job to -- same chains as from host memory; the
    plan is the one hkbcn_plan_batch gives for the same shapes; per-GPU worker times are reported"""
    import ctypes as C
    import torch
    L = hkbcn.lib()
    hkbcn.init()
    nd = L.hkbcn_device_count()
    jobs = [(BC1S, synth.albedo_rgba8(128, 128, 60 + i)) for i in range(3)] + [(BC5, synth.normals_rg8(64, 64, 70))]
    arr = (hkbcn.ImportJob * len(jobs))()
    outs, keep = [], []
    for i, (fmt, img) in enumerate(jobs):
        h, w = img.shape[:2]
        arr[i] = hkbcn.ImportJob(fmt, w, h, 1, hkbcn.MipConfig(), hkbcn.FLAG_SRC_DEVICE, None, None, None, 0)
    plan = (C.c_int * len(jobs))()
    hkbcn.check(L.hkbcn_plan_import_batch(arr, C.c_size_t(len(jobs)), 0, plan))
    assert list(plan) == hkbcn.plan_batch([(f, im.shape[1], im.shape[0]) for f, im in jobs], nd)
    for i, (fmt, img) in enumerate(jobs):
        h, w = img.shape[:2]
        t = torch.from_numpy(np.ascontiguousarray(img).view(np.uint8).ravel().copy()).to(torch.device("cuda", plan[i]))
        dst = np.zeros(L.hkbcn_chain_bytes(fmt, w, h, 1, L.hkbcn_calc_num_mips(w, h)), np.uint8)
        keep.append(t)
        outs.append(dst)
        arr[i].src_level0 = t.data_ptr()
        arr[i].dst = dst.ctypes.data
    torch.cuda.synchronize()
    hkbcn.check(L.hkbcn_import_batch(arr, C.c_size_t(len(jobs))))
    for (fmt, img), got in zip(jobs, outs):
        h, w = img.shape[:2]
        assert np.array_equal(got, hkbcn.import_chain(fmt, img, w, h, L.hkbcn_calc_num_mips(w, h)))
    busy = (C.c_double * 8)()
    assert L.hkbcn_batch_stats(busy, 8) == nd and busy[0] > 0


def test_import_batch_sliced_jobs(tmp_path):
    """hkbcn_import_job.slices: a cube map (six faces, one mip: CubemapImporter) and a 2D array with per-slice cascades next to
    a plain 2D job in one batch; payloads in ImageStorage order (all slices of level 0, then of level 1, ...), files typed
    TEXTURE_CUBE / TEXTURE_2D_ARRAY with depth = slice count"""
    L = hkbcn.lib()
    BC1, BC6U, BC7 = 54, 64, 66
    w = 64
    faces = np.stack([synth.sky_rgba32f(w, w, 0x600 + f, f).reshape(w, w, 4) for f in range(6)])
    layers = np.stack([synth.albedo_rgba8(w, w, 0x610 + s, "mask").reshape(w, w, 4) for s in range(3)])
    flat = synth.albedo_rgba8(w, w, 0x620).reshape(w, w, 4)
    mips = L.hkbcn_calc_num_mips(w, w)
    paths = [str(tmp_path / n).encode() for n in ("cube.texture", "array.texture", "flat.texture")]
    jobs = (hkbcn.ImportJob * 3)()
    for j, (fmt, src, gen, slices, noa) in enumerate(((BC6U, faces, 0, 6, 1), (BC7, layers, 1, 3, 0), (BC1, flat, 1, 0, 0))):
        jobs[j].fmt, jobs[j].width, jobs[j].height, jobs[j].generate_mips, jobs[j].slices = fmt, w, w, gen, slices
        jobs[j].mip.no_alpha = noa
        jobs[j].src_level0 = np.ascontiguousarray(src).ctypes.data_as(C.c_void_p)
        jobs[j].out_path = paths[j]
    keep = [faces, layers, flat]  # the arrays must outlive the call
    hkbcn.check(L.hkbcn_import_batch(jobs, 3))
    assert [jobs[j].status for j in range(3)] == [0, 0, 0] and keep
    # cube: six faces through the reference's CompressBC6h, one after another
    d, payload = hkbcn.read_texture(paths[0].decode())
    assert (d.type, d.depth, d.num_mips, d.fmt, d.storage_flags) == (5, 6, 1, BC6U, 1)
    want = np.concatenate([refshim.compress(BC6U, faces[f], w, w, lib=refshim.ref0()) for f in range(6)])
    assert np.array_equal(payload, want)
    # array: every slice's own chain (device cascade = stbir for this linear format), interleaved level by level
    d, payload = hkbcn.read_texture(paths[1].decode())
    assert (d.type, d.depth, d.num_mips, d.fmt) == (3, 3, mips, BC7)
    chains = [hkbcn.import_chain(BC7, layers[s], w, w, mips) for s in range(3)]
    off, parts = 0, []
    for i in range(mips):
        n = max(w >> i, 4) ** 2
        parts += [c[off: off + n] for c in chains]
        off += n
    assert np.array_equal(payload, np.concatenate(parts))
    d, payload = hkbcn.read_texture(paths[2].decode())
    assert (d.type, d.depth) == (2, 1) and np.array_equal(payload, hkbcn.import_chain(BC1, flat, w, w, mips))


def test_tool_resample_flags(tmp_path):
    """-resample / -resample_edge_mode / -resample_filter of the TextureImporter tool (Main.cpp:145-192): the source is scaled
    on the device by hkbcn_resample_raw_image before the import; 96 x 80 -> 128 x 64 with mips (closest powers of two)"""
    import subprocess
    import os
    tool = os.path.join(os.path.dirname(hkbcn.LIB_PATH), "hk_texture_import")
    w, h = 96, 80
    img = synth.albedo_rgba8(w, h, 0x700, "opaque").reshape(h, w, 4)
    ppm = tmp_path / "src.ppm"
    with open(ppm, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(img[..., :3]).tobytes())
    out = tmp_path / "out.texture"
    r = subprocess.run([tool, "-s", str(ppm), "-o", str(out), "-format", "BC1_UNORM", "-mip_filter", "triangle", "-resample", "100", "70",
                        "-resample_edge_mode", "clamp", "clamp", "-resample_filter", "catmullrom", "catmullrom"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    d, payload = hkbcn.read_texture(str(out))
    assert (d.width, d.height, d.fmt) == (128, 64, 54)
    rgba = np.concatenate([img[..., :3], np.full((h, w, 1), 255, np.uint8)], -1)
    big = refshim.stbir_resize(rgba, 128, 64, alpha_channel=3, flags=0, edge=1, filt=4, space=0)
    levels = ref_cascade(np.ascontiguousarray(big), d.num_mips, 3, False, edge=3, filt=2)
    want = refshim.compress_chain(54, np.concatenate([l.ravel() for l in levels]), 128, 64, d.num_mips)
    assert np.array_equal(payload, want)
    r = subprocess.run([tool, "-s", str(ppm), "-o", str(out), "-format", "BC1_UNORM", "-resample", "5000", "70"], capture_output=True, text=True)
    assert r.returncode != 0 and "too large" in (r.stdout + r.stderr)


def test_import_batch_takes_a_prebuilt_chain():
    """HKBCN_FLAG_SRC_CHAIN: the job hands over its own uncompressed chain (here: the reference's stbir cascade); every level is
    uploaded and encoded, nothing is generated -- the in-memory result equals hkbcn_compress_chain and the reference"""
    L = hkbcn.lib()
    BC7 = 66
    w = h = 128
    mips = L.hkbcn_calc_num_mips(w, h)
    outs, chains, jobs = [], [], (hkbcn.ImportJob * 3)()
    for i in range(3):
        img = synth.albedo_rgba8(w, h, 0x730 + i, "mask").reshape(h, w, 4)
        chain = np.concatenate([l.ravel() for l in ref_cascade(img, mips, 3, False)])
        chains.append(np.ascontiguousarray(chain))
        outs.append(np.zeros(L.hkbcn_chain_bytes(BC7, w, h, 1, mips), np.uint8))
        jobs[i].fmt, jobs[i].width, jobs[i].height, jobs[i].generate_mips = BC7, w, h, 0
        jobs[i].flags = 1 << 18
        jobs[i].src_level0 = chains[i].ctypes.data_as(C.c_void_p)
        jobs[i].out_path = None
        jobs[i].dst = outs[i].ctypes.data_as(C.c_void_p)
    hkbcn.check(L.hkbcn_import_batch(jobs, 3))
    for i in range(3):
        assert jobs[i].status == 0
        assert np.array_equal(outs[i], hkbcn.compress_chain(BC7, chains[i], w, h, mips))
        assert np.array_equal(outs[i], refshim.compress_chain(BC7, chains[i], w, h, mips))
