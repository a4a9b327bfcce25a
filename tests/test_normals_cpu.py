"""Row f4, CPU side: the C restatement of the packers (oracle/oracle_pack.c) against hand-computed known answers and
round trips, so the checker the GPU tests lean on is itself checked.  (The reference has no tests or vectors for these
functions: parity of the restatement is by construction, statement by statement.)"""
import numpy as np

import oracle_c


def test_known_answers():
    # Math::Round(Math::Saturate(v * 0.5 + 0.5) * 255): 0 -> 127.5 -> 128 (half away from zero), 1 -> 255, -1 -> 0
    up = np.array([[0, 0, 1]], np.float32)
    assert oracle_c.pack_normals(0, up).tolist() == [128, 128, 255, 255]
    assert oracle_c.pack_normals(5, up).tolist() == [255, 128, 255, 128]
    assert oracle_c.pack_normals(1, up).tolist() == [128, 128]
    assert oracle_c.pack_normals(1, np.array([[1, -1, 0]], np.float32)).tolist() == [255, 0]
    # spheremap: denom = 1 / sqrt(8 z + 8); z = 1 -> 0.25; x = 0.6 -> 0.65 * 255 = 165.75 -> 166
    assert oracle_c.pack_normals(2, np.array([[0.6, 0.0, 1.0]], np.float32)).tolist() == [166, 128]
    # stereographic: x / (1 + z) * 0.5 + 0.5 with z = 0: x = 0.6 -> 0.8 * 255 = 204
    assert oracle_c.pack_normals(3, np.array([[0.6, 0.8, 0.0]], np.float32)).tolist() == [204, 230]  # 0.9 * 255 = 229.5 -> 230
    # out-of-range components saturate
    assert oracle_c.pack_normals(0, np.array([[3, -3, 0.2]], np.float32)).tolist() == [255, 0, 153, 255]
    # NormalizeVectors: v * 2 - 1, normalise; the zero vector stays zero
    out = oracle_c.normalize_vectors(np.array([[0.5, 0.5, 1.0], [0.5, 0.5, 0.5], [1.0, 0.5, 0.5]], np.float32))
    assert out.tolist() == [[0, 0, 1], [0, 0, 0], [1, 0, 0]]


def test_round_trips_are_within_quantisation():
    rng = np.random.default_rng(2)
    v = rng.normal(size=(20000, 3)).astype(np.float32)
    v[:, 2] = np.abs(v[:, 2]) + 0.05
    v /= np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    for variant, tol in ((0, 0.005), (1, 0.03), (2, 0.02), (3, 0.02), (5, 0.005)):
        back = oracle_c.unpack_normals(variant, oracle_c.pack_normals(variant, v), v.shape[0])
        back /= np.maximum(np.linalg.norm(back, axis=1, keepdims=True), 1e-6)
        err = np.linalg.norm(back - v, axis=1)
        assert np.percentile(err, 99) < tol * 4 and np.median(err) < tol, (variant, float(np.median(err)))


def test_bake_raises_roughness_where_normals_disagree():
    # a shortened averaged normal (|n| < 1) means variance below the texel: roughness goes up, never down
    avg = np.array([[0, 0, 1.0], [0, 0, 0.9], [0, 0, 0.5], [0, 0, 0.0]], np.float32)
    rough = np.array([100, 100, 100, 100], np.uint8)
    for bake in (1, 2):
        out = oracle_c.bake_roughness_level(bake, avg, rough)
        assert out[0] == 100 and out[3] == 100      # r2 == 1 and r2 <= 1e-8 are left alone
        assert out[1] > 100 and out[2] > out[1]


import pytest  # noqa: E402

import refpack  # noqa: E402


@pytest.mark.skipif(not refpack.have(), reason="oracle/_ref/libhkref_pack.so not built")
def test_restatement_equals_the_reference_object_code():
    """oracle/oracle_pack.c against the reference's own functions compiled from the reference tree (oracle/ref_pack_shim.cpp:
    ImageEncoders.cpp:692-878, Image.cpp:2251-2260, :2541-2579 with the engine's own Float3 / Math::): bit-exact on packers,
    unpackers, NormalizeVectors and both roughness bakes -- which pins the restatement, and with it row f4"""
    rng = np.random.default_rng(7)
    v = rng.normal(size=(200000, 3)).astype(np.float32)
    v[::2, 2] = np.abs(v[::2, 2])
    v /= np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    edge = np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, -1, 0], [0, 0, -1], [0.6, 0.8, 0], [0, 0, 0], [2, -3, 0.5],
                     [1e-20, 0, 1], [-0.0, 0.0, 1]], np.float32)
    v = np.concatenate([edge, v]).astype(np.float32)
    with np.errstate(all="ignore"):
        for variant in range(6):
            assert np.array_equal(oracle_c.pack_normals(variant, v), refpack.pack_normals(variant, v)), f"packer {variant}"
            tb = 4 if variant in (0, 5) else 2
            packed = rng.integers(0, 256, 100000 * tb, dtype=np.uint8)
            if tb == 2:
                a, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8))
                packed = np.concatenate([packed, np.stack([a.ravel(), b.ravel()], 1).ravel()])
            n = packed.size // tb
            assert np.array_equal(oracle_c.unpack_normals(variant, packed, n).view(np.uint32), refpack.unpack_normals(variant, packed, n).view(np.uint32)), f"unpacker {variant}"
    u = rng.random((200000, 3), dtype=np.float32)
    u[:3] = [[0.5, 0.5, 0.5], [1, 1, 1], [0, 0, 0]]
    assert np.array_equal(oracle_c.normalize_vectors(u).view(np.uint32), refpack.normalize_vectors(u).view(np.uint32))
    avg = (rng.normal(size=(300000, 3)) * rng.random((300000, 1))).astype(np.float32)  # lengths on both sides of 1
    avg[:4] = [[0, 0, 1], [0, 0, 0.9], [0, 0, 0], [0, 0, 1e-5]]
    rough = rng.integers(0, 256, 300000, dtype=np.uint8)
    for bake in (1, 2):
        assert np.array_equal(oracle_c.bake_roughness_level(bake, avg, rough), refpack.bake_roughness_level(bake, avg, rough)), f"bake {bake}"
