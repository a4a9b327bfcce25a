"""TEST INFRASTRUCTURE: ctypes loader for oracle/liboracle.so (plain-C restatement)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "liboracle.so")
_FMT = {54: 0, 55: 0, 56: 1, 57: 1, 58: 2, 59: 2, 60: 3, 61: 3, 62: 4, 63: 4, 64: 5, 65: 6, 66: 7, 67: 7}
_BLK = {0: 8, 1: 16, 2: 16, 3: 8, 4: 16, 5: 16, 6: 16, 7: 16}
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(PATH)
        L.oracle_compress.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.oracle_decompress.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_size_t, C.c_int]
        L.oracle_encode_bc4_fast.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.oracle_envprobe_prefilter.restype = C.c_size_t
        L.oracle_envprobe_prefilter.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.oracle_irradiance.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.oracle_pack_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.oracle_unpack_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.oracle_normalize_vectors.argtypes = [C.c_void_p, C.c_size_t]
        L.oracle_bake_roughness_level.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        _lib = L
    return _lib


def compress(fmt, src, w, h):
    f = _FMT[fmt]
    src = np.ascontiguousarray(src)
    dst = np.zeros((w // 4) * (h // 4) * _BLK[f], np.uint8)
    rc = lib().oracle_compress(f, src.ctypes.data, dst.ctypes.data, w, h)
    assert rc == 0, "format has no restated encoder"
    return dst


def decompress(fmt, blocks, w, h, f32out=False, row_stride=None):
    f = _FMT[fmt]
    tb = {3: 1, 4: 2, 5: 12 if f32out else 6, 6: 12 if f32out else 6}.get(f, 4)
    if row_stride is None:
        row_stride = (w // 4) * 4 * tb
    blocks = np.ascontiguousarray(blocks)
    dst = np.zeros(row_stride * (h // 4) * 4, np.uint8)
    assert lib().oracle_decompress(f, blocks.ctypes.data, dst.ctypes.data, w, h, row_stride, int(f32out)) == 0
    return dst


def encode_bc4_fast(src, n):
    src = np.ascontiguousarray(src)
    dst = np.zeros(n * 8, np.uint8)
    for i in range(n):
        lib().oracle_encode_bc4_fast(src.ctypes.data + 16 * i, 1, dst.ctypes.data + 8 * i)
    return dst


def pack_normals(variant, normals):
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    out = np.zeros(n.shape[0] * (4 if variant in (0, 5) else 2), np.uint8)
    assert lib().oracle_pack_normals(variant, n.ctypes.data, out.ctypes.data, n.shape[0]) == 0
    return out


def unpack_normals(variant, packed, count):
    p = np.ascontiguousarray(packed, np.uint8)
    out = np.zeros((count, 3), np.float32)
    assert lib().oracle_unpack_normals(variant, p.ctypes.data, out.ctypes.data, count) == 0
    return out


def normalize_vectors(v):
    out = np.array(v, np.float32).reshape(-1, 3).copy()
    lib().oracle_normalize_vectors(out.ctypes.data, out.shape[0])
    return out


def bake_roughness_level(bake, avg_normals, rough):
    a = np.ascontiguousarray(avg_normals, np.float32).reshape(-1, 3)
    r = np.array(rough, np.uint8).ravel().copy()
    lib().oracle_bake_roughness_level(bake, a.ctypes.data, r.ctypes.data, r.size)
    return r


def envprobe_prefilter(cube, src_size, max_lod):
    c = np.ascontiguousarray(cube, np.float32)
    n = sum(6 * ((1 << max_lod) >> l) ** 2 for l in range(max_lod + 1))
    out = np.zeros(n * 4, np.float32)
    assert lib().oracle_envprobe_prefilter(c.ctypes.data, src_size, max_lod, out.ctypes.data) == n
    levels, off = [], 0
    for lod in range(max_lod + 1):
        s = (1 << max_lod) >> lod
        levels.append(out[off:off + 6 * s * s * 4].reshape(6, s, s, 4))
        off += 6 * s * s * 4
    return levels


def irradiance(cube, src_size, out_size):
    c = np.ascontiguousarray(cube, np.float32)
    out = np.zeros((6, out_size, out_size, 4), np.float32)
    lib().oracle_irradiance(c.ctypes.data, src_size, out_size, out.ctypes.data)
    return out
