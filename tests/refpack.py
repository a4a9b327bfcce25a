"""TEST INFRASTRUCTURE: ctypes loader for oracle/_ref/libhkref_pack.so -- the reference's OWN normal-map packers / unpackers
(Hork/Image/ImageEncoders.cpp:692-878), NormalizeVectors (Image.cpp:2251-2260) and roughness update (Image.cpp:2541-2579),
compiled from the reference tree by oracle/Makefile (oracle/ref_pack_shim.cpp).  Same function names as oracle_c, so a test
can take either as its checker; this one is the reference's object code, oracle_c the plain-C restatement."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "_ref", "libhkref_pack.so")
_lib = None


def have():
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(PATH)
        L.ref_pack_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        L.ref_unpack_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        L.ref_normalize_vectors.argtypes = [C.c_void_p, C.c_size_t]
        L.ref_bake_roughness_level.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pack_normals(variant, normals):
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    out = np.zeros(n.shape[0] * (4 if variant in (0, 5) else 2), np.uint8)
    assert lib().ref_pack_normals(variant, _p(n), _p(out), n.shape[0]) == 0
    return out


def unpack_normals(variant, packed, count):
    p = np.ascontiguousarray(packed, np.uint8)
    out = np.zeros((count, 3), np.float32)
    assert lib().ref_unpack_normals(variant, _p(p), _p(out), count) == 0
    return out


def normalize_vectors(v):
    out = np.array(v, np.float32).reshape(-1, 3).copy()
    lib().ref_normalize_vectors(_p(out), out.shape[0])
    return out


def bake_roughness_level(bake, avg_normals, rough):
    n = np.ascontiguousarray(avg_normals, np.float32).reshape(-1, 3)
    r = np.array(rough, np.uint8).ravel().copy()
    lib().ref_bake_roughness_level(bake, _p(n), _p(r), r.size)
    return r
