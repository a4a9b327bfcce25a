"""TEST INFRASTRUCTURE: ctypes loader for oracle/_ref/libhkref.so (the unmodified reference codecs
behind a C shim, built by oracle/Makefile) and oracle/liboracle.so (the plain-C restatement)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libhkref.so")
REF0_PATH = os.path.join(ROOT, "oracle", "_ref", "libhkref0.so")  # same sources, -ftrivial-auto-var-init=zero (oracle/Makefile)
ORACLE_PATH = os.path.join(ROOT, "oracle", "liboracle.so")

# hkbcn / Hk::TEXTURE_FORMAT id -> shim codec id (ref_shim.cpp: F_BC1.. = 0..7)
_FMT = {54: 0, 55: 0, 56: 1, 57: 1, 58: 2, 59: 2, 60: 3, 61: 3, 62: 4, 63: 4, 64: 5, 65: 6, 66: 7, 67: 7}
_SRC_BPP = {0: 4, 1: 4, 2: 4, 3: 1, 4: 2, 5: 16, 6: 16, 7: 4}
_BLK = {0: 8, 1: 16, 2: 16, 3: 8, 4: 16, 5: 16, 6: 16, 7: 16}

_ref = None
_ref0 = None


def have_ref():
    return os.path.exists(REF_PATH)


def _load(path):
    if True:
        L = C.CDLL(path)
        L.ref_compress.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int]
        L.ref_decompress.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_size_t, C.c_int, C.c_int]
        L.ref_encode_block.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.ref_decode_block.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        L.ref_encode_bc4.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.ref_encode_bc5.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.ref_encode_bc6h_f16.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.ref_decode_bc7_bcdec.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_stbir_resize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_int] * 7
        L.ref_half_from_float.restype = C.c_uint16
        L.ref_half_from_float.argtypes = [C.c_uint32]
        L.ref_half_to_float.restype = C.c_uint32
        L.ref_half_to_float.argtypes = [C.c_uint16]
    return L


def ref():
    global _ref
    if _ref is None:
        _ref = _load(REF_PATH)
    return _ref


def ref0():
    """The reference built with every automatic variable zero-initialised: the bit-exact oracle of the BC6H encoder
    (bc6h_enc.h:3134-3150 reads an uninitialised `pad`; see oracle/Makefile)."""
    global _ref0
    if _ref0 is None:
        _ref0 = _load(REF0_PATH)
    return _ref0


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def hw_threads():
    return ref().ref_hw_threads()


def compress(fmt, src, w, h, nthreads=0, lib=None):
    """reference Compress* over the surface; nthreads=0 -> all hardware threads."""
    f = _FMT[fmt]
    src = np.ascontiguousarray(src)
    dst = np.zeros((w // 4) * (h // 4) * _BLK[f], np.uint8)
    (lib or ref()).ref_compress(f, _p(src), _p(dst), w, h, nthreads if nthreads else hw_threads())
    return dst


def encode_bc6h_f16_blocks(src_half, n, signed=False, lib=None):
    """Encode_BC6h_f16 (ImageEncoders.cpp:187-197) over n packed blocks of 16 RGBA16F texels."""
    src_half = np.ascontiguousarray(src_half)
    dst = np.zeros(n * 16, np.uint8)
    L = lib or ref()
    for i in range(n):
        L.ref_encode_bc6h_f16(src_half.ctypes.data + i * 128, dst.ctypes.data + i * 16, int(signed))
    return dst


def compress_chain(fmt, src, w, h, mips, nthreads=0):
    f = _FMT[fmt]
    src = np.ascontiguousarray(src).view(np.uint8).ravel()
    outs, off = [], 0
    for i in range(mips):
        lw, lh = max(w >> i, 4), max(h >> i, 4)
        n = lw * lh * _SRC_BPP[f]
        outs.append(compress(fmt, src[off:off + n], lw, lh, nthreads))
        off += n
    return np.concatenate(outs)


def decompress(fmt, blocks, w, h, f32out=False, nthreads=0, row_stride=None):
    f = _FMT[fmt]
    tb = {3: 1, 4: 2, 5: 12 if f32out else 6, 6: 12 if f32out else 6}.get(f, 4)
    if row_stride is None:
        row_stride = (w // 4) * 4 * tb
    blocks = np.ascontiguousarray(blocks)
    dst = np.zeros(row_stride * (h // 4) * 4, np.uint8)
    ref().ref_decompress(f, _p(blocks), _p(dst), w, h, row_stride, int(f32out), nthreads if nthreads else hw_threads())
    return dst


def encode_bc45_fast(fmt, src, n):
    """non-hq rgbcx::encode_bc4 / encode_bc5 over n packed blocks"""
    f = _FMT[fmt]
    src = np.ascontiguousarray(src)
    bpp, bb = _SRC_BPP[f], _BLK[f]
    dst = np.zeros(n * bb, np.uint8)
    fn = ref().ref_encode_bc4 if f == 3 else ref().ref_encode_bc5
    for i in range(n):
        fn(src.ctypes.data + i * 16 * bpp, dst.ctypes.data + i * bb, 0)
    return dst


def stbir_resize(img, ow, oh, alpha_channel=-1, flags=0, edge=2, filt=4, space=0):
    """img: (h, w, c) uint8 or float32. defaults: wrap edge (STBIR_EDGE_WRAP=2), Mitchell (4), linear."""
    ih, iw, nc = img.shape
    dt = 3 if img.dtype == np.float32 else 0
    img = np.ascontiguousarray(img)
    out = np.zeros((oh, ow, nc), img.dtype)
    ok = ref().ref_stbir_resize(_p(img), iw, ih, _p(out), ow, oh, dt, nc, alpha_channel, flags, edge, filt, space)
    assert ok == 1
    return out


def psnr(a, b, peak=255.0):
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    mse = np.mean((a - b) ** 2)
    return 99.0 if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def encode_bc1_level_blocks(src_rgba, n, level, allow3=False, transparent_black=False):
    """rgbcx::encode_bc1(level, ..) over n packed blocks (Encode_BC1(.., Level, b3ColorMode, bTransparentPixelsForBlack))"""
    src = np.ascontiguousarray(src_rgba, np.uint8).ravel()
    dst = np.zeros(n * 8, np.uint8)
    L = ref()
    L.ref_encode_bc1_level.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int]
    for i in range(n):
        L.ref_encode_bc1_level(src.ctypes.data + i * 64, dst.ctypes.data + i * 8, level, int(allow3), int(transparent_black))
    return dst


def encode_bc7_level_blocks(src_rgba, n, level):
    """bc7enc_compress_block with bc7_params[level] over n packed blocks (Encode_BC7(.., Level))"""
    src = np.ascontiguousarray(src_rgba, np.uint8).ravel()
    dst = np.zeros(n * 16, np.uint8)
    L = ref()
    L.ref_encode_bc7_level.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    for i in range(n):
        L.ref_encode_bc7_level(src.ctypes.data + i * 64, dst.ctypes.data + i * 16, level)
    return dst
