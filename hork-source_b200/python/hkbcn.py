"""ctypes binding of libhkbcn.so (include/hkbcn.h) -- test / bench glue only.

The product is the C-ABI library and the C++ shim in hork-source_b200/host; this module
lets tests/ and bench.py call the same entry points.  Function names mirror
Hk::TextureBlockCompression (reference: Hork/Image/ImageEncoders.h:37-87).

There is no CPU fallback: if the library or a B200 is missing, calls raise HkbcnError.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# HKBCN_LIB: an alternative build of the same library (kernel experiments from tools/), never a different implementation
LIB_PATH = os.environ.get("HKBCN_LIB") or os.path.join(os.path.dirname(_HERE), "libhkbcn.so")

# Hk::TEXTURE_FORMAT values (Hork/Image/Image.h:122-148)
BC1_UNORM, BC1_UNORM_SRGB, BC2_UNORM, BC2_UNORM_SRGB, BC3_UNORM, BC3_UNORM_SRGB = 54, 55, 56, 57, 58, 59
BC4_UNORM, BC4_SNORM, BC5_UNORM, BC5_SNORM, BC6H_UFLOAT, BC6H_SFLOAT, BC7_UNORM, BC7_UNORM_SRGB = 60, 61, 62, 63, 64, 65, 66, 67

FLAG_SRC_F16 = 1 << 0
FLAG_FAST = 1 << 1
FLAG_DST_F32 = 1 << 8
FLAG_BC1_3COLOR = 1 << 9
BUILD_BC1_ORDERINGS, BUILD_STBIR_SRGB = 1, 2
FLAG_BC1_TRANSPARENT_BLACK = 1 << 10
FLAG_SKIP_EXISTING = 1 << 16
FLAG_SRC_DEVICE = 1 << 17

EXPORTS = [
    "hkbcn_init", "hkbcn_shutdown", "hkbcn_device_count", "hkbcn_last_error", "hkbcn_version",
    "hkbcn_block_bytes", "hkbcn_src_texel_bytes", "hkbcn_dst_texel_bytes", "hkbcn_calc_num_mips", "hkbcn_chain_bytes",
    "hkbcn_compress", "hkbcn_decompress", "hkbcn_encode_blocks", "hkbcn_decode_blocks",
    "hkbcn_compress_dev", "hkbcn_decompress_dev", "hkbcn_compress_chain_dev", "hkbcn_compress_chain",
    "hkbcn_slices_src_bytes", "hkbcn_compress_slices_dev", "hkbcn_compress_slices", "hkbcn_compress_chain_multi", "hkbcn_plan_bands",
    "hkbcn_compress_batch", "hkbcn_plan_batch", "hkbcn_launch_count",
    "hkbcn_decoded_chain_bytes", "hkbcn_decompress_slices_dev", "hkbcn_decompress_slices", "hkbcn_decompress_chain_dev", "hkbcn_decompress_chain",
    "hkbcn_decompress_batch", "hkbcn_plan_import_batch", "hkbcn_batch_stats",
    "hkbcn_mip_taps", "hkbcn_chain_src_bytes", "hkbcn_generate_mips_dev", "hkbcn_generate_mips", "hkbcn_import_chain_dev", "hkbcn_import_chain", "hkbcn_alloc_pinned", "hkbcn_free_pinned",
    "hkbcn_texture_header", "hkbcn_write_texture", "hkbcn_read_texture", "hkbcn_import_batch",
    "hkbcn_envprobe_texels", "hkbcn_envprobe_prefilter", "hkbcn_irradiance", "hkbcn_envprobe_prefilter_dev", "hkbcn_irradiance_dev",
    "hkbcn_pack_texel_bytes", "hkbcn_pack_normals", "hkbcn_unpack_normals", "hkbcn_normalize_vectors", "hkbcn_bake_roughness",
    "hkbcn_pack_normals_dev", "hkbcn_unpack_normals_dev", "hkbcn_normalize_vectors_dev", "hkbcn_bake_scratch_bytes", "hkbcn_bake_roughness_dev",
    "hkbcn_static_scratch_bytes", "hkbcn_compress_static",
    "hkbcn_plan_image", "hkbcn_create_image", "hkbcn_create_roughness_map", "hkbcn_resample_raw_image", "hkbcn_resample_raw_image_dev",
    "hkbcn_create_normal_map", "hkbcn_create_normal_and_roughness", "hkbcn_import_pbr_batch", "hkbcn_build_has", "hkbcn_srgb_decode_entry",
]


class HkbcnError(RuntimeError):
    pass


class ResampleParams(C.Structure):
    _fields_ = [("flags", C.c_uint32), ("edge_h", C.c_int), ("edge_v", C.c_int), ("filter_h", C.c_int), ("filter_v", C.c_int),
                ("scaled_width", C.c_uint32), ("scaled_height", C.c_uint32)]


class NormalRoughnessSettings(C.Structure):
    _fields_ = [("normals_rgb32f", C.c_void_p), ("normal_width", C.c_uint32), ("normal_height", C.c_uint32),
                ("roughness_r8", C.c_void_p), ("roughness_width", C.c_uint32), ("roughness_height", C.c_uint32),
                ("pack", C.c_int), ("convert_from_directx", C.c_int), ("edge_mode", C.c_int), ("roughness_bake", C.c_int)]


class ImagePlan(C.Structure):
    _fields_ = [("format", C.c_int), ("storage_flags", C.c_uint32), ("src_channels", C.c_int), ("add_alpha", C.c_int), ("swap_rb", C.c_int),
                ("required_width", C.c_uint32), ("required_height", C.c_uint32)]


class MipConfig(C.Structure):
    _fields_ = [("edge_mode", C.c_int), ("filter", C.c_int), ("no_alpha", C.c_int), ("alpha_premultiplied", C.c_int)]


class Job(C.Structure):
    _fields_ = [("fmt", C.c_int), ("width", C.c_uint32), ("height", C.c_uint32), ("flags", C.c_uint32),
                ("src", C.c_void_p), ("dst", C.c_void_p), ("status", C.c_int)]


class TextureDesc(C.Structure):
    _fields_ = [("type", C.c_int), ("width", C.c_uint32), ("height", C.c_uint32), ("depth", C.c_uint32), ("num_mips", C.c_uint32),
                ("fmt", C.c_int), ("storage_flags", C.c_uint32)]


class ImportJob(C.Structure):
    _fields_ = [("fmt", C.c_int), ("width", C.c_uint32), ("height", C.c_uint32), ("generate_mips", C.c_int), ("mip", MipConfig),
                ("flags", C.c_uint32), ("src_level0", C.c_void_p), ("out_path", C.c_char_p), ("dst", C.c_void_p), ("status", C.c_int), ("slices", C.c_uint32)]


class PbrJob(C.Structure):
    _fields_ = [("fmt", C.c_int), ("width", C.c_uint32), ("height", C.c_uint32), ("maps", C.c_void_p * 4), ("channel", C.c_uint8 * 4),
                ("normals", C.c_void_p), ("roughness_bake", C.c_int), ("generate_mips", C.c_int), ("mip", MipConfig),
                ("out_path", C.c_char_p), ("dst", C.c_void_p), ("status", C.c_int)]


_lib = None


def lib():
    """Load libhkbcn.so (built in-tree by __graft_entry__.build()); fail loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise HkbcnError(f"{LIB_PATH} not built -- run `python __graft_entry__.py build` (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        L.hkbcn_last_error.restype = C.c_char_p
        L.hkbcn_version.restype = C.c_char_p
        L.hkbcn_init.argtypes = [C.c_uint32]
        L.hkbcn_chain_bytes.restype = C.c_size_t
        L.hkbcn_chain_bytes.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_calc_num_mips.restype = C.c_uint32
        L.hkbcn_calc_num_mips.argtypes = [C.c_uint32, C.c_uint32]
        L.hkbcn_compress.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_decompress.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_size_t, C.c_uint32]
        L.hkbcn_encode_blocks.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint32]
        L.hkbcn_decode_blocks.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_uint32]
        L.hkbcn_compress_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.hkbcn_decompress_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_size_t, C.c_uint32, C.c_void_p]
        L.hkbcn_compress_chain_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.hkbcn_compress_chain.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_slices_src_bytes.restype = C.c_size_t
        L.hkbcn_slices_src_bytes.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_compress_slices_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.hkbcn_compress_slices.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_compress_chain_multi.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        L.hkbcn_plan_bands.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.hkbcn_compress_batch.argtypes = [C.POINTER(Job), C.c_size_t]
        L.hkbcn_decompress_batch.argtypes = [C.POINTER(Job), C.c_size_t]
        L.hkbcn_decoded_chain_bytes.restype = C.c_size_t
        L.hkbcn_decoded_chain_bytes.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_decompress_slices_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.hkbcn_decompress_slices.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_decompress_chain_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.hkbcn_decompress_chain.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_plan_batch.argtypes = [C.POINTER(Job), C.c_size_t, C.c_int, C.POINTER(C.c_int)]
        L.hkbcn_mip_taps.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.hkbcn_chain_src_bytes.restype = C.c_size_t
        L.hkbcn_chain_src_bytes.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.hkbcn_generate_mips_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(MipConfig), C.c_uint32, C.c_void_p]
        L.hkbcn_generate_mips.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(MipConfig), C.c_uint32]
        L.hkbcn_import_chain_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(MipConfig), C.c_uint32, C.c_void_p]
        L.hkbcn_import_chain.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(MipConfig), C.c_uint32]
        L.hkbcn_launch_count.restype = C.c_uint64
        L.hkbcn_alloc_pinned.restype = C.c_void_p
        L.hkbcn_texture_header.argtypes = [C.POINTER(TextureDesc), C.c_uint32, C.c_void_p]
        L.hkbcn_write_texture.argtypes = [C.c_char_p, C.POINTER(TextureDesc), C.c_void_p, C.c_size_t]
        L.hkbcn_read_texture.argtypes = [C.c_char_p, C.POINTER(TextureDesc), C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.hkbcn_import_batch.argtypes = [C.POINTER(ImportJob), C.c_size_t]
        L.hkbcn_import_pbr_batch.argtypes = [C.POINTER(PbrJob), C.c_size_t]
        L.hkbcn_build_has.argtypes = [C.c_int]
        L.hkbcn_srgb_decode_entry.argtypes = [C.c_int]
        L.hkbcn_srgb_decode_entry.restype = C.c_float
        L.hkbcn_plan_import_batch.argtypes = [C.POINTER(ImportJob), C.c_size_t, C.c_int, C.POINTER(C.c_int)]
        L.hkbcn_batch_stats.argtypes = [C.POINTER(C.c_double), C.c_int]
        L.hkbcn_envprobe_texels.restype = C.c_size_t
        L.hkbcn_envprobe_texels.argtypes = [C.c_int]
        L.hkbcn_envprobe_prefilter.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        L.hkbcn_irradiance.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.hkbcn_envprobe_prefilter_dev.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
        L.hkbcn_irradiance_dev.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.hkbcn_pack_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.hkbcn_unpack_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.hkbcn_normalize_vectors.argtypes = [C.c_void_p, C.c_size_t]
        L.hkbcn_bake_roughness.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int]
        L.hkbcn_pack_normals_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.hkbcn_unpack_normals_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.hkbcn_normalize_vectors_dev.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.hkbcn_bake_scratch_bytes.restype = C.c_size_t
        L.hkbcn_bake_scratch_bytes.argtypes = [C.c_uint32, C.c_uint32]
        L.hkbcn_bake_roughness_dev.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.hkbcn_alloc_pinned.argtypes = [C.c_size_t]
        L.hkbcn_free_pinned.argtypes = [C.c_void_p]
        L.hkbcn_resample_raw_image.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(ResampleParams), C.c_void_p]
        L.hkbcn_resample_raw_image_dev.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(ResampleParams), C.c_void_p, C.c_void_p]
        L.hkbcn_create_normal_map.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(TextureDesc), C.c_void_p, C.c_size_t]
        L.hkbcn_create_normal_and_roughness.argtypes = [C.POINTER(NormalRoughnessSettings), C.POINTER(TextureDesc), C.c_void_p, C.c_size_t,
                                                        C.POINTER(TextureDesc), C.c_void_p, C.c_size_t]
        L.hkbcn_plan_image.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(ImagePlan)]
        L.hkbcn_create_image.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(MipConfig), C.c_uint32, C.c_uint32, C.POINTER(TextureDesc), C.c_void_p, C.c_size_t]
        L.hkbcn_create_roughness_map.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.POINTER(TextureDesc), C.c_void_p, C.c_size_t]
        L.hkbcn_static_scratch_bytes.restype = C.c_size_t
        L.hkbcn_compress_static.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise HkbcnError(f"hkbcn status {rc}: {lib().hkbcn_last_error().decode()}")


def init(mask=0):
    check(lib().hkbcn_init(mask))


def block_bytes(fmt):
    return lib().hkbcn_block_bytes(fmt)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def compress(fmt, src, width, height, flags=0):
    """CompressBC*(pSrc, pDest, Width, Height): src = numpy array of the tightly packed surface."""
    src = np.ascontiguousarray(src)
    nb = (width // 4) * (height // 4)
    dst = np.zeros(nb * block_bytes(fmt), np.uint8)
    check(lib().hkbcn_compress(fmt, _ptr(src), _ptr(dst), width, height, flags))
    return dst


def compress_chain(fmt, src, width, height, mips, flags=0):
    src = np.ascontiguousarray(src)
    dst = np.zeros(lib().hkbcn_chain_bytes(fmt, width, height, 1, mips), np.uint8)
    check(lib().hkbcn_compress_chain(fmt, _ptr(src), _ptr(dst), width, height, mips, flags))
    return dst


def compress_slices(fmt, src, width, height, slices, mips, flags=0):
    """All slices (cube faces) and mip levels in one launch; src and result in ImageStorage order."""
    src = np.ascontiguousarray(src)
    assert src.nbytes == lib().hkbcn_slices_src_bytes(fmt, width, height, slices, mips, flags), "source size"
    dst = np.zeros(lib().hkbcn_chain_bytes(fmt, width, height, slices, mips), np.uint8)
    check(lib().hkbcn_compress_slices(fmt, _ptr(src), _ptr(dst), width, height, slices, mips, flags))
    return dst


def compress_chain_multi(fmt, src, width, height, mips, n_gpus=0, flags=0):
    """One chain over several GPUs in bands of block rows; n_gpus < 0 = that many bands over the selected GPUs."""
    src = np.ascontiguousarray(src)
    dst = np.zeros(lib().hkbcn_chain_bytes(fmt, width, height, 1, mips), np.uint8)
    check(lib().hkbcn_compress_chain_multi(fmt, _ptr(src), _ptr(dst), width, height, mips, flags, n_gpus))
    return dst


def plan_bands(height, mips, n_gpus, device):
    r0, r1 = (C.c_uint32 * mips)(), (C.c_uint32 * mips)()
    check(lib().hkbcn_plan_bands(height, mips, n_gpus, device, r0, r1))
    return list(r0), list(r1)


def encode_blocks(fmt, src, n, flags=0):
    """Encode_BC*(pSrc, pDest, ...) applied to n packed 4x4 blocks."""
    src = np.ascontiguousarray(src)
    dst = np.zeros(n * block_bytes(fmt), np.uint8)
    check(lib().hkbcn_encode_blocks(fmt, _ptr(src), _ptr(dst), n, flags))
    return dst


def decompress(fmt, blocks, width, height, flags=0, row_stride=None):
    blocks = np.ascontiguousarray(blocks)
    tb = lib().hkbcn_dst_texel_bytes(fmt, flags)
    if row_stride is None:
        row_stride = (width // 4) * 4 * tb
    dst = np.zeros(row_stride * (height // 4) * 4, np.uint8)
    check(lib().hkbcn_decompress(fmt, _ptr(blocks), _ptr(dst), width, height, row_stride, flags))
    return dst


def decode_blocks(fmt, blocks, n, row_stride, flags=0):
    """Decode_BC*(pSrc, pDest, RowStride) applied to n blocks (each into its own 4*row_stride bytes)."""
    blocks = np.ascontiguousarray(blocks)
    dst = np.zeros(n * 4 * row_stride, np.uint8)
    check(lib().hkbcn_decode_blocks(fmt, _ptr(blocks), _ptr(dst), n, row_stride, flags))
    return dst


def decompress_slices(fmt, blocks, width, height, slices, mips, flags=0):
    """whole storage (all slices of all levels, ImageStorage order) -> decoded levels back to back"""
    blocks = np.ascontiguousarray(blocks)
    dst = np.zeros(lib().hkbcn_decoded_chain_bytes(fmt, width, height, slices, mips, flags), np.uint8)
    check(lib().hkbcn_decompress_slices(fmt, _ptr(blocks), _ptr(dst), width, height, slices, mips, flags))
    return dst


def decompress_batch(jobs):
    """jobs: list of (fmt, blocks ndarray, width, height, flags) -> list of decoded ndarrays (tightly packed)"""
    arr = (Job * len(jobs))()
    keep, outs = [], []
    for i, (fmt, blocks, w, h, flags) in enumerate(jobs):
        blocks = np.ascontiguousarray(blocks)
        dst = np.zeros((w // 4) * 4 * (h // 4) * 4 * lib().hkbcn_dst_texel_bytes(fmt, flags), np.uint8)
        keep.append(blocks)
        outs.append(dst)
        arr[i] = Job(fmt, w, h, flags, blocks.ctypes.data, dst.ctypes.data, 0)
    check(lib().hkbcn_decompress_batch(arr, len(jobs)))
    return outs


def compress_batch(jobs):
    """jobs: list of (fmt, src ndarray, width, height, flags) -> list of ndarray outputs."""
    arr = (Job * len(jobs))()
    keep, outs = [], []
    for i, (fmt, src, w, h, flags) in enumerate(jobs):
        src = np.ascontiguousarray(src)
        dst = np.zeros((w // 4) * (h // 4) * block_bytes(fmt), np.uint8)
        keep.append(src)
        outs.append(dst)
        arr[i] = Job(fmt, w, h, flags, src.ctypes.data, dst.ctypes.data, 0)
    check(lib().hkbcn_compress_batch(arr, len(jobs)))
    return outs


def plan_batch(jobs, n_gpus):
    """jobs: list of (fmt, width, height) -> list of GPU indices (host-only planning step of compress_batch)."""
    arr = (Job * len(jobs))()
    for i, (fmt, w, h) in enumerate(jobs):
        arr[i] = Job(fmt, w, h, 0, None, None, 0)
    out = (C.c_int * len(jobs))()
    check(lib().hkbcn_plan_batch(arr, len(jobs), n_gpus, out))
    return list(out)


def mip_taps(in_size, out_size, filter=5, edge=3):
    """(idx[out, T], weight[out, T]) of one axis -- host only"""
    T = C.c_int(0)
    cap = out_size * 64
    idx = np.zeros(cap, np.int32)
    w = np.zeros(cap, np.float32)
    check(lib().hkbcn_mip_taps(in_size, out_size, filter, edge, _ptr(idx), _ptr(w), cap, C.byref(T)))
    return idx[: out_size * T.value].reshape(out_size, T.value), w[: out_size * T.value].reshape(out_size, T.value)


def import_chain(fmt, level0, width, height, mips, cfg=None, flags=0):
    """the import cascade (Image.cpp:1755-1803): level 0 in, compressed chain out; mips generated on the device"""
    level0 = np.ascontiguousarray(level0)
    dst = np.zeros(lib().hkbcn_chain_bytes(fmt, width, height, 1, mips), np.uint8)
    check(lib().hkbcn_import_chain(fmt, _ptr(level0), _ptr(dst), width, height, mips, C.byref(cfg) if cfg else None, flags))
    return dst


def generate_mips(fmt, level0, width, height, mips, cfg=None, flags=0):
    """uncompressed chain (levels back to back, level 0 first) as uint8 bytes"""
    level0 = np.ascontiguousarray(level0)
    dst = np.zeros(lib().hkbcn_chain_src_bytes(fmt, width, height, mips, flags), np.uint8)
    check(lib().hkbcn_generate_mips(fmt, _ptr(level0), _ptr(dst), width, height, mips, C.byref(cfg) if cfg else None, flags))
    return dst


def texture_header(type_, width, height, depth, mips, fmt, storage_flags, payload_bytes):
    """the 30 header bytes of a .texture file (host only)"""
    d = TextureDesc(type_, width, height, depth, mips, fmt, storage_flags)
    out = (C.c_uint8 * 30)()
    check(lib().hkbcn_texture_header(C.byref(d), payload_bytes, out))
    return bytes(out)


def write_texture(path, type_, width, height, depth, mips, fmt, storage_flags, payload):
    payload = np.ascontiguousarray(payload)
    d = TextureDesc(type_, width, height, depth, mips, fmt, storage_flags)
    check(lib().hkbcn_write_texture(os.fsencode(path), C.byref(d), _ptr(payload), C.c_size_t(payload.nbytes)))


def read_texture(path):
    """-> (TextureDesc, payload ndarray)"""
    d = TextureDesc()
    n = C.c_size_t(0)
    check(lib().hkbcn_read_texture(os.fsencode(path), C.byref(d), None, C.c_size_t(0), C.byref(n)))
    payload = np.zeros(n.value, np.uint8)
    check(lib().hkbcn_read_texture(os.fsencode(path), C.byref(d), _ptr(payload), C.c_size_t(n.value), C.byref(n)))
    return d, payload


def import_batch(jobs):
    """jobs: list of dicts {fmt, level0, width, height, mips(bool), cfg(MipConfig or None), flags, out_path or None}
    -> list of compressed chains (ndarray)"""
    arr = (ImportJob * len(jobs))()
    keep, outs = [], []
    for i, j in enumerate(jobs):
        src = np.ascontiguousarray(j["level0"])
        mips = lib().hkbcn_calc_num_mips(j["width"], j["height"]) if j.get("mips") else 1
        dst = np.zeros(lib().hkbcn_chain_bytes(j["fmt"], j["width"], j["height"], 1, mips), np.uint8)
        keep.append(src)
        outs.append(dst)
        cfg = j.get("cfg") or MipConfig()
        path = os.fsencode(j["out_path"]) if j.get("out_path") else None
        arr[i] = ImportJob(j["fmt"], j["width"], j["height"], 1 if j.get("mips") else 0, cfg, j.get("flags", 0),
                           src.ctypes.data, path, dst.ctypes.data, 0)
    check(lib().hkbcn_import_batch(arr, C.c_size_t(len(jobs))))
    for i in range(len(jobs)):
        if arr[i].status:
            raise HkbcnError(f"import job {i} failed with status {arr[i].status}")
    return outs


def pack_normals(variant, normals):
    """PackNormals*_Compatible over Float3 normals (n, 3) -> uint8 bytes"""
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    out = np.zeros(n.shape[0] * lib().hkbcn_pack_texel_bytes(variant), np.uint8)
    check(lib().hkbcn_pack_normals(variant, _ptr(n), _ptr(out), C.c_size_t(n.shape[0])))
    return out


def unpack_normals(variant, packed, count):
    p = np.ascontiguousarray(packed, np.uint8)
    out = np.zeros((count, 3), np.float32)
    check(lib().hkbcn_unpack_normals(variant, _ptr(p), _ptr(out), C.c_size_t(count)))
    return out


def normalize_vectors(v):
    out = np.array(v, np.float32).reshape(-1, 3).copy()
    check(lib().hkbcn_normalize_vectors(_ptr(out), C.c_size_t(out.shape[0])))
    return out


def bake_roughness(normals, rough_chain, width, height, mips, bake, filt=0, edge=0):
    """roughness chain (uint8, levels back to back) updated in a copy and returned"""
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    r = np.array(rough_chain, np.uint8).ravel().copy()
    check(lib().hkbcn_bake_roughness(_ptr(n), _ptr(r), width, height, mips, bake, filt, edge))
    return r


def envprobe_prefilter(cube, src_size, max_lod):
    """cube: (6, n, n, 4) float32 -> list of levels [(6, s, s, 4)] with s = 2^max_lod >> level"""
    c = np.ascontiguousarray(cube, np.float32)
    out = np.zeros(lib().hkbcn_envprobe_texels(max_lod) * 4, np.float32)
    check(lib().hkbcn_envprobe_prefilter(_ptr(c), src_size, max_lod, _ptr(out)))
    levels, off = [], 0
    for lod in range(max_lod + 1):
        s = (1 << max_lod) >> lod
        levels.append(out[off:off + 6 * s * s * 4].reshape(6, s, s, 4))
        off += 6 * s * s * 4
    return levels


def irradiance(cube, src_size, out_size):
    c = np.ascontiguousarray(cube, np.float32)
    out = np.zeros((6, out_size, out_size, 4), np.float32)
    check(lib().hkbcn_irradiance(_ptr(c), src_size, out_size, _ptr(out)))
    return out
