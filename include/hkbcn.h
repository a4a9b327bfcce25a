/*
 * hkbcn.h -- C ABI of the B200-native BCn block-compression path.
 *
 * Drop-in boundary for Hork/Image's TextureBlockCompression entry points
 * (reference: Hork/Image/ImageEncoders.h:37-87).  Plain pointers and sizes only;
 * every function returns 0 on success or a negative hkbcn_status.  There is NO
 * CPU fallback: when no CUDA device / kernel image is usable the call fails with
 * HKBCN_ERR_NO_DEVICE / HKBCN_ERR_CUDA and hkbcn_last_error() says why.
 *
 * Layout conventions (identical to the reference, ImageEncoders.cpp:215-557):
 *   source surface  : tightly packed row-major texels, rowStride = Width * bpp
 *   compressed      : tightly packed blocks, row-major block order
 *   Width/4, Height/4 truncate (ImageEncoders.cpp:224-226): the single-surface calls (hkbcn_compress, _compress_dev,
 *   _compress_batch) accept any Width / Height, keep rowStride = Width * bpp and skip the Width % 4 texels at the end of
 *   each row and the Height % 4 rows at the bottom, exactly as the reference does.  Chains and sliced storage
 *   (ImageStorage, Image.cpp:412-431) need multiples of 4 -- powers of two when mipmapped -- and return HKBCN_ERR_BAD_ARG
 *   otherwise.
 *   Device pointers (*_dev): compressed-block pointers (d_dst of the encoders, d_src of the decoders) must be aligned to
 *   the block size (8 or 16 bytes); texel pointers may sit at any address of their component type.  Violations return
 *   HKBCN_ERR_BAD_ARG instead of faulting.  Work on one stream runs in order; different streams may run concurrently.
 */
#ifndef HKBCN_H
#define HKBCN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#  define HKBCN_API __declspec(dllexport)
#else
#  define HKBCN_API __attribute__((visibility("default")))
#endif

/* Format ids == Hk::TEXTURE_FORMAT enum values (Hork/Image/Image.h:122-148). */
typedef enum hkbcn_format {
    HKBCN_BC1_UNORM      = 54, /* TEXTURE_FORMAT_BC1_UNORM      : RGBA8 in (alpha ignored), 8 B/block  */
    HKBCN_BC1_UNORM_SRGB = 55,
    HKBCN_BC2_UNORM      = 56, /* RGBA8 in, 16 B/block */
    HKBCN_BC2_UNORM_SRGB = 57,
    HKBCN_BC3_UNORM      = 58, /* RGBA8 in, 16 B/block */
    HKBCN_BC3_UNORM_SRGB = 59,
    HKBCN_BC4_UNORM      = 60, /* R8 in, 8 B/block */
    HKBCN_BC4_SNORM      = 61,
    HKBCN_BC5_UNORM      = 62, /* RG8 in, 16 B/block */
    HKBCN_BC5_SNORM      = 63,
    HKBCN_BC6H_UFLOAT    = 64, /* RGBA32F in (alpha ignored), 16 B/block */
    HKBCN_BC6H_SFLOAT    = 65,
    HKBCN_BC7_UNORM      = 66, /* RGBA8 in, 16 B/block */
    HKBCN_BC7_UNORM_SRGB = 67
} hkbcn_format;

typedef enum hkbcn_status {
    HKBCN_OK              =  0,
    HKBCN_ERR_BAD_FORMAT  = -1,
    HKBCN_ERR_BAD_ARG     = -2,
    HKBCN_ERR_NO_DEVICE   = -3, /* no CUDA device, or device is not sm_100 */
    HKBCN_ERR_CUDA        = -4, /* a CUDA runtime call failed; see hkbcn_last_error() */
    HKBCN_ERR_NOT_INIT    = -5,
    HKBCN_ERR_UNSUPPORTED = -6,
    HKBCN_SKIPPED         =  1  /* hkbcn_import_job.status only: the output file already existed (HKBCN_FLAG_SKIP_EXISTING) */
} hkbcn_status;

/* flags for hkbcn_compress* */
enum {
    HKBCN_FLAG_NONE        = 0,
    HKBCN_FLAG_SRC_F16     = 1u << 0, /* BC6H: source texels are RGBA16F (Encode_BC6h_f16, ImageEncoders.cpp:187) */
    HKBCN_FLAG_FAST        = 1u << 1, /* BC4/BC5/BC3-alpha: non-hq encoder (bMaxQuality=false, ImageEncoders.cpp:161-185) */
    /* bits 4..8: quality level of Encode_BC1/2/3(Level) and Encode_BC7(Level) (ImageEncoders.cpp:138-169, 207-212), stored
     * as level + 1; 0 = the preset the Compress* functions bake in (BC1/2/3 level 5, BC7 level 4).  Use HKBCN_FLAG_LEVEL(n).
     * Built: BC1/2/3 levels 0..18 (every level Hork's BC1/2/3_ENCODE_MAX_LEVEL admits, rgbcx.cpp:1692-1811) and BC7 levels
     * 0..4.  BC1/2/3 levels 6..18 and the two 3-colour options need rgbcx's total-ordering tables, which are data of the
     * engine's ThirdParty/bc7enc_rdo and are taken from the engine tree when the library is built (csrc/Makefile: RGBCX_DIR);
     * a library built without them answers HKBCN_ERR_UNSUPPORTED for those. */
    HKBCN_FLAG_LEVEL_MASK  = 31u << 4,
    /* Encode_BC1(.., b3ColorMode, bTransparentPixelsForBlack) (ImageEncoders.cpp:138-143): BC1 only, read at levels >= 5 as in
     * rgbcx::encode_bc1 (levels 0..4 ignore them; Encode_BC2 / Encode_BC3 never pass them).  Blocks may then come out in
     * 3-colour mode (colour0 <= colour1), with selector 3 = transparent black when the second option is on. */
    HKBCN_FLAG_BC1_3COLOR  = 1u << 9,
    HKBCN_FLAG_BC1_TRANSPARENT_BLACK = 1u << 10,
    HKBCN_FLAG_SRC_DEVICE  = 1u << 17, /* hkbcn_import_batch: src_level0 is DEVICE memory on the GPU hkbcn_plan_import_batch assigns the
                                          job to (sources produced on the device never cross PCIe); the default is host memory */
    HKBCN_FLAG_SRC_CHAIN   = 1u << 18, /* hkbcn_import_batch: src_level0 holds the whole uncompressed chain (hkbcn_chain_src_bytes() bytes, e.g. built
                                          by the caller's own cascade): every level is uploaded and encoded, no mips are generated */
    HKBCN_FLAG_SKIP_EXISTING = 1u << 16 /* hkbcn_import_batch: leave a job alone (status HKBCN_SKIPPED) when its out_path already holds a
                                           complete .texture of the job's format, size and mip count -- re-running an interrupted
                                           batch only does what is missing (files are written to a temporary name and renamed) */
};
#define HKBCN_FLAG_LEVEL(n) ((((uint32_t)(n)) + 1u) << 4)
/* flags for hkbcn_decompress* */
enum {
    HKBCN_FLAG_DST_F32     = 1u << 8  /* BC6H: decode to RGB32F (Decode_BC6h_f32) instead of RGB16F (Decode_BC6h_f16) */
};

/* ---- lifetime ---------------------------------------------------------------
 * Replaces the C++ static constructor `CompressionParams compressionParams`
 * (ImageEncoders.cpp:52-68: rgbcx::init + bc7enc_compress_block_init + params).
 * device_mask: bit i set = use CUDA device i; 0 = all visible devices.
 * Called lazily by every other entry point with device_mask = 0. */
HKBCN_API int hkbcn_init(uint32_t device_mask);
HKBCN_API void hkbcn_shutdown(void);
HKBCN_API int hkbcn_device_count(void);          /* devices selected by hkbcn_init */
HKBCN_API const char* hkbcn_last_error(void);
/* What this build of the library carries of the reference's DATA tables (they are read out of the engine tree at build time,
 * csrc/Makefile: RGBCX_DIR, STB_DIR; a build without the tree works, with the stated differences):
 *   HKBCN_BUILD_BC1_ORDERINGS  rgbcx's total-ordering tables: BC1/2/3 levels 6..18 and the 3-colour options (else HKBCN_ERR_UNSUPPORTED)
 *   HKBCN_BUILD_STBIR_SRGB     stbir's float -> sRGB8 segments: mip chains of *_SRGB formats bit-identical to stbir (else within 1 LSB) */
enum { HKBCN_BUILD_BC1_ORDERINGS = 1, HKBCN_BUILD_STBIR_SRGB = 2 };
HKBCN_API int hkbcn_build_has(int feature);
/* entry v of stbir's sRGB -> linear decode table (stbir__srgb_uchar_to_linear_float), as hkbcn_init computes it */
HKBCN_API float hkbcn_srgb_decode_entry(int v);
    /* thread-local message of the last failure */
HKBCN_API const char* hkbcn_version(void);

/* ---- format queries (TexFormat[] table, Image.cpp:49-119) -------------------- */
HKBCN_API int hkbcn_block_bytes(int fmt);        /* 8 or 16; <0 on bad format */
HKBCN_API int hkbcn_src_texel_bytes(int fmt, uint32_t flags); /* 1, 2, 4, 8 or 16 */
HKBCN_API int hkbcn_dst_texel_bytes(int fmt, uint32_t flags); /* decoded texel size: 1,2,4,6,12 */
/* CalcNumMips for a compressed format (Image.cpp:188-216): log2(max(w,h)/4)+1 */
HKBCN_API uint32_t hkbcn_calc_num_mips(uint32_t width, uint32_t height);
/* bytes of a full ImageStorage-layout chain (Image.cpp:363-467): all slices of mip 0, then mip 1, ... */
HKBCN_API size_t hkbcn_chain_bytes(int fmt, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips);

/* ---- whole-surface compress, host pointers ----------------------------------
 * Replaces CompressBC1/2/3/4/5/7(pSrc,pDest,Width,Height) (ImageEncoders.cpp:215,251,287,323,359,524)
 * and CompressBC6h(pSrc,pDest,Width,Height,bSigned) (:395; fmt selects signedness).
 * Quality presets are the reference's hard-coded ones: BC1/2/3 level 5, BC3/4/5 hq,
 * BC7 uber level 4 perceptual.  Synchronous; dst holds (w/4)*(h/4)*block_bytes on return.
 * Every block is encoded (the reference's CompressBC6h drops numBlocks%16 tail blocks,
 * ImageEncoders.cpp:426-427; this does not). */
HKBCN_API int hkbcn_compress(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t flags);

/* Whole-surface decode, host pointers.  The reference only has per-block Decode_BC*
 * (ImageEncoders.cpp:77-136); this is that call applied to every block of the surface,
 * written at dst_row_stride bytes per texel row.  BC6H output is RGB (3 components). */
HKBCN_API int hkbcn_decompress(int fmt, const void* src, void* dst, uint32_t width, uint32_t height,
                               size_t dst_row_stride, uint32_t flags);

/* Decode of whole storages: every slice (cube face, array layer) of every mip level of an ImageStorage (Image.cpp:498-574)
 * with one launch per LEVEL -- the slices of a level are contiguous in the storage and decode as one tall surface -- instead
 * of one call per block (the reference's Decode_BC* loop, ImageEncoders.cpp:77-136) or one launch per surface.  src: the
 * compressed storage (hkbcn_chain_bytes()); dst: the decoded levels back to back, level i = slices x max(w>>i,4) x max(h>>i,4)
 * texels tightly packed (hkbcn_decoded_chain_bytes()).  A batch of same-sized textures is the slices form with num_mips = 1. */
HKBCN_API size_t hkbcn_decoded_chain_bytes(int fmt, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips, uint32_t flags);
HKBCN_API int hkbcn_decompress_slices_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                          uint32_t slices, uint32_t num_mips, uint32_t flags, void* stream);
HKBCN_API int hkbcn_decompress_slices(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t slices,
                                      uint32_t num_mips, uint32_t flags);
HKBCN_API int hkbcn_decompress_chain_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                         uint32_t num_mips, uint32_t flags, void* stream);
HKBCN_API int hkbcn_decompress_chain(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t num_mips, uint32_t flags);

/* ---- per-block entry points (ImageEncoders.h:48-64), host pointers ----------
 * n_blocks contiguous 4x4 blocks (src: n * 16 texels, dst: n * block_bytes). The
 * reference signatures are the n_blocks == 1 case. */
HKBCN_API int hkbcn_encode_blocks(int fmt, const void* src, void* dst, size_t n_blocks, uint32_t flags);
/* dst: n_blocks decoded blocks, each 4 rows of 4 texels at dst_row_stride bytes/row
 * inside its own 4*dst_row_stride bytes (Decode_BC*(pSrc,pDest,RowStride)). */
HKBCN_API int hkbcn_decode_blocks(int fmt, const void* src, void* dst, size_t n_blocks, size_t dst_row_stride, uint32_t flags);

/* ---- device-pointer variants (no PCIe; `stream` is a cudaStream_t or NULL) ---
 * Asynchronous with respect to the host: work is enqueued on `stream` of `device`. */
HKBCN_API int hkbcn_compress_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                 uint32_t flags, void* stream);
HKBCN_API int hkbcn_decompress_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                   size_t dst_row_stride, uint32_t flags, void* stream);

/* Compress a whole pre-built mip chain resident on the device in ONE launch.
 * d_src: levels back to back (level i is max(w>>i,4) x max(h>>i,4) texels, tightly packed);
 * d_dst: ImageStorage layout for one slice (Image.cpp:498-574).  */
HKBCN_API int hkbcn_compress_chain_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                       uint32_t num_mips, uint32_t flags, void* stream);
/* Host-pointer form: src = the same back-to-back level layout in host memory. */
HKBCN_API int hkbcn_compress_chain(int fmt, const void* src, void* dst, uint32_t width, uint32_t height,
                                   uint32_t num_mips, uint32_t flags);

/* Sliced storage (TEXTURE_CUBE faces, array layers) with its mip levels in ONE launch.  Replaces LoadSkyboxImages' per-face
 * loop `CompressBC6h(face.data, subres.data, w, h, false)` (Image.cpp:2033-2035) and the per-slice / per-mip loops of
 * CreateImage (Image.cpp:1753-1803).  src and dst are both in ImageStorage order (Image.cpp:498-574): all `slices` surfaces
 * of level 0, then all of level 1, ...; level i is max(w>>i,4) x max(h>>i,4); width and height multiples of 4. */
HKBCN_API size_t hkbcn_slices_src_bytes(int fmt, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips, uint32_t flags);
HKBCN_API int hkbcn_compress_slices_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                        uint32_t slices, uint32_t num_mips, uint32_t flags, void* stream);
HKBCN_API int hkbcn_compress_slices(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t slices,
                                    uint32_t num_mips, uint32_t flags);

/* One large surface (or pre-built chain) over several GPUs: GPU d encodes band d -- whole block rows -- of every level and
 * copies it to its final offset in dst; no exchange between GPUs.  n_gpus: 0 = all GPUs selected by
 * hkbcn_init, n > 0 = the first n of them, n < 0 = -n bands dealt round-robin over them.  hkbcn_compress and
 * hkbcn_compress_chain take this path by themselves for surfaces of 4 Mpix and more when several GPUs are selected.
 * hkbcn_plan_bands (host only): block-row range [row0[i], row1[i]) of level i that `device` of `n_gpus` encodes. */
HKBCN_API int hkbcn_compress_chain_multi(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t num_mips,
                                         uint32_t flags, int n_gpus);
HKBCN_API int hkbcn_plan_bands(uint32_t height, uint32_t num_mips, int n_gpus, int device, uint32_t* row0, uint32_t* row1);

/* ---- device-side mip cascade ------------------------------------------------
 * The importers build every mip level on the CPU from the previous one with stbir_resize and compress it
 * (Image.cpp:1755-1803 for BC1-5/BC7, :1852-1899 for BC6H; ImageMipmapConfig, Image.h:361-371).  These entry points do the
 * same cascade on the device, so only level 0 crosses PCIe.  Against stbir: bit-identical for uint8 formats, linear and sRGB (sRGB
 * within 1 LSB in a build without stbir's segment table, see hkbcn_build_has) and for RGBA32F chains (the same bit patterns). */
typedef struct hkbcn_mip_config {
    int edge_mode;            /* IMAGE_RESAMPLE_EDGE_MODE (Image.h:330): 1 clamp, 2 reflect, 3 wrap, 4 zero; 0 = default (wrap) */
    int filter;               /* IMAGE_RESAMPLE_FILTER (Image.h:339): 1 box, 2 triangle, 3 cubic b-spline, 4 catmull-rom, 5 mitchell; 0 = default (mitchell) */
    int no_alpha;             /* IMAGE_STORAGE_NO_ALPHA: the fourth channel is colour, no alpha weighting */
    int alpha_premultiplied;  /* IMAGE_STORAGE_ALPHA_PREMULTIPLIED */
} hkbcn_mip_config;
/* The resampling taps of one axis (host only, no device needed): output texel i = sum over k < *taps_per_output of
 * weight[i * T + k] * input[idx[i * T + k]], in ascending input order -- stbir's downsampling contributors after
 * normalisation and edge mapping (stb_image_resize.h:1024-1200).  capacity = entries available in idx / weight. */
HKBCN_API int hkbcn_mip_taps(int in_size, int out_size, int filter, int edge_mode, int* idx, float* weight, int capacity,
                             int* taps_per_output);
/* bytes of the uncompressed chain (levels back to back) for `fmt`'s source texel type */
HKBCN_API size_t hkbcn_chain_src_bytes(int fmt, uint32_t width, uint32_t height, uint32_t num_mips, uint32_t flags);
/* d_chain: hkbcn_chain_src_bytes() bytes on `device` with level 0 filled in; levels 1.. are written after it.
 * The source texel type and the sRGB / alpha handling follow `fmt` as in the importers (cfg may be NULL: defaults). */
HKBCN_API int hkbcn_generate_mips_dev(int device, int fmt, void* d_chain, uint32_t width, uint32_t height, uint32_t num_mips,
                                      const hkbcn_mip_config* cfg, uint32_t flags, void* stream);
/* host pointers: level 0 in, the uncompressed chain (level 0 included) out, hkbcn_chain_src_bytes() bytes */
HKBCN_API int hkbcn_generate_mips(int fmt, const void* src_level0, void* dst_chain, uint32_t width, uint32_t height,
                                  uint32_t num_mips, const hkbcn_mip_config* cfg, uint32_t flags);
/* generate the mips in d_chain, then compress the whole chain into d_dst (ImageStorage layout): the import cascade */
HKBCN_API int hkbcn_import_chain_dev(int device, int fmt, void* d_chain, void* d_dst, uint32_t width, uint32_t height,
                                     uint32_t num_mips, const hkbcn_mip_config* cfg, uint32_t flags, void* stream);
/* host pointers: level 0 in (width x height source texels), full compressed chain out */
HKBCN_API int hkbcn_import_chain(int fmt, const void* src_level0, void* dst, uint32_t width, uint32_t height, uint32_t num_mips,
                                 const hkbcn_mip_config* cfg, uint32_t flags);

/* ---- CreateImage from a raw image: format selection + channel pass + cascade + compress -------------------------------
 * Replaces the compressed branch of Hk::CreateImage(RawImage const&, ImageMipmapConfig const*, IMAGE_STORAGE_FLAGS,
 * IMAGE_IMPORT_FLAGS) (Image.cpp:866-1289).  raw_format: Hk::RAW_IMAGE_FORMAT values (RawImage.h:38-55); storage_flags:
 * IMAGE_STORAGE_FLAGS (Image.h:221-226); import_flags: IMAGE_IMPORT_FLAGS (Image.h:546-553). */
enum hkbcn_raw_format {
    HKBCN_RAW_R8 = 1, HKBCN_RAW_R8_ALPHA = 2, HKBCN_RAW_RGB8 = 3, HKBCN_RAW_BGR8 = 4, HKBCN_RAW_RGBA8 = 5, HKBCN_RAW_BGRA8 = 6,
    HKBCN_RAW_R32F = 7, HKBCN_RAW_R32_ALPHA_F = 8, HKBCN_RAW_RGB32F = 9, HKBCN_RAW_BGR32F = 10, HKBCN_RAW_RGBA32F = 11, HKBCN_RAW_BGRA32F = 12
};
/* storage_flags: HKBCN_STORAGE_NO_ALPHA / HKBCN_STORAGE_ALPHA_PREMULTIPLIED (below, with hkbcn_texture_desc) */
enum { HKBCN_IMPORT_STORE_HDRI_AS_HALF_FLOAT = 1, HKBCN_IMPORT_USE_COMPRESSION = 2, HKBCN_IMPORT_ALLOW_HDRI_COMPRESSION = 4,
       HKBCN_IMPORT_ASSUME_8BIT_RGB_IMAGES_ARE_SRGB = 8 };
typedef struct hkbcn_image_plan {
    int      format;          /* the block-compressed TEXTURE_FORMAT the reference picks (Image.cpp:878-1056); 0 = it stores the image uncompressed */
    uint32_t storage_flags;   /* storage_flags as the reference leaves them (RGB sources add NO_ALPHA) */
    int      src_channels;    /* channels of the raw image */
    int      add_alpha;       /* a fourth channel of 255 / 1.0 is appended before compression */
    int      swap_rb;         /* R and B are exchanged before compression (BGR / BGRA sources) */
    uint32_t required_width;  /* the size the reference resamples to first (Image.cpp:1064-1068): closest power of two, at least 4, */
    uint32_t required_height; /*   when mipmapped; the next multiple of 4 otherwise */
} hkbcn_image_plan;
/* ---- ResampleRawImage (Image.cpp:2188-2243; the -resample pre-pass of the importers, Tools/TextureImporter/Main.cpp:145-200, and
 * CreateImage's resample to a power of two, Image.cpp:1064-1096): stbir_resize of a raw image to scaled_width x scaled_height,
 * up or down, each axis with its own edge mode and filter, on the device.  Bit-identical to stbir for linear data (uint8 and
 * float, with and without alpha weighting) and for sRGB uint8 (as the mip cascade).  Float images in the sRGB colour
 * space are refused (no caller in the reference asks for them). */
enum { HKBCN_RESAMPLE_HAS_ALPHA = 1, HKBCN_RESAMPLE_ALPHA_PREMULTIPLIED = 2, HKBCN_RESAMPLE_COLORSPACE_SRGB = 4 }; /* RAW_IMAGE_RESAMPLE_FLAG, Image.h:518-524 */
typedef struct hkbcn_resample_params {
    uint32_t flags;
    int      edge_h, edge_v;       /* IMAGE_RESAMPLE_EDGE_MODE, 0 = wrap (RawImageResampleParams' default) */
    int      filter_h, filter_v;   /* IMAGE_RESAMPLE_FILTER, 0 = mitchell */
    uint32_t scaled_width, scaled_height;
} hkbcn_resample_params;
/* host pointers, device 0, synchronous; dst: scaled_width * scaled_height texels of the source's raw format */
HKBCN_API int hkbcn_resample_raw_image(int raw_format, const void* src, uint32_t width, uint32_t height,
                                       const hkbcn_resample_params* params, void* dst);
/* device pointers on `device`, asynchronous on `stream` (tap tables are built and uploaded on first use of a size pair) */
HKBCN_API int hkbcn_resample_raw_image_dev(int device, int raw_format, const void* d_src, uint32_t width, uint32_t height,
                                           const hkbcn_resample_params* params, void* d_dst, void* stream);

/* host only.  mipmapped: whether a mip config will be passed.  Unknown raw formats: HKBCN_ERR_BAD_FORMAT. */
HKBCN_API int hkbcn_plan_image(int raw_format, uint32_t width, uint32_t height, int mipmapped, uint32_t storage_flags,
                               uint32_t import_flags, hkbcn_image_plan* plan);
/* The whole call on device 0: raw image in (host, tightly packed), compressed ImageStorage payload out (host).  cfg != NULL:
 * full mip chain (CalcNumMips levels) built on the device with cfg's edge mode and filter; alpha weighting follows the
 * storage flags as in ImageStorage::GenerateMipmaps (Image.cpp:767-774).  desc receives type / size / mips / format / flags
 * for hkbcn_write_texture.  capacity: bytes available at dst (hkbcn_chain_bytes(plan.format, w, h, 1, mips) are needed).
 * An image whose size is not plan.required_width x required_height is first resampled to it on the device, with the mip
 * config's edge mode and filter (wrap / mitchell without one) and the alpha / sRGB handling of Image.cpp:1074-1093, as the
 * reference does; desc then carries the resampled size and capacity must cover it.
 * Returns HKBCN_ERR_UNSUPPORTED when the reference would not compress (plan.format == 0). */
struct hkbcn_texture_desc;
HKBCN_API int hkbcn_create_image(int raw_format, const void* src, uint32_t width, uint32_t height, const hkbcn_mip_config* cfg,
                                 uint32_t storage_flags, uint32_t import_flags, struct hkbcn_texture_desc* desc, void* dst,
                                 size_t capacity);
/* Replaces Hk::CreateRoughnessMap(pRoughnessMap, Width, Height, bUseCompression = true, bMipmapped, edge mode, filter)
 * (Image.cpp:2399-2456): R8 in, BC4 chain out.  Sizes the reference refuses (not a power of two / multiple of 4) return
 * HKBCN_ERR_BAD_ARG with its message. */
HKBCN_API int hkbcn_create_roughness_map(const uint8_t* roughness, uint32_t width, uint32_t height, int mipmapped, int edge_mode,
                                         int filter, struct hkbcn_texture_desc* desc, void* dst, size_t capacity);

/* ---- batch API: independent surfaces sharded over the selected GPUs ---------
 * No collective: each job is sent whole to one GPU (greedy by pixel-cost), results land
 * directly in the caller's dst ("host gather").  Replaces one-process-per-texture use of
 * the importers (Tools/TextureImporter/Main.cpp:236). */
typedef struct hkbcn_job {
    int         fmt;
    uint32_t    width, height;
    uint32_t    flags;
    const void* src;   /* host */
    void*       dst;   /* host */
    int         status;/* out */
} hkbcn_job;
HKBCN_API int hkbcn_compress_batch(hkbcn_job* jobs, size_t n_jobs);
/* the same for decode: job.src = blocks, job.dst = texels tightly packed ((width / 4) * 4 texels per row), job.flags = HKBCN_FLAG_DST_F32 */
HKBCN_API int hkbcn_decompress_batch(hkbcn_job* jobs, size_t n_jobs);
/* The planning step of hkbcn_compress_batch on its own (host only, no device needed): device_of_job[i] = index (0..n_gpus-1)
 * of the GPU job i would be sent to.  Greedy longest-processing-time by pixel count x per-format cost weight; deterministic. */
HKBCN_API int hkbcn_plan_batch(const hkbcn_job* jobs, size_t n_jobs, int n_gpus, int* device_of_job);

/* ---- .texture files and the importer batch -----------
 * File layout written by the importers: AssetUtils::CreateTexture (Hork/Resources/Resource_Texture.cpp:104-109) =
 * MakeResourceMagic(RESOURCE_TEXTURE = 4, Version = 1) (ResourceBase.h:72-75: 'H' 'k' type version) followed by
 * ImageStorage::Write (Hork/Image/Image.cpp:833-844): u8 type, u32 width, u32 height, u32 depth / slice count,
 * u32 mip count, u8 format, u32 flags, u32 payload size, payload.  All integers little-endian
 * (Core/BinaryStream.h:408-412).  30 header bytes.  These three entry points are host-only (no device needed). */
#define HKBCN_TEXTURE_HEADER_BYTES 30
#define HKBCN_TEXTURE_2D 2            /* TEXTURE_TYPE (Image.h:40-51) */
#define HKBCN_TEXTURE_2D_ARRAY 3
#define HKBCN_TEXTURE_CUBE 5
#define HKBCN_STORAGE_NO_ALPHA 1      /* IMAGE_STORAGE_FLAGS (Image.h:221-226) */
#define HKBCN_STORAGE_ALPHA_PREMULTIPLIED 2
typedef struct hkbcn_texture_desc {
    int      type;          /* TEXTURE_TYPE */
    uint32_t width, height;
    uint32_t depth;         /* depth or slice count: 1 for 2D, 6 for a cube */
    uint32_t num_mips;
    int      fmt;           /* TEXTURE_FORMAT = enum hkbcn_format */
    uint32_t storage_flags; /* IMAGE_STORAGE_FLAGS */
} hkbcn_texture_desc;
HKBCN_API int hkbcn_texture_header(const hkbcn_texture_desc* desc, uint32_t payload_bytes, void* dst30);
HKBCN_API int hkbcn_write_texture(const char* path, const hkbcn_texture_desc* desc, const void* payload, size_t payload_bytes);
/* payload may be NULL (header only); *payload_bytes receives the stored size, at most `capacity` bytes are copied */
HKBCN_API int hkbcn_read_texture(const char* path, hkbcn_texture_desc* desc, void* payload, size_t capacity, size_t* payload_bytes);

/* One texture of an import batch: what one run of Tools/TextureImporter does (Main.cpp:58-253 -> CreateImage, Image.cpp:1700-1803),
 * from level 0 already in memory to the .texture file.  File decoding (stb_image / tinyexr / webp) stays with the caller.
 * generate_mips: 0 = level 0 only (no -mip_* flag given), 1 = full chain with `mip` (zeroes = the importer's defaults). */
typedef struct hkbcn_import_job {
    int              fmt;
    uint32_t         width, height;   /* multiples of 4; powers of two when generate_mips (Image.cpp:415-416) */
    int              generate_mips;
    hkbcn_mip_config mip;             /* no_alpha / alpha_premultiplied are also stored as the file's IMAGE_STORAGE_FLAGS */
    uint32_t         flags;           /* HKBCN_FLAG_* for the source texels */
    const void*      src_level0;      /* host */
    const char*      out_path;        /* .texture file to write; NULL = keep the chain in `dst` only */
    void*            dst;             /* optional host buffer for the compressed chain (hkbcn_chain_bytes), may be NULL */
    int              status;          /* out */
    uint32_t         slices;          /* 0 or 1: a 2D texture.  6: a cube map (TEXTURE_CUBE: CubemapImporter, Tools/CubemapImporter/Main.cpp:58-145 ->
                                         LoadSkyboxImages, Image.cpp:1918-2040), any other count: a 2D array.  src_level0 then holds the slices'
                                         level 0 back to back, each slice gets its own cascade when generate_mips is set (ImageStorage::
                                         GenerateMipmaps slice by slice, Image.cpp:806-819), and the payload is in ImageStorage order: all slices
                                         of level 0, then of level 1, ... */
} hkbcn_import_job;
/* Jobs are sharded over the selected GPUs as in hkbcn_compress_batch (whole textures, no collective); each GPU's worker runs
 * upload -> mip cascade -> encode -> download, and the files are written by a separate thread behind it. */
HKBCN_API int hkbcn_import_batch(hkbcn_import_job* jobs, size_t n_jobs);
/* The planning step of hkbcn_import_batch on its own (host only; deterministic, the plan hkbcn_import_batch itself uses):
 * device_of_job[i] = index (0 .. n_gpus-1, in hkbcn_init's device order) of the GPU job i runs on.  n_gpus = 0: the selected GPUs. */
HKBCN_API int hkbcn_plan_import_batch(const hkbcn_import_job* jobs, size_t n_jobs, int n_gpus, int* device_of_job);
/* Per-GPU wall time (ms) of the workers of the last hkbcn_import_batch / hkbcn_compress_batch call of this process, in
 * hkbcn_init's device order; returns the number of devices (at most `capacity` values are written). */
HKBCN_API int hkbcn_batch_stats(double* busy_ms, int capacity);

/* ---- PBRImporter's ORMX texture as a batch job (Tools/PBRImporter/Main.cpp:284-735) ----------------------------------
 * One job = one ORMX texture: channel k of the RGBA8 image is channel `channel[k]` of `maps[k]` (occlusion, roughness,
 * metallic, displacement-or-emission; RGBA8 images of width x height in host memory), a missing map leaves the default
 * {255, 255, 0, 255} (CreateORMX, Main.cpp:86-190); the mips are those of an RGBA8 storage with IMAGE_STORAGE_NO_ALPHA
 * (:596-613); with `normals` (width x height normalised Float3, the output of hkbcn_normalize_vectors) and a bake, the
 * roughness channel of every level >= 1 is updated from the triangle-filtered normals (:616-680); the chain is compressed to
 * BC1 / BC3 / BC7 (the ORMX_BC1 / ORMX_BC3 / ORMX_BC7 presets, :683-735) and written as a .texture with NO_ALPHA.  Every
 * step runs on the GPU the job is sent to; only the maps go up and the compressed chain comes back.  Jobs are dealt over
 * the selected GPUs in order (job i -> GPU i mod n), one worker thread per GPU. */
typedef struct hkbcn_pbr_job {
    int              fmt;             /* HKBCN_BC1_UNORM, HKBCN_BC3_UNORM or HKBCN_BC7_UNORM */
    uint32_t         width, height;   /* multiples of 4; powers of two when generate_mips */
    const void*      maps[4];         /* host RGBA8 images or NULL; at least one */
    uint8_t          channel[4];      /* 0..3 */
    const float*     normals;         /* host, optional */
    int              roughness_bake;  /* 0, HKBCN_ROUGHNESS_BAKE_VMF or HKBCN_ROUGHNESS_BAKE_TOKSVIG; needs normals and generate_mips */
    int              generate_mips;
    hkbcn_mip_config mip;             /* edge mode and filter of the cascade (no_alpha is forced on) */
    const char*      out_path;        /* .texture file to write; NULL = keep the chain in `dst` only */
    void*            dst;             /* optional host buffer for the compressed chain (hkbcn_chain_bytes), may be NULL */
    int              status;          /* out */
} hkbcn_pbr_job;
HKBCN_API int hkbcn_import_pbr_batch(hkbcn_pbr_job* jobs, size_t n_jobs);

/* ---- normal maps and roughness anti-aliasing ----------
 * The element-wise work of the PBRImporter path between load and the BC5 / BC4 / BC1 / BC3 encoders:
 *   hkbcn_pack_normals    PackNormals*_Compatible (Hork/Image/ImageEncoders.cpp:693-796): Float3 normals -> RG8 / RGBA8
 *   hkbcn_unpack_normals  UnpackNormal*_Compatible (:798-878) applied to `count` texels
 *   hkbcn_normalize_vectors  NormalizeVectors (Hork/Image/Image.cpp:2251-2260): v = v * 2 - 1, then normalise
 *   hkbcn_bake_roughness  the vMF / Toksvig update of the roughness mips from the averaged normals
 *                         (CreateNormalAndRoughness, Image.cpp:2526-2580 == Tools/PBRImporter/Main.cpp:616-680)
 * Packers, unpackers, the normalisation and the bake are bit-exact with the reference's object code (tests/test_gpu_normals.py;
 * SURVEY.md 8f allows the bake +-1 LSB, what is measured and asserted is identity). */
enum hkbcn_normal_pack {
    HKBCN_NORMALS_RGBA_BC1 = 0,          /* NORMAL_MAP_PACK_RGBA_BC1_COMPATIBLE: x y z 255 */
    HKBCN_NORMALS_RG_BC5 = 1,            /* x y */
    HKBCN_NORMALS_SPHEREMAP_BC5 = 2,
    HKBCN_NORMALS_STEREOGRAPHIC_BC5 = 3,
    HKBCN_NORMALS_PARABOLOID_BC5 = 4,
    HKBCN_NORMALS_RGBA_BC3 = 5           /* 255 y z x */
};
#define HKBCN_ROUGHNESS_BAKE_VMF 1
#define HKBCN_ROUGHNESS_BAKE_TOKSVIG 2
HKBCN_API int hkbcn_pack_texel_bytes(int variant); /* 2 or 4; <0 on a bad variant */
/* host pointers; normals are tightly packed Float3 (12 bytes per texel) */
HKBCN_API int hkbcn_pack_normals(int variant, const float* normals_xyz, void* dst, size_t count);
HKBCN_API int hkbcn_unpack_normals(int variant, const void* src, float* normals_xyz, size_t count);
HKBCN_API int hkbcn_normalize_vectors(float* xyz, size_t count);
/* roughness_chain: the R8 mip chain (levels back to back, as hkbcn_generate_mips lays out BC4's source) of a width x height
 * roughness map; normals_xyz: the normalised normal map of the same size.  Level i >= 1 is updated from the normals averaged
 * down level by level with `filter` / `edge_mode` (hkbcn_mip_config values; 0 = the importers' triangle / wrap). */
HKBCN_API int hkbcn_bake_roughness(const float* normals_xyz, uint8_t* roughness_chain, uint32_t width, uint32_t height,
                                   uint32_t num_mips, int bake, int filter, int edge_mode);
/* device pointers, asynchronous on `stream`; d_scratch: hkbcn_bake_scratch_bytes() bytes for the averaged normals */
HKBCN_API int hkbcn_pack_normals_dev(int device, int variant, const float* d_normals_xyz, void* d_dst, size_t count, void* stream);
HKBCN_API int hkbcn_unpack_normals_dev(int device, int variant, const void* d_src, float* d_normals_xyz, size_t count, void* stream);
HKBCN_API int hkbcn_normalize_vectors_dev(int device, float* d_xyz, size_t count, void* stream);
HKBCN_API size_t hkbcn_bake_scratch_bytes(uint32_t width, uint32_t height);
HKBCN_API int hkbcn_bake_roughness_dev(int device, const float* d_normals_xyz, uint8_t* d_roughness_chain, void* d_scratch,
                                       uint32_t width, uint32_t height, uint32_t num_mips, int bake, int filter, int edge_mode,
                                       void* stream);

/* The two composed importer calls around them, on device 0 (host pointers, synchronous).  dst == NULL: only the descriptor(s)
 * are filled in (size query: hkbcn_chain_bytes(desc.fmt, desc.width, desc.height, 1, desc.num_mips) bytes are needed).
 * hkbcn_create_normal_map = Hk::CreateNormalMap(Float3 const*, Width, Height, Pack, bUseCompression = true, bMipmapped,
 * ResampleEdgeMode, ResampleFilter) (Image.cpp:2262-2352): normalised normals -> pack -> cascade of the packed texels ->
 * BC1 / BC5 / BC3 by packing.  Sizes the reference refuses return HKBCN_ERR_BAD_ARG with its message. */
HKBCN_API int hkbcn_create_normal_map(const float* normals_xyz, uint32_t width, uint32_t height, int pack, int mipmapped, int edge_mode,
                                      int filter, struct hkbcn_texture_desc* desc, void* dst, size_t capacity);
/* hkbcn_create_normal_and_roughness = Hk::CreateNormalAndRoughness(NormalRoughnessImportSettings) (Image.cpp:2459-2638) with
 * bCompressNormals and bCompressRoughness_BC4 set, from the two images already decoded (the reference loads them from files
 * as RGB32F and R8): the roughness map is resampled to powers of two, the normal map (green inverted for DirectX maps) to the
 * roughness map's size, normalised; the roughness chain's levels >= 1 are updated from the averaged normals (vMF / Toksvig)
 * and compressed to BC4; the normal map is packed, mipped and compressed at the powers of two of its own size.  All
 * resampling uses `edge_mode` and the triangle filter, as the reference does. */
typedef struct hkbcn_normal_roughness_settings {
    const float*   normals_rgb32f; uint32_t normal_width, normal_height;       /* Float3 per texel, as loaded (not yet normalised) */
    const uint8_t* roughness_r8;   uint32_t roughness_width, roughness_height;
    int pack;                  /* hkbcn_normal_pack */
    int convert_from_directx;  /* bConvertFromDirectXNormalMap: g = 1 - g first */
    int edge_mode;             /* 0 = wrap */
    int roughness_bake;        /* HKBCN_ROUGHNESS_BAKE_VMF / _TOKSVIG */
} hkbcn_normal_roughness_settings;
HKBCN_API int hkbcn_create_normal_and_roughness(const hkbcn_normal_roughness_settings* settings, struct hkbcn_texture_desc* normal_desc,
                                                void* normal_dst, size_t normal_capacity, struct hkbcn_texture_desc* roughness_desc,
                                                void* roughness_dst, size_t roughness_capacity);

/* ---- environment probes: GGX-prefiltered reflection chain and irradiance cube ----
 * The reference computes these by rendering into layered cubemaps through OpenGL (EnvProbeGenerator::Generate,
 * Hork/RenderUtils/EnvProbeGenerator.cpp:182-240 with envprobegen.frag; IrradianceGenerator::Generate,
 * IrradianceGenerator.cpp:160-200 with irradiancegen.frag).  Here the shader arithmetic runs as CUDA kernels on plain buffers.
 * cube: six RGBA32F faces of src_size x src_size, face-major (+X -X +Y -Y +Z -Z), rows top to bottom, linear HDR.
 * Prefilter output: levels back to back, level L = six faces of (2^max_lod >> L)^2 RGBA32F texels, roughness L / max_lod
 * (hkbcn_envprobe_texels() texels in all; alpha = 1).  The reference uses max_lod = 7 (128^2 .. 1^2).
 * Irradiance output: six faces of out_size^2 (the reference uses 32).  Every level >= 4 texels wide can go straight into
 * hkbcn_compress(HKBCN_BC6H_UFLOAT, ...) face by face. */
HKBCN_API size_t hkbcn_envprobe_texels(int max_lod);
HKBCN_API int hkbcn_envprobe_prefilter(const float* cube_rgba32f, uint32_t src_size, int max_lod, float* dst_rgba32f);
HKBCN_API int hkbcn_irradiance(const float* cube_rgba32f, uint32_t src_size, uint32_t out_size, float* dst_rgba32f);
HKBCN_API int hkbcn_envprobe_prefilter_dev(int device, const float* d_cube, uint32_t src_size, int max_lod, float* d_dst, void* stream);
HKBCN_API int hkbcn_irradiance_dev(int device, const float* d_cube, uint32_t src_size, uint32_t out_size, float* d_dst, void* stream);

/* ---- instrumentation --------------------------------------------------------
 * Number of kernels this library launched since init (bench.py's gpu_launches). */
HKBCN_API uint64_t hkbcn_launch_count(void);
/* Pinned host memory helpers so callers can avoid a pageable-memory staging copy. */
HKBCN_API void* hkbcn_alloc_pinned(size_t bytes);
HKBCN_API void hkbcn_free_pinned(void* p);

/* ---- compress without any allocation ----------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, kernels and bytes out as hkbcn_compress (host pointers, one surface, synchronous, Width % 4 != 0 allowed),
 * on library device `device`.  The device staging of the texels and of the blocks and the encode kernel's tile counters live
 * in ONE file-scope __device__ array of HKBCN_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit;
 * the call runs no cudaMalloc / cudaFree / cudaHostAlloc and creates no stream or event.  Consequences:
 *   - hkbcn_init must have been called (it allocates streams and tables): otherwise HKBCN_ERR_NOT_INIT;
 *   - 256 + source bytes (rounded up to 256) + block bytes (+ for BC7 a work list of 4 bytes per block) must fit:
 *     1024 x 1024 RGBA8 -> BC7 or 512 x 512 RGBA32F -> BC6H
 *     do; anything larger is refused with HKBCN_ERR_BAD_ARG and a message that names both sizes -- use hkbcn_compress;
 *   - all calls on one device share the array, which keeps the previous call's contents: the counters are cleared at the
 *     start of each call, and calls from several threads are serialised on a per-device mutex.  Work that reaches the array
 *     any other way while a call is running would race with it. */
#define HKBCN_STATIC_SCRATCH_BYTES (6u << 20)
HKBCN_API size_t hkbcn_static_scratch_bytes(void);
HKBCN_API int hkbcn_compress_static(int device, int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t flags);

#ifdef __cplusplus
}
#endif
#endif /* HKBCN_H */
