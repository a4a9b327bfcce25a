// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// ref_shim.cpp: C entry points over the *unmodified* reference codec sources
// (compiled where they lie under /root/reference by oracle/Makefile into
// oracle/_ref/libhkref.so).  Nothing in here is arithmetic: it only restates
//   * the wrapper argument choices of Hork/Image/ImageEncoders.cpp:138-212
//     (BC1/2/3 level 5, BC3/4/5 hq, BC7 uber level 4, BC6H half macros), and
//   * the row-major 4x4 gather loops of ImageEncoders.cpp:215-557,
// which cannot be compiled as-is because ImageEncoders.cpp drags in Hork/Core.
//
// Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference arm, as the checker and the timed CPU baseline.

#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <thread>
#include <vector>

#include <bc7enc_rdo/rgbcx.h>
#include <bc7enc_rdo/bc7decomp.h>
#include <bc7enc_rdo/bc7enc.h>

#define BCDEC_IMPLEMENTATION
#include <bcdec/bcdec.h>

extern "C" uint16_t half_from_float(uint32_t f);      // Hork/Core/half.c:802
extern "C" uint32_t fast_half_to_float(uint16_t h);   // Hork/Core/half.c:309

// ImageEncoders.cpp:43-44
#define BC6H_HALF_TO_FLOAT(h) fast_half_to_float(h)
#define BC6H_FLOAT_TO_HALF(f) half_from_float(f)
#define BC6H_ENC_IMPLEMENTATION
#include <bc6h_enc/bc6h_enc.h>

#define STB_IMAGE_RESIZE_IMPLEMENTATION
#define STBIR_MAX_CHANNELS 4
#include <stb/stb_image_resize.h>

namespace {

// ImageEncoders.cpp:52-68 (static CompressionParams)
struct Params {
    bc7enc_compress_block_params bc7[BC7ENC_MAX_UBER_LEVEL + 1];
    Params() {
        rgbcx::init();
        bc7enc_compress_block_init();
        for (uint32_t i = 0; i <= BC7ENC_MAX_UBER_LEVEL; i++) {
            bc7enc_compress_block_params_init(&bc7[i]);
            bc7[i].m_uber_level = i;
        }
    }
};
Params g_params;

// Format ids == Hk::TEXTURE_FORMAT enum values (Hork/Image/Image.h:122-148)
enum {
    F_BC1 = 0, F_BC2, F_BC3, F_BC4, F_BC5, F_BC6H_U, F_BC6H_S, F_BC7
};

inline float f16tof32(uint16_t h) { uint32_t r = fast_half_to_float(h); float f; memcpy(&f, &r, 4); return f; }

// ImageEncoders.cpp:138-212
void encode_block(int fmt, const void* src, void* dst)
{
    switch (fmt) {
    case F_BC1: rgbcx::encode_bc1(5, dst, (const uint8_t*)src, false, false); break;
    case F_BC2: {
        const uint8_t* pixels = (const uint8_t*)src;
        uint8_t* block = (uint8_t*)dst;
        for (int i = 0; i < 8; ++i, pixels += 8) {
            int q1 = std::min(int(pixels[3] * (15.0f / 255.0f) + 0.5f), 15);
            int q2 = std::min(int(pixels[7] * (15.0f / 255.0f) + 0.5f), 15);
            *block++ = (uint8_t)(q1 | (q2 << 4));
        }
        rgbcx::encode_bc1(5, block, (const uint8_t*)src, false, false);
        break;
    }
    case F_BC3: rgbcx::encode_bc3_hq(5, dst, (const uint8_t*)src); break;
    case F_BC4: rgbcx::encode_bc4_hq(dst, (const uint8_t*)src, 1); break;
    case F_BC5: rgbcx::encode_bc5_hq(dst, (const uint8_t*)src, 0, 1, 2); break;
    case F_BC6H_U: bc6h_enc::EncodeBC6HU(dst, src); break;
    case F_BC6H_S: bc6h_enc::EncodeBC6HS(dst, src); break;
    case F_BC7: bc7enc_compress_block(dst, src, &g_params.bc7[BC7ENC_MAX_UBER_LEVEL]); break;
    }
}

int src_bpp(int fmt) { return fmt == F_BC4 ? 1 : fmt == F_BC5 ? 2 : (fmt == F_BC6H_U || fmt == F_BC6H_S) ? 16 : 4; }
int blk_bytes(int fmt) { return (fmt == F_BC1 || fmt == F_BC4) ? 8 : 16; }

// gather (ImageEncoders.cpp:229-246 and siblings) + per-block encode over [b0,b1)
void compress_range(int fmt, const uint8_t* src, uint8_t* dst, uint32_t nbx, size_t rowStride, uint32_t b0, uint32_t b1)
{
    const int bpp = src_bpp(fmt);
    const int brs = 4 * bpp;
    const int bb = blk_bytes(fmt);
    alignas(16) uint8_t block[4 * 4 * 16];
    for (uint32_t i = b0; i < b1; i++) {
        uint32_t bx = i % nbx, by = i / nbx;
        const uint8_t* p = src + (size_t)by * 4 * rowStride + (size_t)bx * brs;
        for (int r = 0; r < 4; r++, p += rowStride) memcpy(block + brs * r, p, brs);
        encode_block(fmt, block, dst + (size_t)i * bb);
    }
}

} // namespace

extern "C" {

int ref_hw_threads(void) { int n = (int)std::thread::hardware_concurrency(); return n > 0 ? n : 1; }

// per-block entry points (ImageEncoders.h:57-64 with the presets Compress* bakes in)
void ref_encode_block(int fmt, const void* src, void* dst) { encode_block(fmt, src, dst); }

// generic BC1 level / flags access (ImageEncoders.cpp:138-143)
void ref_encode_bc1_level(const void* src, void* dst, uint32_t level, int allow3, int transparentBlack)
{
    rgbcx::encode_bc1(level, dst, (const uint8_t*)src, allow3 != 0, transparentBlack != 0);
}
void ref_encode_bc4(const void* src, void* dst, int hq)
{
    if (hq) rgbcx::encode_bc4_hq(dst, (const uint8_t*)src, 1); else rgbcx::encode_bc4(dst, (const uint8_t*)src, 1);
}
void ref_encode_bc5(const void* src, void* dst, int hq)
{
    if (hq) rgbcx::encode_bc5_hq(dst, (const uint8_t*)src, 0, 1, 2); else rgbcx::encode_bc5(dst, (const uint8_t*)src, 0, 1, 2);
}
void ref_encode_bc7_level(const void* src, void* dst, uint32_t level)
{
    bc7enc_compress_block(dst, src, &g_params.bc7[level > BC7ENC_MAX_UBER_LEVEL ? BC7ENC_MAX_UBER_LEVEL : level]);
}
// ImageEncoders.cpp:187-197
void ref_encode_bc6h_f16(const void* src, void* dst, int bSigned)
{
    float block[64];
    const uint16_t* h = (const uint16_t*)src;
    for (int i = 0; i < 64; i++) block[i] = f16tof32(h[i]);
    encode_block(bSigned ? F_BC6H_S : F_BC6H_U, block, dst);
}

// ImageEncoders.cpp:77-136
void ref_decode_block(int fmt, const void* src, void* dst, size_t rowStride, int f32out)
{
    switch (fmt) {
    case F_BC1: bcdec_bc1(src, dst, (int)rowStride); break;
    case F_BC2: bcdec_bc2(src, dst, (int)rowStride); break;
    case F_BC3: bcdec_bc3(src, dst, (int)rowStride); break;
    case F_BC4: bcdec_bc4(src, dst, (int)rowStride); break;
    case F_BC5: bcdec_bc5(src, dst, (int)rowStride); break;
    case F_BC6H_U:
    case F_BC6H_S:
        if (f32out) bcdec_bc6h_float(src, dst, (int)(rowStride / sizeof(float)), fmt == F_BC6H_S);
        else bcdec_bc6h_half(src, dst, (int)(rowStride / sizeof(uint16_t)), fmt == F_BC6H_S);
        break;
    case F_BC7:
        if (rowStride == 16) bc7decomp::unpack_bc7(src, (bc7decomp::color_rgba*)dst);
        else {
            bc7decomp::color_rgba tmp[4][4];
            bc7decomp::unpack_bc7(src, &tmp[0][0]);
            uint8_t* d = (uint8_t*)dst;
            for (int i = 0; i < 4; i++, d += rowStride) memcpy(d, tmp[i], 16);
        }
        break;
    }
}

// alternative BC7 decoder in the tree (bcdec.h:908), compiled out of Hork by '#if 1'
void ref_decode_bc7_bcdec(const void* src, void* dst, size_t rowStride) { bcdec_bc7(src, dst, (int)rowStride); }

// Whole-surface compress: the reference's Compress* loops, all blocks encoded
// (ImageEncoders.cpp:215-557).  nthreads<=1: the as-shipped single-thread loop.
// nthreads>1: contiguous block ranges per thread, the decomposition CompressBC6h
// uses (ImageEncoders.cpp:426-486) -- except that the tail numBlocks%nthreads
// blocks ARE encoded here (the reference drops them, SURVEY.md 7 "defect").
void ref_compress(int fmt, const void* src, void* dst, uint32_t w, uint32_t h, int nthreads)
{
    uint32_t nbx = w / 4, nby = h / 4, nb = nbx * nby;
    size_t rowStride = (size_t)w * src_bpp(fmt);
    if (nthreads <= 1 || nb < (uint32_t)nthreads) {
        compress_range(fmt, (const uint8_t*)src, (uint8_t*)dst, nbx, rowStride, 0, nb);
        return;
    }
    std::vector<std::thread> th;
    uint32_t per = (nb + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        uint32_t b0 = std::min(nb, per * t), b1 = std::min(nb, per * (t + 1));
        if (b0 < b1) th.emplace_back(compress_range, fmt, (const uint8_t*)src, (uint8_t*)dst, nbx, rowStride, b0, b1);
    }
    for (auto& t : th) t.join();
}

// Whole-surface decode (the reference has per-block Decode_* only; this is the
// obvious loop over them, used as checker for the surface decode kernels).
// BC6H output is RGB (3 comps/pixel) half or float, as Decode_BC6h_f16/f32.
static void decompress_range(int fmt, const uint8_t* src, uint8_t* dst, uint32_t nbx, size_t dstRowStride, int f32out, uint32_t b0, uint32_t b1)
{
    int bb = blk_bytes(fmt);
    int opp = (fmt == F_BC4) ? 1 : (fmt == F_BC5) ? 2 : (fmt == F_BC6H_U || fmt == F_BC6H_S) ? (f32out ? 12 : 6) : 4;
    for (uint32_t i = b0; i < b1; i++) {
        uint32_t bx = i % nbx, by = i / nbx;
        ref_decode_block(fmt, src + (size_t)i * bb, dst + (size_t)by * 4 * dstRowStride + (size_t)bx * 4 * opp, dstRowStride, f32out);
    }
}
void ref_decompress(int fmt, const void* src, void* dst, uint32_t w, uint32_t h, size_t dstRowStride, int f32out, int nthreads)
{
    uint32_t nbx = w / 4, nby = h / 4, nb = nbx * nby;
    if (nthreads <= 1 || nb < (uint32_t)nthreads) {
        decompress_range(fmt, (const uint8_t*)src, (uint8_t*)dst, nbx, dstRowStride, f32out, 0, nb);
        return;
    }
    std::vector<std::thread> th;
    uint32_t per = (nb + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        uint32_t b0 = std::min(nb, per * t), b1 = std::min(nb, per * (t + 1));
        if (b0 < b1) th.emplace_back(decompress_range, fmt, (const uint8_t*)src, (uint8_t*)dst, nbx, dstRowStride, f32out, b0, b1);
    }
    for (auto& t : th) t.join();
}

// Mip cascade step (Image.cpp:1787-1796 for uint8, :1881-1893 for float):
// level i from level i-1 through stbir_resize.
// datatype: 0 = uint8, 3 = float (stbir_datatype); edge: stbir_edge; filter: stbir_filter;
// space: 0 linear, 1 sRGB.  alpha_channel -1 = none.
int ref_stbir_resize(const void* in, int iw, int ih, void* out, int ow, int oh, int datatype,
                     int nch, int alpha_channel, int flags, int edge, int filter, int space)
{
    int es = (datatype == 3) ? 4 : 1;
    return stbir_resize(in, iw, ih, iw * nch * es, out, ow, oh, ow * nch * es, (stbir_datatype)datatype, nch, alpha_channel, flags,
                        (stbir_edge)edge, (stbir_edge)edge, (stbir_filter)filter, (stbir_filter)filter, (stbir_colorspace)space, NULL);
}

uint16_t ref_half_from_float(uint32_t f) { return half_from_float(f); }
uint32_t ref_half_to_float(uint16_t h) { return fast_half_to_float(h); }

// Table dumps for pinning the restatement's table generators (a12)
void ref_bc7_params(uint32_t level, uint32_t* out /*[12]*/)
{
    const bc7enc_compress_block_params& p = g_params.bc7[level];
    out[0] = p.m_mode_mask; out[1] = p.m_max_partitions; out[2] = p.m_weights[0]; out[3] = p.m_weights[1];
    out[4] = p.m_weights[2]; out[5] = p.m_weights[3]; out[6] = p.m_uber_level; out[7] = p.m_perceptual;
    out[8] = p.m_try_least_squares; out[9] = p.m_mode17_partition_estimation_filterbank;
    out[10] = p.m_force_alpha; out[11] = p.m_force_selectors;
}

} // extern "C"
