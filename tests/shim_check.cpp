// TEST INFRASTRUCTURE.  Runs the drop-in itself: links hork-source_b200/libhkbcn_hork.so (namespace
// Hk::TextureBlockCompression over the C ABI), takes the seven Compress* entry points through the function-pointer
// type Image.cpp:1718 uses, calls them and the sixteen per-block functions on malloc'd (pageable) buffers -- what
// Image.cpp's HeapBlob memory is to the CUDA runtime -- and compares every result with the reference's own code in
// oracle/_ref (dlopen'ed: libhkref.so, and libhkref0.so for the BC6H encoder, see oracle/Makefile).
//
//   shim_check <libhkref.so> <libhkref0.so>      prints "shim_check: <n> checks, <m> differ"; exit code = min(m, 255)
#include "../hork-source_b200/host/HkTextureBlockCompression.h"

#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace TBC = Hk::TextureBlockCompression;

typedef void (*ref_compress_t)(int, const void*, void*, uint32_t, uint32_t, int);
typedef void (*ref_encode_block_t)(int, const void*, void*);
typedef void (*ref_decode_block_t)(int, const void*, void*, size_t, int);
typedef void (*ref_encode_bc45_t)(const void*, void*, int);
typedef void (*ref_encode_bc6h_f16_t)(const void*, void*, int);
typedef uint16_t (*ref_half_from_float_t)(uint32_t);
typedef void (*ref_encode_bc1_level_t)(const void*, void*, uint32_t, int, int);
typedef void (*ref_encode_bc7_level_t)(const void*, void*, uint32_t);

static uint32_t g_rng = 12345;
static uint32_t rnd() { g_rng = g_rng * 1664525u + 1013904223u; return g_rng >> 8; }

static int g_checks = 0, g_bad = 0;
static void expect_equal(const char* what, const void* a, const void* b, size_t n)
{
    g_checks++;
    if (memcmp(a, b, n) != 0) {
        g_bad++;
        size_t i = 0;
        while (((const uint8_t*)a)[i] == ((const uint8_t*)b)[i]) i++;
        fprintf(stderr, "shim_check: %s differs from the reference at byte %zu of %zu\n", what, i, n);
    }
}

// the load-bearing signature: CreateImage stores Compress* in exactly this type (Image.cpp:1718, :2271)
typedef void (*CompressFn)(void const*, void*, uint32_t, uint32_t);

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: shim_check libhkref.so libhkref0.so\n"); return 2; }
    void* R = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    void* R0 = dlopen(argv[2], RTLD_NOW | RTLD_LOCAL);
    if (!R || !R0) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    ref_compress_t ref_compress = (ref_compress_t)dlsym(R, "ref_compress");
    ref_compress_t ref0_compress = (ref_compress_t)dlsym(R0, "ref_compress");
    ref_encode_block_t ref_encode_block = (ref_encode_block_t)dlsym(R, "ref_encode_block");
    ref_encode_block_t ref0_encode_block = (ref_encode_block_t)dlsym(R0, "ref_encode_block");
    ref_decode_block_t ref_decode_block = (ref_decode_block_t)dlsym(R, "ref_decode_block");
    ref_encode_bc45_t ref_encode_bc4 = (ref_encode_bc45_t)dlsym(R, "ref_encode_bc4");
    ref_encode_bc45_t ref_encode_bc5 = (ref_encode_bc45_t)dlsym(R, "ref_encode_bc5");
    ref_encode_bc6h_f16_t ref0_encode_bc6h_f16 = (ref_encode_bc6h_f16_t)dlsym(R0, "ref_encode_bc6h_f16");
    ref_half_from_float_t ref_half_from_float = (ref_half_from_float_t)dlsym(R, "ref_half_from_float");
    ref_encode_bc1_level_t ref_encode_bc1_level = (ref_encode_bc1_level_t)dlsym(R, "ref_encode_bc1_level");
    ref_encode_bc7_level_t ref_encode_bc7_level = (ref_encode_bc7_level_t)dlsym(R, "ref_encode_bc7_level");
    if (!ref_encode_bc1_level || !ref_encode_bc7_level) { fprintf(stderr, "missing level symbols in the reference shim\n"); return 2; }
    if (!ref_compress || !ref0_compress || !ref_encode_block || !ref0_encode_block || !ref_decode_block || !ref_encode_bc4 || !ref_encode_bc5 ||
        !ref0_encode_bc6h_f16 || !ref_half_from_float) { fprintf(stderr, "missing symbol in the reference shim\n"); return 2; }

    // ---- whole surfaces through the function-pointer type, pageable memory, widths that are and are not multiples of 4
    struct { const char* name; CompressFn fn; int ref_fmt; int bpp; int bb; } surf[] = {
        {"CompressBC1", &TBC::CompressBC1, 0, 4, 8},  {"CompressBC2", &TBC::CompressBC2, 1, 4, 16}, {"CompressBC3", &TBC::CompressBC3, 2, 4, 16},
        {"CompressBC4", &TBC::CompressBC4, 3, 1, 8},  {"CompressBC5", &TBC::CompressBC5, 4, 2, 16}, {"CompressBC7", &TBC::CompressBC7, 7, 4, 16}};
    const uint32_t dims[][2] = {{256, 128}, {64, 64}, {22, 8}, {37, 19}, {4, 4}};
    for (auto& s : surf)
        for (auto& d : dims) {
            const uint32_t w = d[0], h = d[1];
            const size_t sbytes = (size_t)w * h * s.bpp, dbytes = (size_t)(w / 4) * (h / 4) * s.bb;
            uint8_t* src = (uint8_t*)malloc(sbytes);
            uint8_t* got = (uint8_t*)malloc(dbytes + 1);
            uint8_t* want = (uint8_t*)malloc(dbytes + 1);
            for (uint32_t y = 0; y < h; y++)
                for (uint32_t x = 0; x < w; x++)
                    for (int c = 0; c < s.bpp; c++) {
                        int v = (int)((x * 3 + y * 5 + c * 40) & 255) + (int)(rnd() % 23) - 11;
                        if (c == 3) v = ((x / 8 + y / 8) & 1) ? 255 : 128 + (int)(rnd() % 100); // opaque and translucent blocks
                        src[((size_t)y * w + x) * s.bpp + c] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
                    }
            memset(got, 0xCD, dbytes + 1);
            memset(want, 0xCD, dbytes + 1);
            s.fn(src, got, w, h);
            ref_compress(s.ref_fmt, src, want, w, h, 4);
            char what[64];
            snprintf(what, sizeof(what), "%s %ux%u", s.name, w, h);
            expect_equal(what, got, want, dbytes + 1); // + the guard byte: nothing written past the last block
            free(src); free(got); free(want);
        }
    {   // CompressBC6h (five arguments, not storable in CompressFn): unsigned and signed, 32 x 16
        const uint32_t w = 32, h = 16;
        float* src = (float*)malloc((size_t)w * h * 16);
        for (uint32_t i = 0; i < w * h; i++) {
            const float base = 0.05f + 3.0f * (float)(i / w) / h;
            src[4 * i + 0] = base * (0.5f + (rnd() % 1000) / 1000.0f);
            src[4 * i + 1] = base * (0.7f + (rnd() % 1000) / 1500.0f);
            src[4 * i + 2] = base * (1.0f + (rnd() % 1000) / 2000.0f) + ((i % 97) == 0 ? 4000.0f : 0.0f);
            src[4 * i + 3] = 1.0f;
        }
        const size_t dbytes = (size_t)(w / 4) * (h / 4) * 16;
        uint8_t* got = (uint8_t*)malloc(dbytes);
        uint8_t* want = (uint8_t*)malloc(dbytes);
        for (int sgn = 0; sgn < 2; sgn++) {
            if (sgn) for (uint32_t i = 0; i < w * h; i += 3) src[4 * i] = -src[4 * i];
            TBC::CompressBC6h(src, got, w, h, sgn != 0);
            ref0_compress(sgn ? 6 : 5, src, want, w, h, 4);
            expect_equal(sgn ? "CompressBC6h signed" : "CompressBC6h unsigned", got, want, dbytes);
        }
        // ---- per-block encoders (ImageEncoders.h:57-64), the presets and variants the device path implements
        uint8_t blk[64], g16[16], w16[16];
        for (int t = 0; t < 24; t++) {
            for (int i = 0; i < 64; i++) blk[i] = (uint8_t)((t * 9 + i * 3 + rnd() % 31) & 255);
            if (t & 1) for (int i = 3; i < 64; i += 4) blk[i] = 255;
            TBC::Encode_BC1(blk, g16, 5, false, false); ref_encode_block(0, blk, w16); expect_equal("Encode_BC1", g16, w16, 8);
            TBC::Encode_BC2(blk, g16, 5);               ref_encode_block(1, blk, w16); expect_equal("Encode_BC2", g16, w16, 16);
            TBC::Encode_BC3(blk, g16, 5, true);         ref_encode_block(2, blk, w16); expect_equal("Encode_BC3", g16, w16, 16);
            TBC::Encode_BC7(blk, g16, 4);               ref_encode_block(7, blk, w16); expect_equal("Encode_BC7", g16, w16, 16);
            // quality levels (row a15): rgbcx levels 0..3 of Encode_BC1 (the 3-colour flags are only read from level 5 up),
            // Encode_BC3 without bMaxQuality, bc7enc uber levels 0..3
            const uint32_t lvl = (uint32_t)(t & 3);
            TBC::Encode_BC1(blk, g16, lvl, true, true); ref_encode_bc1_level(blk, w16, lvl, 1, 1); expect_equal("Encode_BC1 level 0..3", g16, w16, 8);
            TBC::Encode_BC7(blk, g16, lvl);             ref_encode_bc7_level(blk, w16, lvl);       expect_equal("Encode_BC7 level 0..3", g16, w16, 16);
            TBC::Encode_BC1(blk, g16, 4, true, true);   ref_encode_bc1_level(blk, w16, 4, 1, 1);   expect_equal("Encode_BC1 level 4", g16, w16, 8);
            // the levels with the total-orderings search and the 3-colour options (read from level 5 up)
            const uint32_t hi = 5 + (uint32_t)(t * 7 % 14);
            if (t & 2) for (int i = 0; i < 16; i += 3) blk[4 * i] = blk[4 * i + 1] = blk[4 * i + 2] = (uint8_t)(i & 3); // near-black texels
            TBC::Encode_BC1(blk, g16, hi, (t & 1) != 0, (t & 2) != 0); ref_encode_bc1_level(blk, w16, hi, t & 1, (t & 2) >> 1); expect_equal("Encode_BC1 level 5..18 with options", g16, w16, 8);
            TBC::Encode_BC2(blk, g16, hi); ref_encode_bc1_level(blk, w16 + 8, hi, 0, 0); expect_equal("Encode_BC2 level 5..18 (colour)", g16 + 8, w16 + 8, 8);
            {
                uint8_t al[16];
                for (int i = 0; i < 16; i++) al[i] = blk[4 * i + 3];
                TBC::Encode_BC3(blk, g16, lvl, false);
                ref_encode_bc4(al, w16, 0);
                ref_encode_bc1_level(blk, w16 + 8, lvl, 0, 0);
                expect_equal("Encode_BC3 bMaxQuality = false", g16, w16, 16);
            }
            for (int hq = 0; hq < 2; hq++) {
                TBC::Encode_BC4(blk, g16, hq != 0); ref_encode_bc4(blk, w16, hq); expect_equal(hq ? "Encode_BC4 hq" : "Encode_BC4", g16, w16, 8);
                TBC::Encode_BC5(blk, g16, hq != 0); ref_encode_bc5(blk, w16, hq); expect_equal(hq ? "Encode_BC5 hq" : "Encode_BC5", g16, w16, 16);
            }
        }
        uint16_t half[64];
        for (int t = 0; t < 4; t++) {
            const float* f = src + 64 * t;
            for (int i = 0; i < 64; i++) { uint32_t u; float v = f[i] < 0 ? -f[i] : f[i]; memcpy(&u, &v, 4); half[i] = ref_half_from_float(u); }
            TBC::Encode_BC6h_f32(f, g16, false); ref0_encode_block(5, f, w16); expect_equal("Encode_BC6h_f32", g16, w16, 16);
            TBC::Encode_BC6h_f32(f, g16, true);  ref0_encode_block(6, f, w16); expect_equal("Encode_BC6h_f32 signed", g16, w16, 16);
            TBC::Encode_BC6h_f16(half, g16, false); ref0_encode_bc6h_f16(half, w16, 0); expect_equal("Encode_BC6h_f16", g16, w16, 16);
        }
        free(src); free(got); free(want);
    }
    // ---- per-block decoders on random 128-bit blocks, odd row strides
    {
        uint8_t blk[16];
        static uint8_t got[4 * 80], want[4 * 80];
        for (int t = 0; t < 40; t++) {
            for (int i = 0; i < 16; i++) blk[i] = (uint8_t)rnd();
            if (t < 8) blk[0] = (uint8_t)(1u << t); // every BC7 mode
            const size_t stride = 48 + (size_t)(t % 5) * 8;
            #define DEC(NAME, CALL, FMT, F32, ROWB)                                             \
                memset(got, 0xEE, sizeof(got)); memset(want, 0xEE, sizeof(want));                \
                CALL; ref_decode_block(FMT, blk, want, stride, F32); expect_equal(NAME, got, want, 3 * stride + ROWB);
            DEC("Decode_BC1", TBC::Decode_BC1(blk, got, stride), 0, 0, 16)
            DEC("Decode_BC2", TBC::Decode_BC2(blk, got, stride), 1, 0, 16)
            DEC("Decode_BC3", TBC::Decode_BC3(blk, got, stride), 2, 0, 16)
            DEC("Decode_BC4", TBC::Decode_BC4(blk, got, stride), 3, 0, 4)
            DEC("Decode_BC5", TBC::Decode_BC5(blk, got, stride), 4, 0, 8)
            DEC("Decode_BC7", TBC::Decode_BC7(blk, got, stride), 7, 0, 16)
            DEC("Decode_BC6h_f16", TBC::Decode_BC6h_f16(blk, got, stride, false), 5, 0, 24)
            DEC("Decode_BC6h_f16 signed", TBC::Decode_BC6h_f16(blk, got, stride, true), 6, 0, 24)
            DEC("Decode_BC6h_f32", TBC::Decode_BC6h_f32(blk, got, stride, false), 5, 1, 48)
            DEC("Decode_BC6h_f32 signed", TBC::Decode_BC6h_f32(blk, got, stride, true), 6, 1, 48)
            #undef DEC
        }
    }
    printf("shim_check: %d checks, %d differ\n", g_checks, g_bad);
    return g_bad > 255 ? 255 : g_bad;
}
