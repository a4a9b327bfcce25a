// TEST / DEBUG INFRASTRUCTURE -- not part of the product, never linked into libhkbcn.so.
//
// Compiles the per-block BC7 encoder of hork-source_b200/csrc/bc7_thread.cuh for the CPU (it is plain serial
// code per thread) and compares it block by block with the reference's bc7enc_compress_block from
// oracle/_ref/libhkref.so, so that deviations can be chased without a GPU.  Build + run: tools/bc7_host_check.sh
#define HKBCN_HOST_EMU 1
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <dlfcn.h>
#include <algorithm>
#define __device__
#define __forceinline__ inline
#define __noinline__
struct uint4 { uint32_t x, y, z, w; };
struct float4 { float x, y, z, w; };
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
static inline float __ull2float_rn(unsigned long long v) { return (float)v; }
using std::max;
using std::min;
#include "../hork-source_b200/csrc/bc7_host_tables.h"

using namespace hkbcn;

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s libhkref.so blocks.bin [verbose]\n", argv[0]); return 2; }
    void* h = dlopen(argv[1], RTLD_NOW);
    if (!h) { fprintf(stderr, "%s\n", dlerror()); return 2; }
    auto ref_encode_block = (void (*)(int, const void*, void*))dlsym(h, "ref_encode_block");
    FILE* f = fopen(argv[2], "rb");
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t* buf = (uint8_t*)malloc(sz);
    if (fread(buf, 1, sz, f) != (size_t)sz) return 2;
    fclose(f);
    static bc7::Tables T;
    if (!bc7::fill_tables(&T)) { fprintf(stderr, "weights mismatch\n"); return 2; }
    long n = sz / 64, diff = 0;
    if (argc > 4) {
        // detailed view of one block: per-mode errors of this encoder
        long i = atol(argv[4]);
        uint32_t px[16];
        memcpy(px, buf + 64 * i, 64);
        bc7::Block blk{px, 1};
        bc7::Sub r6[1], r5[1], r7[2], r1[2], r6o[1];
        uint64_t s6, s5, s7, s1, as5;
        bc7::color_cell_compression<6, true>(blk, 0, T, r6, s6);
        bc7::color_cell_compression<5, false>(blk, 0, T, r5, s5);
        int alo, ahi;
        uint32_t aerr = bc7::mode5_alpha(blk, T, alo, ahi, as5);
        uint32_t p7 = bc7::estimate_partition<true>(blk);
        bc7::color_cell_compression<7, true>(blk, kShape2[p7], T, r7, s7);
        printf("alpha path: err6 %u (lo %08x hi %08x pb %u) | err5 %u + a %u (a %d..%d) | part7 %u err7 %u + %u (lo %08x/%08x hi %08x/%08x pb %u/%u sels %016llx)\n", r6[0].err, r6[0].lo,
               r6[0].hi, r6[0].pb, r5[0].err, aerr, alo, ahi, p7, r7[0].err, r7[1].err, r7[0].lo, r7[1].lo, r7[0].hi, r7[1].hi, r7[0].pb, r7[1].pb, (unsigned long long)s7);
        uint32_t p1 = bc7::estimate_partition<false>(blk);
        bc7::color_cell_compression<6, false>(blk, 0, T, r6o, s6);
        bc7::color_cell_compression<1, false>(blk, kShape2[p1], T, r1, s1);
        printf("opaque path: err6 %u | part1 %u err1 %u + %u\n", r6o[0].err, p1, r1[0].err, r1[1].err);
        return 0;
    }
    for (long i = 0; i < n; i++) {
        uint32_t px[16];
        memcpy(px, buf + 64 * i, 64);
        bc7::Block blk{px, 1};
        bool alpha = false;
        for (int k = 0; k < 16; k++) alpha |= (px[k] >> 24) != 255;
        // HKBCN_BC7_LEVEL (environment): uber level 0..4 of Encode_BC7 (default 4, the CompressBC7 preset)
        static const int level = getenv("HKBCN_BC7_LEVEL") ? atoi(getenv("HKBCN_BC7_LEVEL")) : 4;
        uint4 o = alpha ? bc7::encode_alpha(blk, T, level) : bc7::encode_opaque(blk, T, level);
        uint8_t r[16];
        if (level == 4) ref_encode_block(7, px, r);
        else {
            static auto ref_level = (void (*)(const void*, void*, uint32_t))dlsym(h, "ref_encode_bc7_level");
            ref_level(px, r, (uint32_t)level);
        }
        if (memcmp(&o, r, 16)) {
            diff++;
            if (argc > 3) {
                printf("block %ld differs (alpha=%d)\n  mine:", i, alpha);
                for (int k = 0; k < 16; k++) printf(" %02x", ((uint8_t*)&o)[k]);
                printf("\n  ref: ");
                for (int k = 0; k < 16; k++) printf(" %02x", r[k]);
                printf("\n  px:");
                for (int k = 0; k < 16; k++) printf(" %08x", px[k]);
                printf("\n");
            }
        }
    }
    printf("%ld blocks, %ld differ\n", n, diff);
    return 0;
}
