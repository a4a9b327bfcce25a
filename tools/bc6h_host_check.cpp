// TEST / DEBUG INFRASTRUCTURE -- not part of the product, never linked into libhkbcn.so.
//
// Compiles the per-block BC6H encoder of hork-source_b200/csrc/bc6h_thread.cuh for the CPU and compares it block by
// block with the reference's bc6h_enc::EncodeBC6HU / S from oracle/_ref/libhkref.so (and, with `half`, the restated
// float->half conversion with Hork/Core/half.c over every float bit pattern).
// usage: bc6h_host_check libhkref.so blocks_rgba32f.bin [signed] [verbose] | bc6h_host_check libhkref.so half
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
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
static inline int __popc(uint32_t v) { return __builtin_popcount(v); }
static inline int __ffs(uint32_t v) { return __builtin_ffs((int)v); }
static inline int __clz(uint32_t v) { return v ? __builtin_clz(v) : 32; }
static inline uint32_t __brev(uint32_t v) { uint32_t r = 0; for (int i = 0; i < 32; i++) r |= ((v >> i) & 1u) << (31 - i); return r; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
using std::max;
using std::min;
#include "../hork-source_b200/csrc/bc6h_thread.cuh"
#include "../hork-source_b200/csrc/bc6h_layout.h"

using namespace hkbcn;

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s libhkref.so blocks.bin|half [signed] [verbose]\n", argv[0]); return 2; }
    void* h = dlopen(argv[1], RTLD_NOW);
    if (!h) { fprintf(stderr, "%s\n", dlerror()); return 2; }
    if (!strcmp(argv[2], "half")) {
        auto ref_h = (uint16_t (*)(uint32_t))dlsym(h, "ref_half_from_float");
        uint64_t bad = 0;
        for (uint64_t f = 0; f < (1ull << 32); f += 1) {
            if (bc6h::half_from_float_bits((uint32_t)f) != ref_h((uint32_t)f)) {
                if (bad++ < 5) printf("half mismatch at %08x: %04x vs %04x\n", (uint32_t)f, bc6h::half_from_float_bits((uint32_t)f), ref_h((uint32_t)f));
            }
        }
        printf("half_from_float: %llu mismatches over 2^32 inputs\n", (unsigned long long)bad);
        return bad != 0;
    }
    auto ref_encode_block = (void (*)(int, const void*, void*))dlsym(h, "ref_encode_block");
    const bool isSigned = argc > 3 && !strcmp(argv[3], "signed");
    const bool verbose = argc > 4;
    FILE* f = fopen(argv[2], "rb");
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t* buf = (uint8_t*)malloc(sz);
    if (fread(buf, 1, sz, f) != (size_t)sz) return 2;
    fclose(f);
    static bc6h::Layout L;
    bc6h_parse_layout(L.tok);
    long n = sz / 256, diff = 0;
    for (long i = 0; i < n; i++) {
        const float* src = (const float*)(buf + 256 * i);
        float hdr[48];
        int ip[48];
        for (int k = 0; k < 16; k++)
            for (int c = 0; c < 3; c++) {
                hdr[3 * k + c] = src[4 * k + c];
                ip[3 * k + c] = bc6h::f16_to_int(bc6h::half_from_float_bits(__float_as_uint(src[4 * k + c])), isSigned);
            }
        bc6h::Block blk{hdr, ip, 1};
        uint4 o = bc6h::encode_block(blk, L, isSigned);
        uint8_t r[16];
        ref_encode_block(isSigned ? 6 : 5, src, r);
        if (memcmp(&o, r, 16)) {
            diff++;
            if (verbose) {
                printf("block %ld differs\n  mine:", i);
                for (int k = 0; k < 16; k++) printf(" %02x", ((uint8_t*)&o)[k]);
                printf("\n  ref: ");
                for (int k = 0; k < 16; k++) printf(" %02x", r[k]);
                printf("\n");
            }
        }
    }
    printf("%ld blocks, %ld differ\n", n, diff);
    return 0;
}
