// TEST / DEBUG INFRASTRUCTURE -- not part of the product, never linked into libhkbcn.so.
//
// Compiles the per-block BC6H encoder of hork-source_b200/csrc/bc6h_thread.cuh for the CPU and compares it block by
// block with the reference's bc6h_enc::EncodeBC6HU / S from oracle/_ref/libhkref.so (and, with `half`, the restated
// float->half conversion with Hork/Core/half.c over every float bit pattern).
// usage: bc6h_host_check libhkref.so blocks_rgba32f.bin [signed|x] [verbose|x] [coop] | bc6h_host_check libhkref.so half
// coop: run the task-parallel kernel's pipeline (bc6h_coop.cuh) instead of the thread-per-block one
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
#include "../hork-source_b200/csrc/bc6h_coop.cuh"
#include "../hork-source_b200/csrc/bc6h_layout.h"

using namespace hkbcn;

// The task-parallel kernel's pipeline (bc6h_coop.cuh: rough, rank, and per round setup -> optimise -> finish) run serially
// for one block: same per-lane functions as the kernel, the warp's collective parts (task lists, the best-candidate commit)
// replaced by plain loops.
static uint4 encode_block_coop(const float* hdr4, const int* tq4, const bc6h::Layout& L, bool isSigned, uint32_t nan_sign)
{
    using namespace bc6h;
    float tf4[64];
    for (int i = 0; i < 64; i++) tf4[i] = (float)tq4[i];
    Block blk{hdr4, tq4, 4, 1};
    blk.nan_sign = nan_sign;
    float rough[32];
    uint8_t order[32];
    for (int s = 0; s < 32; s++) {
        Ep e;
        float err = rough_region<8>(blk, ~kShape2[s] & 0xFFFFu, isSigned, e);
        err += rough_region<8>(blk, kShape2[s], isSigned, e);
        rough[s] = err;
        order[s] = (uint8_t)s;
    }
    rank_shapes(rough, order);
    Ep unq8[8][2], unq1[1];
    for (int i = 0; i < 8; i++) {
        rough_region<8>(blk, ~kShape2[order[i]] & 0xFFFFu, isSigned, unq8[i][0]);
        rough_region<8>(blk, kShape2[order[i]], isSigned, unq8[i][1]);
    }
    rough_region<16>(blk, 0xFFFFu, isSigned, unq1[0]);
    float best_err = FLT_MAX;
    int best_ord = 0x7FFFFFFF;
    uint4 out = make_uint4(0, 0, 0, 0);
    static TaskRec rec[COOP_MAX_TASKS];
    for (int round = 0; round < 3; round++) {
        int n = 0;
        if (round < 2) {
            for (int gi = 0; gi < 3; gi++)
                for (int rank = 0; rank < 8; rank++) {
                    const int g = round * 3 + gi;
                    if (coop_setup<8, 2>(blk, 0, kGroupMode0[g], kGroupModes[g], rank, isSigned, order[rank], unq8[rank], rec[n])) n++;
                }
        } else {
            for (int m = 13; m >= 10; m--)
                if (coop_setup<16, 1>(blk, 0, m, 1, 0, isSigned, 0, unq1, rec[n])) n++;
        }
        for (int ch = 0; ch < 3; ch++)
            for (int t = 0; t < n; t++) {
                const int mode0 = (rec[t].meta >> 8) & 15;
                const int prec = kModes[mode0].prec[0][0];
                const uint32_t shapeIdx = (rec[t].meta >> 20) & 31u;
                for (int p = 0; p < (round < 2 ? 2 : 1); p++) {
                    const uint32_t mask = round < 2 ? (p ? kShape2[shapeIdx] : 0xFFFFu) : 0xFFFFu;
#define RUN(N, NR, CH, SG)                                                                                   \
    {                                                                                                        \
        OptChan<N, SG> M;                                                                           \
        M.ch = CH;                                                                                           \
        M.zero2 = f2_pack(0.0f, 0.0f);                                                                       \
        Ep e;                                                                                                \
        rec_load_ep<NR>(rec[t], p, e);                                                                       \
        M.start(tf4, mask, prec, e, rec[t].cur_err[p]);                                                      \
        while (M.busy) {                                                                                     \
            if (!M.in_trial) M.advance();                                                                    \
            if (M.busy && M.in_trial) M.eval_texel();                                                        \
        }                                                                                                    \
        e.A[CH] = M.a;                                                                                       \
        e.B[CH] = M.b;                                                                                       \
        rec_store_ep<NR>(rec[t], p, e);                                                                      \
        rec[t].cur_err[p] = M.opt_err;                                                                       \
    }
#define RUN_CH(N, NR, SG)                                                                                    \
    if (ch == 0) RUN(N, NR, 0, SG) else if (ch == 1) RUN(N, NR, 1, SG) else RUN(N, NR, 2, SG)
                    if (round < 2) { if (isSigned) { RUN_CH(8, 2, true) } else { RUN_CH(8, 2, false) } }
                    else { if (isSigned) { RUN_CH(16, 1, true) } else { RUN_CH(16, 1, false) } }
                }
            }
        for (int t = 0; t < n; t++) {
            const int rank = (rec[t].meta >> 4) & 15;
            if (round < 2) coop_finish<8, 2>(blk, L, rec[t], isSigned, unq8[rank], best_err, best_ord, out);
            else coop_finish<16, 1>(blk, L, rec[t], isSigned, unq1, best_err, best_ord, out);
        }
    }
    return out;
}

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
    const bool verbose = argc > 4 && !strcmp(argv[4], "verbose");
    bool coop = false;
    for (int a = 3; a < argc; a++) coop = coop || !strcmp(argv[a], "coop");
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
        bc6h::Block blk{hdr, ip, 3, 1};
        uint4 o;
        if (coop) {
            float hdr4[64];
            int ip4[64];
            for (int k = 0; k < 16; k++)
                for (int c = 0; c < 4; c++) {
                    hdr4[4 * k + c] = c < 3 ? hdr[3 * k + c] : 0.0f;
                    ip4[4 * k + c] = c < 3 ? ip[3 * k + c] : 0;
                }
            o = encode_block_coop(hdr4, ip4, L, isSigned, 0x80000000u);
        } else o = bc6h::encode_block(blk, L, isSigned);
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
#ifdef HKBCN_HOST_COUNT
    printf("trial length histogram (texels evaluated):");
    { long t = 0; for (int i = 0; i < 17; i++) t += bc6h::g_hist[i]; for (int i = 0; i < 17; i++) printf(" %d:%.1f%%", i, 100.0 * bc6h::g_hist[i] / t); printf("\n"); }
    printf("pair evaluations per block: rough %.0f, two-region trials %.0f (%.0f trials), one-region trials %.0f (%.0f trials), assign %.0f\n", (double)bc6h::g_cnt[0] / n,
           (double)bc6h::g_cnt[1] / n, (double)bc6h::g_cnt[4] / n, (double)bc6h::g_cnt[2] / n, (double)bc6h::g_cnt[5] / n, (double)bc6h::g_cnt[3] / n);
#endif
    printf("%ld blocks, %ld differ\n", n, diff);
    return 0;
}
