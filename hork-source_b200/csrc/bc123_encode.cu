// bc123_encode.cu -- BC1 / BC2 / BC3 surface encoders (sm_100a).
//
// Replaces TextureBlockCompression::CompressBC1/2/3 (Hork/Image/ImageEncoders.cpp:215-320):
//   BC1: Encode_BC1(.., 5, false, false)         -> rgbcx::encode_bc1 level 5   (see bc1_block.cuh)
//   BC2: 4-bit explicit alpha (ImageEncoders.cpp:152-157) + the BC1 colour block
//   BC3: Encode_BC3(.., 5, bMaxQuality = true)   -> rgbcx::encode_bc3_hq = encode_bc4_hq on alpha
//        (stride 4, radius 3, all modes; see bc4_search.cuh) + the BC1 colour block
// Bit-exact contract for all three.
//
// Mapping: BC1/BC2 one lane per block; BC3 a 2-lane sub-warp per block for the alpha search (lane
// parity = BC4 alpha mode; a tile's two halves one after the other), then one lane per block for
// the colour blocks.  128-thread CTAs, tiles of 128 blocks staged by TMA bulk copies.  Algorithmic HBM bytes:
// 4.5 B/px (BC1), 5 B/px (BC2/BC3).
#include "hkbcn_common.cuh"
#include "hkbcn_kernels.h"
#include "bc4_search.cuh"
#include "bc1_block.cuh"
#include "bc1_hq.cuh"
#include "build/bc1_orderings.inc"
#if HK_HAVE_BC1_ORDERINGS
// rgbcx's total-ordering tables as one blob (tools/gen_bc1_orderings.py: "HKO1" | histograms4 969 x 4 u8 | histograms3 153 x 3 u8 |
// pad | best4 969 x 128 u16 | best3 153 x 32 u8), embedded by the assembler
__asm__(".section .rodata\n"
        ".balign 16\n"
        ".hidden hk_bc1_orderings_blob\n"
        ".global hk_bc1_orderings_blob\n"
        "hk_bc1_orderings_blob:\n"
        ".incbin \"build/bc1_orderings.bin\"\n"
        ".previous\n");
extern "C" const unsigned char hk_bc1_orderings_blob[];
#endif
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <string.h>

namespace hkbcn {

// HQ: the colour block goes through the error-tracking branch of bc1_hq.cuh (levels 6..18, level 5 with a 3-colour flag)
template <int CODEC, bool HQ>
__global__ void __launch_bounds__(128) bc123_encode_kernel(const __grid_constant__ EncodeParams P)
{
    constexpr int NT = 128;
    constexpr int TB = NT; // BC3 too: the alpha search runs on lane pairs over the tile's two halves, the colour blocks on all lanes
    constexpr int BW = 16;
    constexpr int ROWB = TB * BW;
    __shared__ __align__(128) uint8_t stage[4 * ROWB];
    __shared__ __align__(8) uint2 s_alpha[CODEC == C_BC3 ? TB : 1];
    __shared__ uint8_t s_list[CODEC == C_BC3 ? TB : 1];
    __shared__ uint32_t s_nsearch;
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        s_nsearch = 0;
    }
    __syncthreads();
    uint32_t phase = 0;
    const int k = threadIdx.x;
    // HKBCN_FLAG_LEVEL: rgbcx level + 1 in bits 4..8 (0 = the Compress* preset, level 5); HKBCN_FLAG_FAST: non-hq BC3 alpha
    const int level = ((P.flags >> 4) & 31u) ? (int)((P.flags >> 4) & 31u) - 1 : 5;
    const bool fastAlpha = (P.flags & 2u) != 0;
    // Encode_BC2 / Encode_BC3 pass (false, false) for the 3-colour options (ImageEncoders.cpp:158, rgbcx.cpp:2897-2911)
    uint32_t to4 = 1, to3 = 1;
    const uint32_t rflags = HQ ? bc1_level_flags(level, CODEC == C_BC1 && (P.flags & (1u << 9)), CODEC == C_BC1 && (P.flags & (1u << 10)), to4, to3) : 0u;
    auto colour_block = [&](const uint32_t* px16) { return HQ ? bc1_hq_encode_block(px16, rflags, to4, to3) : bc1_encode_block(px16, level); };

    for (uint32_t tile = blockIdx.x; tile < P.totalTiles; tile += gridDim.x) {
        const Level& L = P.lv[find_level(P, tile)];
        const uint32_t b0 = (tile - L.tileBase) * TB;
        const uint32_t nb = min((uint32_t)TB, L.nblocks - b0);
        stage_tile<4, TB, NT>(L, b0, nb, stage, &bar, phase);

        const bool valid = (uint32_t)k < nb;
        uint32_t px[16];
        if (valid) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
                uint4 v = *(const uint4*)(stage + r * ROWB + k * BW);
                px[r * 4 + 0] = v.x; px[r * 4 + 1] = v.y; px[r * 4 + 2] = v.z; px[r * 4 + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) px[i] = 0;
        }

        if (CODEC == C_BC1) {
            if (valid) *(uint2*)(L.dst + (size_t)(b0 + k) * 8) = colour_block(px);
        } else if (CODEC == C_BC2) {
            if (valid) {
                uint32_t a[2] = {0, 0};
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    // min(int(a * (15.0f / 255.0f) + 0.5f), 15)   (ImageEncoders.cpp:154-155)
                    int q = __float2int_rz(__fadd_rn(__fmul_rn((float)(px[i] >> 24), 15.0f / 255.0f), 0.5f));
                    q = min(q, 15);
                    a[i >> 3] |= (uint32_t)q << (4 * (i & 7));
                }
                uint2 c = colour_block(px);
                *(uint4*)(L.dst + (size_t)(b0 + k) * 16) = make_uint4(a[0], a[1], c.x, c.y);
            }
        } else { // BC3
            // ---- alpha (rgbcx::encode_bc4_hq / encode_bc4 on channel 3).  Every lane first looks at its own block: a solid alpha
            // block (and, without bMaxQuality, every block: rgbcx::encode_bc3 = the non-hq encode_bc4, rgbcx.cpp:2886-2903) is
            // finished on the spot, the others are appended to a list.  The radius-3 search then runs on lane pairs (lane parity =
            // BC4 mode) over the LIST, 64 entries a round: opaque regions of a texture no longer leave their lane pairs idle.
            {
                uint32_t a2[8];
                int minv = 255, maxv = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int p0 = px[2 * i] >> 24, p1 = px[2 * i + 1] >> 24;
                    a2[i] = (uint32_t)p0 | ((uint32_t)p1 << 16);
                    minv = min(minv, min(p0, p1));
                    maxv = max(maxv, max(p0, p1));
                }
                const bool search = valid && !fastAlpha && minv != maxv;
                if (valid && !search) s_alpha[k] = fastAlpha ? bc4_fast(a2, minv, maxv) : bc4_solid((uint32_t)minv);
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, search);
                uint32_t base = 0;
                if ((threadIdx.x & 31u) == 0 && m) base = atomicAdd(&s_nsearch, (uint32_t)__popc(m));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (search) s_list[base + __popc(m & ((1u << (threadIdx.x & 31u)) - 1u))] = (uint8_t)k;
            }
            __syncthreads();
            const uint32_t nsearch = s_nsearch;
#pragma unroll 1
            for (uint32_t e0 = 0; e0 < nsearch; e0 += TB / 2) {
                const uint32_t e = e0 + (threadIdx.x >> 1);
                const int sub = (int)(threadIdx.x & 1u);
                const bool va = e < nsearch;
                const int ka = va ? (int)s_list[e] : 0;
                uint32_t a2[8];
                int minv = 255, maxv = 0;
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    const uint4 v = *(const uint4*)(stage + r * ROWB + ka * BW);
                    const int p0 = v.x >> 24, p1 = v.y >> 24, p2 = v.z >> 24, p3 = v.w >> 24;
                    a2[2 * r] = (uint32_t)p0 | ((uint32_t)p1 << 16);
                    a2[2 * r + 1] = (uint32_t)p2 | ((uint32_t)p3 << 16);
                    minv = min(minv, min(min(p0, p1), min(p2, p3)));
                    maxv = max(maxv, max(max(p0, p1), max(p2, p3)));
                }
                uint32_t key = 0xFFFFFFFFu;
                if (va) key = bc4hq_search_mode(a2, minv, maxv, sub);
                key = min(key, __shfl_xor_sync(0xFFFFFFFFu, key, 1));
                if (va && sub == 0) s_alpha[ka] = bc4hq_emit(a2, minv, maxv, key & 127u);
            }
            __syncthreads();
            if (threadIdx.x == 0) s_nsearch = 0; // everybody has read it; the barrier that ends the tile orders this before the next appends
            // ---- colour: one lane per block, every lane busy
            if (valid) {
                const uint2 al = s_alpha[k];
                const uint2 c = colour_block(px);
                *(uint4*)(L.dst + (size_t)(b0 + k) * 16) = make_uint4(al.x, al.y, c.x, c.y);
            }
        }
        __syncthreads();
    }
}

uint32_t bc123_tile_blocks(int) { return 128; }

// ---- a12: table generation (host) -------------------------------------------
// prepare_bc1_single_color_table (rgbcx.cpp:485-527) in cBC1Ideal mode: for each 8-bit value
// the (hi, lo) 5/6-bit pair whose 2/3-1/3 interpolant is closest, with a 3% endpoint-spread
// penalty and ties broken towards hi == lo.
static void gen_match(uint16_t* out, int size)
{
    uint8_t expand[64];
    for (int i = 0; i < size; i++) expand[i] = (uint8_t)(size == 32 ? ((i << 3) | (i >> 2)) : ((i << 2) | (i >> 4)));
    for (int i = 0; i < 256; i++) {
        int lowest = 256, bh = 0, bl = 0;
        for (int lo = 0; lo < size; lo++) {
            for (int hi = 0; hi < size; hi++) {
                int lo_e = expand[lo], hi_e = expand[hi];
                int v = (hi_e * 2 + lo_e) / 3;
                int e = abs(v - i) + (abs(hi_e - lo_e) * 3) / 100;
                if (e < lowest || (e == lowest && lo == hi)) { bh = hi; bl = lo; lowest = e; }
            }
        }
        out[i] = (uint16_t)(bh | (bl << 8));
    }
}
// g_midpoint5/6 (rgbcx.cpp:693-694): midpoint between consecutive expanded 5/6-bit levels
// in [0,1], written in the reference as 6-decimal literals; the last entry is 1e+37.
static void gen_mid(float* out, int size)
{
    for (int i = 0; i < size - 1; i++) {
        int a = size == 32 ? ((i << 3) | (i >> 2)) : ((i << 2) | (i >> 4));
        int b = size == 32 ? (((i + 1) << 3) | ((i + 1) >> 2)) : (((i + 1) << 2) | ((i + 1) >> 4));
        char buf[32];
        snprintf(buf, sizeof(buf), "%.6f", (a + b) / 2.0 / 255.0);
        out[i] = strtof(buf, nullptr);
    }
    out[size - 1] = 1e+37f;
}

// the same search keeping the error, for the 2/3-1/3 interpolant (prepare_bc1_single_color_table, :485-522) or the midpoint
// (prepare_bc1_single_color_table_half, :447-483): hi | lo << 8 | e << 16
static void gen_match_e(uint32_t* out, int size, bool half)
{
    uint8_t expand[64];
    for (int i = 0; i < size; i++) expand[i] = (uint8_t)(size == 32 ? ((i << 3) | (i >> 2)) : ((i << 2) | (i >> 4)));
    for (int i = 0; i < 256; i++) {
        int lowest = 256, bh = 0, bl = 0;
        for (int lo = 0; lo < size; lo++) {
            for (int hi = 0; hi < size; hi++) {
                const int lo_e = expand[lo], hi_e = expand[hi];
                const int v = half ? (hi_e + lo_e) / 2 : (hi_e * 2 + lo_e) / 3;
                const int e = abs(v - i) + (abs(hi_e - lo_e) * 3) / 100;
                if (e < lowest || (e == lowest && lo == hi)) { bh = hi; bl = lo; lowest = e; }
            }
        }
        out[i] = (uint32_t)(bh | (bl << 8) | (lowest << 16));
    }
}
// compute_selector_factors4 / 3 (rgbcx.cpp:529-571): the inverse of the 2 x 2 normal matrix of a selector histogram
static void selector_factors(const uint8_t* hist, int n, float* out)
{
    static const uint32_t w4[4] = {0x000009, 0x010204, 0x040201, 0x090000}, w3[3] = {0x000004, 0x040000, 0x010101};
    uint32_t acc = 0;
    for (int s = 0; s < n; s++) acc += (n == 4 ? w4[s] : w3[s]) * hist[s];
    const float z00 = (float)((acc >> 16) & 0xFF), z10 = (float)((acc >> 8) & 0xFF), z11 = (float)(acc & 0xFF), z01 = z10;
    volatile float a = z00 * z11, b = z01 * z10; // two rounded products, as the reference's baseline x86-64 build computes them
    float det = a - b;
    if (fabsf(det) < 1e-8f) det = 0.0f;
    else det = ((n == 4 ? 3.0f : 2.0f) / 255.0f) / det;
    out[0] = z11 * det;
    out[1] = -z10 * det;
    out[2] = z00 * det;
}

bool bc1_hq_available() { return HK_HAVE_BC1_ORDERINGS != 0; }

cudaError_t bc1_upload_tables()
{
    uint16_t m5[256], m6[256];
    float mid5[32], mid6[64];
    gen_match(m5, 32);
    gen_match(m6, 64);
    gen_mid(mid5, 32);
    gen_mid(mid6, 64);
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(g_bc1_match5, m5, sizeof(m5))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_bc1_match6, m6, sizeof(m6))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_bc1_mid5, mid5, sizeof(mid5))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_bc1_mid6, mid6, sizeof(mid6))) != cudaSuccess) return e;
    uint32_t t[256];
    gen_match_e(t, 32, false);
    if ((e = cudaMemcpyToSymbol(g_hq_eq5, t, sizeof(t))) != cudaSuccess) return e;
    gen_match_e(t, 64, false);
    if ((e = cudaMemcpyToSymbol(g_hq_eq6, t, sizeof(t))) != cudaSuccess) return e;
    gen_match_e(t, 32, true);
    if ((e = cudaMemcpyToSymbol(g_hq_half5, t, sizeof(t))) != cudaSuccess) return e;
    gen_match_e(t, 64, true);
    if ((e = cudaMemcpyToSymbol(g_hq_half6, t, sizeof(t))) != cudaSuccess) return e;
#if HK_HAVE_BC1_ORDERINGS
    // rgbcx::init (:591-625): histogram -> index, and the least-squares factors of every histogram
    static uint16_t hash4[4096], hash3[256];
    static float fac4[HQ_N4 * 3], fac3[HQ_N3 * 3];
    memset(hash4, 0, sizeof(hash4));
    memset(hash3, 0, sizeof(hash3));
    const unsigned char* blob = hk_bc1_orderings_blob;
    if (memcmp(blob, "HKO1", 4) != 0 || HK_BC1_ORDERINGS_BYTES != 4 + HQ_N4 * 4 + HQ_N3 * 3 + 1 + HQ_N4 * HQ_MAX4 * 2 + HQ_N3 * HQ_MAX3) return cudaErrorInvalidValue;
    const uint8_t* orders4 = blob + 4;
    const uint8_t* orders3 = orders4 + HQ_N4 * 4;
    const uint8_t* best4 = orders3 + HQ_N3 * 3 + 1; // u16 little-endian, 2-byte aligned within the 16-byte aligned blob
    const uint8_t* best3 = best4 + HQ_N4 * HQ_MAX4 * 2;
    for (int i = 0; i < HQ_N4; i++) {
        const uint8_t* h = orders4 + 4 * i;
        if (h[0] != 16 && h[1] != 16 && h[2] != 16 && h[3] != 16) hash4[h[0] | (h[1] << 4) | (h[2] << 8)] = (uint16_t)i;
        selector_factors(h, 4, fac4 + 3 * i);
    }
    for (int i = 0; i < HQ_N3; i++) {
        const uint8_t* h = orders3 + 3 * i;
        if (h[0] != 16 && h[1] != 16 && h[2] != 16) hash3[h[0] | (h[1] << 4)] = (uint16_t)i;
        selector_factors(h, 3, fac3 + 3 * i);
    }
    if ((e = cudaMemcpyToSymbol(g_hq_orders4, orders4, HQ_N4 * 4)) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_best4, best4, HQ_N4 * HQ_MAX4 * 2)) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_hash4, hash4, sizeof(hash4))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_fac4, fac4, sizeof(fac4))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_orders3, orders3, HQ_N3 * 3)) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_best3, best3, HQ_N3 * HQ_MAX3)) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_hash3, hash3, sizeof(hash3))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(g_hq_fac3, fac3, sizeof(fac3))) != cudaSuccess) return e;
#endif
    return cudaSuccess;
}

// the branch of rgbcx::encode_bc1 a launch takes: levels 6..18 always track the block error, level 5 when a 3-colour option
// is on (BC1 only -- Encode_BC2 / Encode_BC3 never pass one); below 5 the options are not read (rgbcx.cpp:1700-1726)
bool bc123_needs_hq(int codec, uint32_t flags)
{
    const int level = ((flags >> 4) & 31u) ? (int)((flags >> 4) & 31u) - 1 : 5;
    return level >= 6 || (level == 5 && codec == C_BC1 && (flags & ((1u << 9) | (1u << 10))) != 0);
}

cudaError_t launch_bc123_encode(int codec, const EncodeParams& P, int sm_count, cudaStream_t stream)
{
    if (P.totalTiles == 0) return cudaSuccess;
    uint32_t g = (uint32_t)sm_count * 8;
    int grid = (int)(P.totalTiles < g ? P.totalTiles : g);
    if (bc123_needs_hq(codec, P.flags)) {
        if (!bc1_hq_available()) return cudaErrorNotSupported;
        if (codec == C_BC1) bc123_encode_kernel<C_BC1, true><<<grid, 128, 0, stream>>>(P);
        else if (codec == C_BC2) bc123_encode_kernel<C_BC2, true><<<grid, 128, 0, stream>>>(P);
        else bc123_encode_kernel<C_BC3, true><<<grid, 128, 0, stream>>>(P);
        return cudaGetLastError();
    }
    if (codec == C_BC1) bc123_encode_kernel<C_BC1, false><<<grid, 128, 0, stream>>>(P);
    else if (codec == C_BC2) bc123_encode_kernel<C_BC2, false><<<grid, 128, 0, stream>>>(P);
    else bc123_encode_kernel<C_BC3, false><<<grid, 128, 0, stream>>>(P);
    return cudaGetLastError();
}

} // namespace hkbcn
