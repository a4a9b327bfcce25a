// bcn_decode.cu -- BC1..BC7 / BC6H surface decoders (sm_100a).
//
// Bit-exact contract with the decoders Hork wires behind Decode_BC*
// (Hork/Image/ImageEncoders.cpp:77-136):
//   BC1/2/3  bcdec__color_block / __sharp_alpha_block / __smooth_alpha_block (ThirdParty/bcdec/bcdec.h:93-210)
//   BC4/5    bcdec__smooth_alpha_block                                         (bcdec.h:170-210, 264-272)
//   BC6H     bcdec_bc6h_half / bcdec_bc6h_float                                 (bcdec.h:375-902)
//   BC7      bc7decomp::unpack_bc7                                              (ThirdParty/bc7enc_rdo/bc7decomp.cpp:594-634)
//
// HBM-bound: 8/16 B in, 16..192 B out per block.  One thread decodes one block; the 32
// threads of a warp hold 32 horizontally adjacent blocks, so each of the 4 row stores of a
// warp is one contiguous run (32 x 16 B = 512 B for RGBA8) written with 128-bit stores --
// no shared-memory transpose needed.  BC4 packs 4 blocks per thread to reach 128-bit stores.
#include <atomic>
#include "hkbcn_common.cuh"
#include "hkbcn_kernels.h"
#include "bcn_tables.cuh"
#include "bc6h_layout.h"
#include "bcn_decode_lean.cuh"
#include "bc7_decode_lean.cuh"
#include "bc6h_decode_lean.cuh"
#include <cuda_fp16.h>

namespace hkbcn {

// ---------------------------------------------------------------- BC1 colour block
// bcdec.h:93-151.  out[i] = 0xAABBGGRR of pixel i.
__device__ __forceinline__ void decode_color_block(uint2 blk, bool onlyOpaque, uint32_t out[16])
{
    uint32_t c0 = blk.x & 0xFFFF, c1 = blk.x >> 16;
    uint32_t r0 = (((c0 >> 11) & 0x1F) * 527 + 23) >> 6;
    uint32_t g0 = (((c0 >> 5) & 0x3F) * 259 + 33) >> 6;
    uint32_t b0 = ((c0 & 0x1F) * 527 + 23) >> 6;
    uint32_t r1 = (((c1 >> 11) & 0x1F) * 527 + 23) >> 6;
    uint32_t g1 = (((c1 >> 5) & 0x3F) * 259 + 33) >> 6;
    uint32_t b1 = ((c1 & 0x1F) * 527 + 23) >> 6;
    uint32_t ref[4];
    ref[0] = 0xFF000000u | (b0 << 16) | (g0 << 8) | r0;
    ref[1] = 0xFF000000u | (b1 << 16) | (g1 << 8) | r1;
    if (c0 > c1 || onlyOpaque) {
        ref[2] = 0xFF000000u | (((2 * b0 + b1 + 1) / 3) << 16) | (((2 * g0 + g1 + 1) / 3) << 8) | ((2 * r0 + r1 + 1) / 3);
        ref[3] = 0xFF000000u | (((b0 + 2 * b1 + 1) / 3) << 16) | (((g0 + 2 * g1 + 1) / 3) << 8) | ((r0 + 2 * r1 + 1) / 3);
    } else {
        ref[2] = 0xFF000000u | (((b0 + b1 + 1) >> 1) << 16) | (((g0 + g1 + 1) >> 1) << 8) | ((r0 + r1 + 1) >> 1);
        ref[3] = 0;
    }
    uint32_t idx = blk.y;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t s = (idx >> (2 * i)) & 3;
        out[i] = s == 0 ? ref[0] : s == 1 ? ref[1] : s == 2 ? ref[2] : ref[3];
    }
}

// bcdec.h:170-210: 8-entry alpha palette + 16 x 3-bit indices.
__device__ __forceinline__ void decode_smooth_alpha(uint2 blk, uint32_t out[16])
{
    uint32_t a0 = blk.x & 0xFF, a1 = (blk.x >> 8) & 0xFF;
    uint32_t a[8];
    a[0] = a0;
    a[1] = a1;
    if (a0 > a1) {
        a[2] = (6 * a0 + a1 + 1) / 7;
        a[3] = (5 * a0 + 2 * a1 + 1) / 7;
        a[4] = (4 * a0 + 3 * a1 + 1) / 7;
        a[5] = (3 * a0 + 4 * a1 + 1) / 7;
        a[6] = (2 * a0 + 5 * a1 + 1) / 7;
        a[7] = (a0 + 6 * a1 + 1) / 7;
    } else {
        a[2] = (4 * a0 + a1 + 1) / 5;
        a[3] = (3 * a0 + 2 * a1 + 1) / 5;
        a[4] = (2 * a0 + 3 * a1 + 1) / 5;
        a[5] = (a0 + 4 * a1 + 1) / 5;
        a[6] = 0;
        a[7] = 255;
    }
    // pack the palette into two registers so the per-pixel pick is a byte-permute
    uint32_t plo = a[0] | (a[1] << 8) | (a[2] << 16) | (a[3] << 24);
    uint32_t phi = a[4] | (a[5] << 8) | (a[6] << 16) | (a[7] << 24);
    uint64_t idx = (((uint64_t)blk.y << 32) | blk.x) >> 16;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t s = (uint32_t)(idx >> (3 * i)) & 7;
        out[i] = __byte_perm(plo, phi, s) & 0xFF;
    }
}

// ---------------------------------------------------------------- BC7
struct Bits128 {
    uint64_t lo, hi;
    uint32_t pos;
    __device__ __forceinline__ uint32_t read(uint32_t n)
    {
        // n in [0, 8]
        uint64_t v;
        if (pos >= 64) v = hi >> (pos - 64);
        else if (pos == 0) v = lo;
        else v = (lo >> pos) | (hi << (64 - pos));
        pos += n;
        return (uint32_t)v & ((1u << n) - 1u);
    }
};

__device__ __forceinline__ uint32_t bc7_weight(uint32_t bits, uint32_t i)
{
    // g_bc7_weights2/3/4 (bc7decomp.cpp:31-33), nibble/byte packed
    if (bits == 2) return (0x402B1500u >> (8 * i)) & 0xFF;                        // 0,21,43,64
    if (bits == 3) return (uint32_t)((0x40372E251B120900ull >> (8 * i)) & 0xFF);  // 0,9,18,27,37,46,55,64
    const uint64_t lo = 0x1E1A15110D090400ull, hi = 0x403C37332F2B2622ull;        // 0,4,9,...,64
    return (uint32_t)(((i < 8 ? lo : hi) >> (8 * (i & 7))) & 0xFF);
}

// bc7decomp.cpp:168-634 semantics; out[i] = 0xAABBGGRR.
__device__ void decode_bc7_block(uint4 blk, uint32_t out[16])
{
    uint32_t b0 = blk.x & 0xFF;
    if (b0 == 0) {
#pragma unroll
        for (int i = 0; i < 16; i++) out[i] = 0;
        return;
    }
    const uint32_t mode = __ffs(b0) - 1;
    // per-mode format constants, nibble packed (index = mode)
    const uint32_t NS = (0x21112323u >> (4 * mode)) & 0xF;   // subsets          3,2,3,2,1,1,1,2
    const uint32_t PB = (0x60006664u >> (4 * mode)) & 0xF;   // partition bits   4,6,6,6,0,0,0,6
    const uint32_t CB = (0x57757564u >> (4 * mode)) & 0xF;   // colour bits      4,6,5,7,5,7,7,5
    const uint32_t AB = (0x57860000u >> (4 * mode)) & 0xF;   // alpha bits       0,0,0,0,6,8,7,5
    const uint32_t IB = (0x24222233u >> (4 * mode)) & 0xF;   // index bits       3,3,2,2,2,2,4,2
    const uint32_t IB2 = (0x00230000u >> (4 * mode)) & 0xF;  // 2nd index bits   0,0,0,0,3,2,0,0
    const bool uniqueP = (0xC9u >> mode) & 1;                // modes 0,3,6,7
    const bool sharedP = mode == 1;

    Bits128 bs;
    bs.lo = ((uint64_t)blk.y << 32) | blk.x;
    bs.hi = ((uint64_t)blk.w << 32) | blk.z;
    bs.pos = mode + 1;
    const uint32_t part = bs.read(PB);
    uint32_t rot = 0, idxSel = 0;
    if (mode == 4 || mode == 5) rot = bs.read(2);
    if (mode == 4) idxSel = bs.read(1);

    const uint32_t NE = NS * 2;
    uint32_t ep[6][4]; // [endpoint][channel]
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint32_t nb = c < 3 ? CB : AB;
#pragma unroll
        for (int e = 0; e < 6; e++) {
            if ((uint32_t)e < NE) ep[e][c] = nb ? bs.read(nb) : 255u;
        }
    }
    uint32_t pbits[6];
#pragma unroll
    for (int e = 0; e < 6; e++) pbits[e] = 0;
    if (uniqueP) {
#pragma unroll
        for (int e = 0; e < 6; e++)
            if ((uint32_t)e < NE) pbits[e] = bs.read(1);
    } else if (sharedP) {
        uint32_t p0 = bs.read(1), p1 = bs.read(1);
        pbits[0] = pbits[1] = p0;
        pbits[2] = pbits[3] = p1;
    }
    const bool hasP = uniqueP || sharedP;
#pragma unroll
    for (int e = 0; e < 6; e++) {
        if ((uint32_t)e >= NE) continue;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t nb = c < 3 ? CB : AB;
            if (nb == 0) continue; // alpha stays 255
            uint32_t v = ep[e][c], tb = nb;
            if (hasP) {
                v = (v << 1) | pbits[e];
                tb = nb + 1;
            }
            v <<= (8 - tb);
            v |= v >> tb;
            ep[e][c] = v;
        }
    }

    // subsets and anchors
    uint32_t subsetMask = 0; // 2 bits per pixel
    uint32_t anchor1 = 0xFF, anchor2 = 0xFF;
    if (NS == 2) {
        uint32_t m = kShape2[part];
#pragma unroll
        for (int i = 0; i < 16; i++) subsetMask |= ((m >> i) & 1u) << (2 * i);
        anchor1 = kAnchor2[part];
    } else if (NS == 3) {
        subsetMask = kShape3[part];
        anchor1 = kAnchor3a[part];
        anchor2 = kAnchor3b[part];
    }

    uint32_t idx1[16], idx2[16];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t n = IB - ((i == 0 || (uint32_t)i == anchor1 || (uint32_t)i == anchor2) ? 1u : 0u);
        idx1[i] = bs.read(n);
    }
    if (IB2) {
#pragma unroll
        for (int i = 0; i < 16; i++) idx2[i] = bs.read(IB2 - (i == 0 ? 1u : 0u));
    }

#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t s = (subsetMask >> (2 * i)) & 3;
        uint32_t l[4], h[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            l[c] = s == 0 ? ep[0][c] : s == 1 ? ep[2][c] : ep[4][c];
            h[c] = s == 0 ? ep[1][c] : s == 1 ? ep[3][c] : ep[5][c];
        }
        uint32_t cw, aw;
        if (IB2 == 0) {
            cw = aw = bc7_weight(IB, idx1[i]);
        } else if (idxSel) {
            // mode 4, index selection bit set: colour uses the 3-bit set, alpha the 2-bit set
            cw = bc7_weight(IB2, idx2[i]);
            aw = bc7_weight(IB, idx1[i]);
        } else {
            cw = bc7_weight(IB, idx1[i]);
            aw = bc7_weight(IB2, idx2[i]);
        }
        uint32_t r = (l[0] * (64 - cw) + h[0] * cw + 32) >> 6;
        uint32_t g = (l[1] * (64 - cw) + h[1] * cw + 32) >> 6;
        uint32_t b = (l[2] * (64 - cw) + h[2] * cw + 32) >> 6;
        uint32_t a = (l[3] * (64 - aw) + h[3] * aw + 32) >> 6;
        if (rot == 1) { uint32_t t = a; a = r; r = t; }
        else if (rot == 2) { uint32_t t = a; a = g; g = t; }
        else if (rot == 3) { uint32_t t = a; a = b; b = t; }
        out[i] = r | (g << 8) | (b << 16) | (a << 24);
    }
}

// ---------------------------------------------------------------- BC6H
// header bit layouts: bc6h_layout.h
__device__ uint16_t g_bc6h_layout[14 * BC6H_MAX_TOKENS];

cudaError_t bc7dec_upload_tables();
cudaError_t decode_upload_tables()
{
    uint16_t tab[14][BC6H_MAX_TOKENS];
    bc6h_parse_layout(tab);
    cudaError_t e = cudaMemcpyToSymbol(g_bc6h_layout, tab, sizeof(tab));
    if (e != cudaSuccess) return e;
    return bc7dec_upload_tables();
}

// bcdec.h:289-329
__device__ __forceinline__ int bc6h_unquantize(int val, int bits, bool isSigned)
{
    int unq;
    if (!isSigned) {
        if (bits >= 15) unq = val;
        else if (val == 0) unq = 0;
        else if (val == ((1 << bits) - 1)) unq = 0xFFFF;
        else unq = ((val << 16) + 0x8000) >> bits;
    } else {
        if (bits >= 16) unq = val;
        else {
            int s = 0;
            if (val < 0) { s = 1; val = -val; }
            if (val == 0) unq = 0;
            else if (val >= ((1 << (bits - 1)) - 1)) unq = 0x7FFF;
            else unq = ((val << 15) + 0x4000) >> (bits - 1);
            if (s) unq = -unq;
        }
    }
    return unq;
}
// bcdec.h:331-346
__device__ __forceinline__ uint32_t bc6h_finish(int val, bool isSigned)
{
    if (!isSigned) return (uint32_t)((val * 31) >> 6) & 0xFFFF;
    val = (val < 0) ? -(((-val) * 31) >> 5) : (val * 31) >> 5;
    uint32_t s = 0;
    if (val < 0) { s = 0x8000; val = -val; }
    return (s | (uint32_t)val) & 0xFFFF;
}
__device__ __forceinline__ int sext(int v, int bits) { return (v << (32 - bits)) >> (32 - bits); }

// bcdec__half_to_float_quick (bcdec.h:348-373): exact half -> float, Inf/NaN payload kept.
__device__ __forceinline__ uint32_t half_bits_to_float_bits(uint32_t h)
{
    uint32_t em = h & 0x7FFF, sign = (h & 0x8000) << 16;
    uint32_t e = em >> 10;
    uint32_t o;
    if (e == 31) o = (em << 13) + ((127 - 15) << 23) + ((128 - 16) << 23);
    else if (e == 0) o = __float_as_uint((float)em * 5.9604644775390625e-8f); // m * 2^-24, exact
    else o = (em << 13) + ((127 - 15) << 23);
    return o | sign;
}

// out[i*3+c] = half bits of pixel i channel c.  `lay` = layout table in shared memory.
__device__ void decode_bc6h_block(uint4 blk, bool isSigned, const uint16_t* lay, uint16_t out[48])
{
    Bits128 bs;
    bs.lo = ((uint64_t)blk.y << 32) | blk.x;
    bs.hi = ((uint64_t)blk.w << 32) | blk.z;
    bs.pos = 0;
    uint32_t code = bs.read(2);
    if (code > 1) code |= bs.read(3) << 2;
    // 5-bit code -> mode index 0..13 (-1 reserved)
    int mode;
    if (code < 2) mode = (int)code;
    else if ((code & 3) == 2) mode = 2 + (int)(code >> 2);      // 00010,00110,...,11110 -> 2..9
    else if (code < 16) mode = 10 + (int)(code >> 2);           // 00011,00111,01011,01111 -> 10..13
    else mode = -1;                                             // 10011,10111,11011,11111
    if (mode < 0) {
#pragma unroll
        for (int i = 0; i < 48; i++) out[i] = 0;
        return;
    }
    int f[13];
#pragma unroll
    for (int i = 0; i < 13; i++) f[i] = 0;
    const uint16_t* t = lay + mode * BC6H_MAX_TOKENS;
#pragma unroll 1
    for (int k = 0; k < BC6H_MAX_TOKENS; k++) {
        uint32_t tk = t[k];
        if (tk == 0xFFFF) break;
        uint32_t field = tk & 15, lsb = (tk >> 4) & 15, cnt = (tk >> 8) & 15, rev = tk >> 12;
        uint32_t v = bs.read(cnt);
        if (rev) v = __brev(v) >> (32 - cnt);
        v <<= lsb;
        // dynamic register-array index avoided: predicated select over the 13 fields
#pragma unroll
        for (int i = 0; i < 13; i++)
            if (field == (uint32_t)i) f[i] |= (int)v;
    }
    // actual_bits_count (bcdec.h:376-381), nibble packed per mode (W-1 stored for W to fit 16)
    const int W = (int)((0xFBA957778AAA69ull >> (4 * mode)) & 0xF) + 1; // 10,7,11,11,11,9,8,8,8,6,10,11,12,16
    const int dR = (int)((0x489A6556544565ull >> (4 * mode)) & 0xF);     // 5,6,5,4,4,5,6,5,5,6,10,9,8,4
    const int dG = (int)((0x489A6565545465ull >> (4 * mode)) & 0xF);     // 5,6,4,5,4,5,5,6,5,6,10,9,8,4
    const int dB = (int)((0x489A6655554465ull >> (4 * mode)) & 0xF);     // 5,6,4,4,5,5,5,5,6,6,10,9,8,4
    const bool two = mode < 10;
    const int part = two ? f[12] : 0;
    const int ne = two ? 4 : 2;

    if (isSigned) {
        f[0] = sext(f[0], W);
        f[4] = sext(f[4], W);
        f[8] = sext(f[8], W);
    }
    const bool transformed = (mode != 9 && mode != 10);
    if (transformed || isSigned) {
#pragma unroll
        for (int i = 1; i < 4; i++) {
            if (i < ne) {
                f[i] = sext(f[i], dR);
                f[4 + i] = sext(f[4 + i], dG);
                f[8 + i] = sext(f[8 + i], dB);
            }
        }
    }
    if (transformed) {
        const int mask = (1 << W) - 1;
#pragma unroll
        for (int i = 1; i < 4; i++) {
            if (i < ne) {
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    int v = (f[c * 4 + i] + f[c * 4]) & mask;
                    if (isSigned) v = sext(v, W);
                    f[c * 4 + i] = v;
                }
            }
        }
    }
    int ep[3][4];
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int i = 0; i < 4; i++) ep[c][i] = (i < ne) ? bc6h_unquantize(f[c * 4 + i], W, isSigned) : 0;

    const uint32_t shape = two ? kShape2[part] : 0;
    const uint32_t anchor = two ? kAnchor2[part] : 0xFF;
    const uint32_t ib = two ? 3 : 4;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t n = ib - ((i == 0 || (uint32_t)i == anchor) ? 1u : 0u);
        uint32_t idx = bs.read(n);
        int w = (int)bc7_weight(ib, idx); // aWeight3 / aWeight4 == the BC7 3/4-bit weights
        int s = (shape >> i) & 1;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            int a = s ? ep[c][2] : ep[c][0];
            int b = s ? ep[c][3] : ep[c][1];
            int v = (a * (64 - w) + b * w + 32) >> 6;
            out[i * 3 + c] = (uint16_t)bc6h_finish(v, isSigned);
        }
    }
}

// ---------------------------------------------------------------- kernels
// One thread = one block.  VEC = destination rows may be written with vector stores.
template <int CODEC, bool VEC>
__global__ void __launch_bounds__(256) decode_kernel(const DecodeParams P)
{
    __shared__ uint16_t s_lay[CODEC == C_BC6H ? 14 * BC6H_MAX_TOKENS : 1];
    if (CODEC == C_BC6H) {
        for (int i = threadIdx.x; i < 14 * BC6H_MAX_TOKENS; i += blockDim.x) s_lay[i] = g_bc6h_layout[i];
        __syncthreads();
    }
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < P.nblocks; b += gridDim.x * blockDim.x) {
        const uint32_t by = b / P.nbx, bx = b - by * P.nbx;
        uint8_t* drow = P.dst + (size_t)by * 4 * P.dstRowStride;

        if (CODEC == C_BC1 || CODEC == C_BC2 || CODEC == C_BC3 || CODEC == C_BC7) {
            uint32_t px[16];
            if (CODEC == C_BC1) {
                uint2 blk = __ldg((const uint2*)P.src + b);
                decode_color_block(blk, false, px);
            } else if (CODEC == C_BC7) {
                uint4 blk = __ldg((const uint4*)P.src + b);
                decode_bc7_block(blk, px);
            } else {
                uint4 blk = __ldg((const uint4*)P.src + b);
                decode_color_block(make_uint2(blk.z, blk.w), true, px);
                if (CODEC == C_BC2) {
                    // bcdec__sharp_alpha_block (bcdec.h:153-168): 4-bit alpha * 17
                    uint64_t a = ((uint64_t)blk.y << 32) | blk.x;
#pragma unroll
                    for (int i = 0; i < 16; i++) px[i] = (px[i] & 0x00FFFFFFu) | ((((uint32_t)(a >> (4 * i)) & 0xF) * 17u) << 24);
                } else {
                    uint32_t al[16];
                    decode_smooth_alpha(make_uint2(blk.x, blk.y), al);
#pragma unroll
                    for (int i = 0; i < 16; i++) px[i] = (px[i] & 0x00FFFFFFu) | (al[i] << 24);
                }
            }
            uint8_t* d = drow + (size_t)bx * 16;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                if (VEC) *(uint4*)(d + r * P.dstRowStride) = make_uint4(px[r * 4], px[r * 4 + 1], px[r * 4 + 2], px[r * 4 + 3]);
                else {
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        uint8_t* q = d + r * P.dstRowStride + j * 4;
                        uint32_t v = px[r * 4 + j];
                        q[0] = v; q[1] = v >> 8; q[2] = v >> 16; q[3] = v >> 24;
                    }
                }
            }
        } else if (CODEC == C_BC4) {
            uint32_t v[16];
            decode_smooth_alpha(__ldg((const uint2*)P.src + b), v);
            uint8_t* d = drow + (size_t)bx * 4;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                uint32_t w = v[r * 4] | (v[r * 4 + 1] << 8) | (v[r * 4 + 2] << 16) | (v[r * 4 + 3] << 24);
                if (VEC) *(uint32_t*)(d + r * P.dstRowStride) = w;
                else {
                    uint8_t* q = d + r * P.dstRowStride;
                    q[0] = w; q[1] = w >> 8; q[2] = w >> 16; q[3] = w >> 24;
                }
            }
        } else if (CODEC == C_BC5) {
            uint4 blk = __ldg((const uint4*)P.src + b);
            uint32_t v0[16], v1[16];
            decode_smooth_alpha(make_uint2(blk.x, blk.y), v0);
            decode_smooth_alpha(make_uint2(blk.z, blk.w), v1);
            uint8_t* d = drow + (size_t)bx * 8;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                uint32_t w0 = v0[r * 4] | (v1[r * 4] << 8) | (v0[r * 4 + 1] << 16) | (v1[r * 4 + 1] << 24);
                uint32_t w1 = v0[r * 4 + 2] | (v1[r * 4 + 2] << 8) | (v0[r * 4 + 3] << 16) | (v1[r * 4 + 3] << 24);
                if (VEC) *(uint2*)(d + r * P.dstRowStride) = make_uint2(w0, w1);
                else {
                    uint8_t* q = d + r * P.dstRowStride;
                    q[0] = w0; q[1] = w0 >> 8; q[2] = w0 >> 16; q[3] = w0 >> 24;
                    q[4] = w1; q[5] = w1 >> 8; q[6] = w1 >> 16; q[7] = w1 >> 24;
                }
            }
        } else { // BC6H -> RGB16F / RGB32F
            uint16_t h[48];
            decode_bc6h_block(__ldg((const uint4*)P.src + b), P.is_signed != 0, s_lay, h);
            if (!P.f32out) {
                uint8_t* d = drow + (size_t)bx * 24;
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    uint8_t* q = d + r * P.dstRowStride;
                    if (VEC) {
#pragma unroll
                        for (int j = 0; j < 3; j++) {
                            const uint16_t* s = h + r * 12 + j * 4;
                            ((uint2*)q)[j] = make_uint2(s[0] | ((uint32_t)s[1] << 16), s[2] | ((uint32_t)s[3] << 16));
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 12; j++) { q[2 * j] = h[r * 12 + j] & 0xFF; q[2 * j + 1] = h[r * 12 + j] >> 8; }
                    }
                }
            } else {
                uint8_t* d = drow + (size_t)bx * 48;
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    uint8_t* q = d + r * P.dstRowStride;
                    if (VEC) {
#pragma unroll
                        for (int j = 0; j < 3; j++) {
                            const uint16_t* s = h + r * 12 + j * 4;
                            ((uint4*)q)[j] = make_uint4(half_bits_to_float_bits(s[0]), half_bits_to_float_bits(s[1]),
                                                         half_bits_to_float_bits(s[2]), half_bits_to_float_bits(s[3]));
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 12; j++) {
                            uint32_t v = half_bits_to_float_bits(h[r * 12 + j]);
                            uint8_t* qq = q + 4 * j;
                            qq[0] = v; qq[1] = v >> 8; qq[2] = v >> 16; qq[3] = v >> 24;
                        }
                    }
                }
            }
        }
    }
}


// Persistent grids: D7_WAVES x (CTAs that fit on the device at once), so every CTA is resident from the start and all of
// them finish within one tile of each other (a grid of 1.6 or 3.2 waves leaves the last partial wave half empty).
#ifndef D7_WAVES
#define D7_WAVES 1
#endif
template <class K>
static uint32_t resident_grid(K kernel, int smem, int sm_count, uint32_t nTiles, int* cache, uint32_t waves = D7_WAVES)
{
    if (*cache == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, 256, (size_t)smem) != cudaSuccess || n < 1) n = 1;
        *cache = n;
    }
    const uint32_t g = (uint32_t)sm_count * (uint32_t)*cache * waves;
    return nTiles < g ? nTiles : g;
}
// Share of a CTA's tiles that is strided rather than claimed.  Measured (BC7 / BC6H, 16384^2): all claimed 82.6 / 73.3 % of
// HBM peak, half 82.4 / 70.9, three quarters 77.0 / 66.5, all strided 72.2 / 65.0 -- so everything after the first tile is
// claimed; the knob stays for tools/build_variant.sh.
#ifndef D7_STATIC_NUM
#define D7_STATIC_NUM 0
#define D7_STATIC_DEN 1
#endif
static uint32_t static_rounds(uint32_t nTiles, uint32_t grid)
{
    const uint32_t k = (nTiles / grid) * D7_STATIC_NUM / D7_STATIC_DEN;
    return k < 1 ? 1 : k; // tile 0 of a CTA is always its blockIdx
}


// ---------------------------------------------------------------- lean BC1..BC5 kernels (bcn_decode_lean.cuh)
// One thread = one 16-byte-per-row work item: 1 block (BC1/2/3), 2 adjacent blocks (BC5) or 4 adjacent blocks (BC4), so
// every row store is one 128-bit streaming store and a warp's 32 stores form one 512-byte run.  Used when the
// destination is 16-byte aligned and the block row divides into whole items; decode_kernel below is the general path.
template <int CODEC>
__global__ void __launch_bounds__(256) decode_lean_kernel(const DecodeParams P, const uint32_t itemsPerRow, const uint32_t nItems)
{
    constexpr int K = (CODEC == C_BC4) ? 4 : (CODEC == C_BC5) ? 2 : 1;      // blocks per item
    constexpr int IB = K * ((CODEC == C_BC1 || CODEC == C_BC4) ? 8 : 16);   // source bytes per item
    const uint32_t stride = gridDim.x * 256u;
    uint32_t item = blockIdx.x * 256u + threadIdx.x;
    uint32_t by = item / itemsPerRow, ix = item - by * itemsPerRow;
    const uint32_t dby = stride / itemsPerRow, dix = stride - dby * itemsPerRow;
    for (; item < nItems; item += stride) {
        uint4 o[4];
        if (CODEC == C_BC1) {
            const uint2 b = __ldcs((const uint2*)(P.src + (size_t)item * IB));
            uint32_t px[16];
            lean::decode_bc1(b.x, b.y, px);
#pragma unroll
            for (int r = 0; r < 4; r++) o[r] = make_uint4(px[4 * r], px[4 * r + 1], px[4 * r + 2], px[4 * r + 3]);
        } else if (CODEC == C_BC2 || CODEC == C_BC3) {
            const uint4 b = __ldcs((const uint4*)(P.src + (size_t)item * IB));
            uint32_t px[16];
            if (CODEC == C_BC2) lean::decode_bc2(b.x, b.y, b.z, b.w, px);
            else lean::decode_bc3(b.x, b.y, b.z, b.w, px);
#pragma unroll
            for (int r = 0; r < 4; r++) o[r] = make_uint4(px[4 * r], px[4 * r + 1], px[4 * r + 2], px[4 * r + 3]);
        } else if (CODEC == C_BC4) {
            const uint4 b0 = __ldcs((const uint4*)(P.src + (size_t)item * IB)), b1 = __ldcs((const uint4*)(P.src + (size_t)item * IB) + 1);
            uint32_t a[4][4];
            lean::alpha_rows(b0.x, b0.y, a[0]);
            lean::alpha_rows(b0.z, b0.w, a[1]);
            lean::alpha_rows(b1.x, b1.y, a[2]);
            lean::alpha_rows(b1.z, b1.w, a[3]);
#pragma unroll
            for (int r = 0; r < 4; r++) o[r] = make_uint4(a[0][r], a[1][r], a[2][r], a[3][r]);
        } else { // BC5
            const uint4 b0 = __ldcs((const uint4*)(P.src + (size_t)item * IB)), b1 = __ldcs((const uint4*)(P.src + (size_t)item * IB) + 1);
            uint32_t a[4][4];
            lean::alpha_rows(b0.x, b0.y, a[0]);
            lean::alpha_rows(b0.z, b0.w, a[1]);
            lean::alpha_rows(b1.x, b1.y, a[2]);
            lean::alpha_rows(b1.z, b1.w, a[3]);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                lean::bc5_row(a[0][r], a[1][r], o[r].x, o[r].y);
                lean::bc5_row(a[2][r], a[3][r], o[r].z, o[r].w);
            }
        }
        uint8_t* d = P.dst + (size_t)by * 4 * P.dstRowStride + (size_t)ix * 16;
#pragma unroll
        for (int r = 0; r < 4; r++) __stcs((uint4*)(d + r * P.dstRowStride), o[r]);
        ix += dix;
        by += dby;
        if (ix >= itemsPerRow) { ix -= itemsPerRow; by++; }
    }
}

template <int CODEC>
static bool try_launch_lean(const DecodeParams& P, int sm_count, cudaStream_t stream, cudaError_t& err)
{
    constexpr uint32_t K = (CODEC == C_BC4) ? 4 : (CODEC == C_BC5) ? 2 : 1;
    if (((size_t)P.dst & 15) || (P.dstRowStride & 15) || ((size_t)P.src & 15) || (P.nbx % K)) return false;
    const uint32_t ipr = P.nbx / K, nItems = P.nblocks / K;
    // One item per thread, no grid cap: CTAs run in address order, so the blocks in flight at any moment cover a contiguous
    // stretch of the surface.  Measured at 16384^2 (BC1 / BC3 / BC4 / BC5, % of HBM peak): grid-stride over one resident
    // wave 79 / 81 / 79 / 81, four waves 88 / 90 / 88 / 90, sixteen 93 / 95 / 93 / 95, uncapped 97 / 99 / 93 / 97.
    (void)sm_count;
    const uint32_t ctas = (nItems + 255) / 256;
    decode_lean_kernel<CODEC><<<ctas, 256, 0, stream>>>(P, ipr, nItems);
    err = cudaGetLastError();
    return true;
}


// ---------------------------------------------------------------- lean BC7 kernel (bc7_decode_lean.cuh)
// A CTA takes a tile of 256 consecutive blocks: coalesced 16-byte loads into shared memory, a counting sort of the
// tile's blocks by mode (warp-aggregated with match_any, so a warp runs one mode's code path instead of up to eight),
// one block per thread decoded by the mode's specialised routine into a shared-memory image of the tile's four texel
// rows, and the rows leave as TMA bulk stores (cp.async.bulk shared -> global, 4 KB each) issued by one thread -- measured:
// writing the sorted blocks straight from the decoding lanes (scattered 16-byte stores) costs 10 us of 60 at 8192^2.
// Sort buffers are double-buffered by tile parity; three barriers per tile.
__device__ lean7::Tables g_bc7dec_tables;
__device__ lean6::FTab g_bc6h_ftab;
constexpr int D7_NT = 256;

// modes 0, 2, 4: the general decoder, kept out of line so that it does not set the lean kernel's register budget.
// Called only from warp-converged code in a loop of its own after the tile loop (every lane of the warp calls, `store`
// selects the lanes that hold such a block): a call from a divergent branch inside the tile loop returned to lean-path
// lanes with clobbered warp-uniform registers, and a converged call inside the loop made the loop spill.
__device__ __noinline__ void decode_bc7_general(uint4 blk, uint8_t* d, size_t rowStride, bool store)
{
    uint32_t px[16];
    decode_bc7_block(blk, px);
    if (store) {
#pragma unroll
        for (int r = 0; r < 4; r++) *(uint4*)(d + r * rowStride) = make_uint4(px[4 * r], px[4 * r + 1], px[4 * r + 2], px[4 * r + 3]);
    }
}

__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

// ROWS: every tile is one run of 256 blocks inside a block row (nbx % 256 == 0) and rows are 16-byte aligned -> bulk stores
// Tile scheduling (BC7 and BC6H kernels): the grid is one resident wave; a CTA's first tile is its blockIdx, every later
// one comes from a device counter (P.sched[0]) claimed two tiles ahead by thread 0, so the atomic's latency sits behind
// the sort and SMs that run faster (nearer memory partitions, fewer divergent warps) simply take more tiles.
// Static striding measured 5-8 % slower at one wave, and several waves of short-lived CTAs lose the prefetch.
// P.sched[1] counts finished CTAs; the last one zeroes both words for the next launch.
// The claim made at the top of tile i is for tile i + 2; its result is parked in shared memory after the sort's second
// barrier (s_next[(i + 1) & 1] = the tile after tile i + 1), so no register carries it through the decode.
// claim_issue returns a ticket, claim_value turns it into the tile index; the two are apart (top of the tile / after the
// sort's second barrier) so that the atomic's round trip is waited for at the second, not the first.
// The counter is a float (exact to 2^24 tiles = 4 G blocks; its zero is the integer zero) on purpose: an integer add of a
// constant -- atomicAdd, atom.add, atom.inc alike -- is warp-aggregated by ptxas, and the SHFL that hands the result back
// waits for the atomic on the spot.
__device__ __forceinline__ uint32_t claim_issue(const DecodeParams& P, uint32_t i, uint32_t kStatic)
{
    if (i < kStatic) return blockIdx.x + i * gridDim.x;
    uint32_t old;
    asm volatile("atom.global.add.f32 %0, [%1], 0f3F800000;" : "=r"(old) : "l"(P.sched) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t claim_value(uint32_t ticket, uint32_t i, uint32_t kStatic)
{
    return i < kStatic ? ticket : kStatic * gridDim.x + (uint32_t)__uint_as_float(ticket);
}
__device__ __forceinline__ void finish_claims(const DecodeParams& P)
{
    __threadfence();
    if (atomicAdd(P.sched + 1, 1u) == gridDim.x - 1u) {
        P.sched[0] = 0;
        P.sched[1] = 0;
    }
}
// ROWS kernels: block row of a tile = tile / (nbx / 256) as a multiply-high by ceil(2^32 / d); exact while tile * d < 2^32
// (checked by the launcher); rowMagic 0 = one tile per row
__device__ __forceinline__ uint32_t tile_row(uint32_t tile, uint32_t rowMagic) { return rowMagic ? __umulhi(tile, rowMagic) : tile; }
static uint32_t row_magic(uint32_t nbx)
{
    const uint32_t d = nbx / D7_NT;
    return d <= 1 ? 0u : (uint32_t)(0x100000000ull / d) + 1u;
}
static bool row_magic_ok(uint32_t nbx, uint32_t nTiles) { return (unsigned long long)nTiles * (nbx / D7_NT) < 0x100000000ull; }
__device__ __forceinline__ void fetch_tile(const DecodeParams& P, uint32_t tile, void* s_dst, uint64_t* bar)
{
    const uint32_t n = min((uint32_t)D7_NT, P.nblocks - tile * D7_NT) * 16u;
    mbar_arrive_expect_tx(bar, n);
    bulk_g2s(s_dst, P.src + (size_t)tile * D7_NT * 16, n, bar);
}

#ifndef D7_OCC
#define D7_OCC 5
#endif
template <bool ROWS>
__global__ void __launch_bounds__(D7_NT, D7_OCC) decode_bc7_lean_kernel(const DecodeParams P, const uint32_t nTiles, const uint32_t kStatic, const uint32_t rowMagic)
{
    __shared__ uint4 s_blk[2][D7_NT];
    __shared__ __align__(128) uint4 s_stage[4][D7_NT];
    __shared__ __align__(16) uint32_t s_pal[16 * D7_NT];
    __shared__ lean7::Tables s_T;
    __shared__ __align__(16) uint32_t s_cnt[2][12];
    __shared__ uint8_t s_order[2][D7_NT];
    __shared__ __align__(8) uint64_t s_bar[2]; // s_bar[p] completes when the bulk copy into s_blk[p] has landed
    __shared__ uint32_t s_next[2];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    {
        const uint32_t* g = (const uint32_t*)&g_bc7dec_tables;
        uint32_t* d = (uint32_t*)&s_T;
        for (uint32_t i = tid; i < sizeof(lean7::Tables) / 4; i += D7_NT) d[i] = g[i];
    }
    if (tid < 24) (&s_cnt[0][0])[tid] = 0;
    if (tid == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        fence_mbar_init();
        fetch_tile(P, blockIdx.x, &s_blk[0][0], &s_bar[0]); // grid <= nTiles
        s_next[0] = claim_value(claim_issue(P, 1, kStatic), 1, kStatic);
    }
    const lean7::Pal pal{(char*)s_pal, tid * 4u};
    uint32_t tile = blockIdx.x, it = 0; // it: tiles done by this CTA; buffer parity it & 1, mbarrier phase (it >> 1) & 1
    __syncthreads();
    // Sort, decode and store the lean-mode blocks of one tile; returns how many blocks of modes 0, 2, 4 it holds.
    auto lean_tile = [&](const uint32_t tile, const uint32_t it) -> uint32_t {
        const uint32_t par = it & 1u;
        const uint32_t t0 = tile * D7_NT;
        const bool valid = t0 + tid < P.nblocks;
        uint32_t claimed = 0;
        if (tid == 0) {
            // the tile after this one is fetched by a bulk copy while this one is sorted and decoded (no registers held,
            // no load latency in front of the sort); s_blk[par ^ 1] was last read before the previous tile's final barrier
            const uint32_t nxt = s_next[par];
            if (nxt < nTiles) {
                fetch_tile(P, nxt, &s_blk[par ^ 1u][0], &s_bar[par ^ 1u]);
                claimed = claim_issue(P, it + 2, kStatic);
            }
        }
        mbar_wait(&s_bar[par], (it >> 1) & 1u);
        uint32_t mode = 9; // 9 = no block
        if (valid) {
            // sort key: modes 0, 2, 4 (general decoder) first, then 1, 3, 5, 6, 7 and the reserved mode
            const uint32_t low = s_blk[par][tid].x & 0xFFu;
            mode = low ? (0x76524130u >> (4u * ((uint32_t)__ffs(low) - 1u))) & 15u : 8u;
        }
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, mode);
        const uint32_t leader = (uint32_t)__ffs(peers) - 1u, rank = (uint32_t)__popc(peers & ((1u << lane) - 1u));
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&s_cnt[par][mode], (uint32_t)__popc(peers));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        // the previous tile's bulk stores must have read s_stage before this tile's decode overwrites it (two barriers on)
        if (ROWS && tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();
        if (valid) {
            const uint4 c0 = *(const uint4*)&s_cnt[par][0], c1 = *(const uint4*)&s_cnt[par][4];
            const uint32_t c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
            uint32_t start = base + rank;
#pragma unroll
            for (uint32_t m = 0; m < 8; m++) start += (m < mode) ? c[m] : 0u;
            s_order[par][start] = (uint8_t)tid;
        }
        __syncthreads();
        if (tid == 0) s_next[par ^ 1u] = s_next[par] < nTiles ? claim_value(claimed, it + 2, kStatic) : nTiles;
        const uint32_t nGeneral = s_cnt[par][0] + s_cnt[par][1] + s_cnt[par][2]; // sorted slots [0, nGeneral)
        if (tid >= 3 && tid < 12) s_cnt[par][tid] = 0; // next used two tiles from now, behind barriers; [0..2] after the next barrier
        const uint32_t nvalid = min((uint32_t)D7_NT, P.nblocks - t0);
        if (tid >= nGeneral && tid < nvalid) {
            const uint32_t k = s_order[par][tid];
            const uint4 blk = s_blk[par][k];
            uint32_t px[16];
            lean7::decode_bc7(blk.x, blk.y, blk.z, blk.w, s_T, pal, px);
#pragma unroll
            for (int r = 0; r < 4; r++) s_stage[r][k] = make_uint4(px[4 * r], px[4 * r + 1], px[4 * r + 2], px[4 * r + 3]);
        }
        __syncthreads();
        if (tid < 3) s_cnt[par][tid] = 0;
        if (ROWS) {
            if (tid == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                const uint32_t by = tile_row(tile, rowMagic), bx = t0 - by * P.nbx;
                uint8_t* d = P.dst + (size_t)by * 4 * P.dstRowStride + (size_t)bx * 16;
#pragma unroll
                for (int r = 0; r < 4; r++) bulk_s2g(d + r * P.dstRowStride, &s_stage[r][0], nvalid * 16u);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        } else if (valid) {
            const uint32_t b = t0 + tid, by = b / P.nbx, bx = b - by * P.nbx;
            uint8_t* d = P.dst + (size_t)by * 4 * P.dstRowStride + (size_t)bx * 16;
#pragma unroll
            for (int r = 0; r < 4; r++) __stcs((uint4*)(d + r * P.dstRowStride), s_stage[r][tid]);
        }
        return nGeneral;
    };
    // Hot loop: tiles of the modes encoders emit.  It ends at the first tile that also holds blocks of modes 0, 2, 4 (none
    // in anything the reference's encoder presets produce); from there the CTA runs the second loop, which adds the
    // general decoder after each tile.  The two loops are separate on purpose: a __noinline__ call from a divergent
    // branch returned to the other lanes with clobbered warp-uniform registers, and a converged call anywhere inside the
    // hot loop's nest made ptxas keep the sixteen texels of every block in local memory.
    bool general = false;
    while (tile < nTiles) {
        if (lean_tile(tile, it)) { // uniform
            general = true;
            break;
        }
        tile = s_next[it & 1u]; // rewritten two tiles from now
        it++;
    }
    if (general) {
        for (;;) {
            // the tile's own stores (which carried stale texels for these blocks) are complete first
            if (ROWS && tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            __syncthreads();
            const uint32_t b = tile * D7_NT + tid;
            const bool valid = b < P.nblocks;
            uint4 blk = make_uint4(0, 0, 0, 0);
            if (valid) blk = s_blk[it & 1u][tid]; // still there: the next bulk copy into it is issued by the next tile
            const uint32_t low = blk.x & 0xFFu;
            const uint32_t m = low ? (uint32_t)__ffs(low) - 1u : 8u;
            const bool mine = valid && (m == 0 || m == 2 || m == 4);
            if (__any_sync(0xFFFFFFFFu, mine)) { // every lane calls
                const uint32_t by = b / P.nbx, bx = b - by * P.nbx;
                decode_bc7_general(blk, P.dst + (size_t)by * 4 * P.dstRowStride + (size_t)bx * 16, P.dstRowStride, mine);
            }
            __syncthreads();
            tile = s_next[it & 1u];
            it++;
            if (tile >= nTiles) break;
            lean_tile(tile, it);
        }
    }
    if (tid == 0) {
        if (ROWS) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); // stores complete, not only read
        finish_claims(P);
    }
}

// ---------------------------------------------------------------- lean BC6H kernel (bc6h_decode_lean.cuh)
// Same shape as the BC7 kernel: 256-block tiles, counting sort by mode (14 modes + reserved), one block per thread through
// the mode's specialised routine, texel rows staged in shared memory (RGB16F: 24 B per block row, RGB32F: 48 B) and written
// with TMA bulk stores.
#ifndef BC6D_OCC
#define BC6D_OCC 5
#endif
template <bool F32>
struct Bc6hSink {
    uint8_t* col; // this block's first row in the stage
    __device__ __forceinline__ void row(int r, const uint32_t w[6]) const
    {
        if (!F32) {
            uint2* d = (uint2*)(col + r * (D7_NT * 24));
            d[0] = make_uint2(w[0], w[1]);
            d[1] = make_uint2(w[2], w[3]);
            d[2] = make_uint2(w[4], w[5]);
        } else {
            // exact half -> float (bcdec__half_to_float_quick, bcdec.h:348-373; decoder outputs are always finite)
            float4* d = (float4*)(col + r * (D7_NT * 48));
#pragma unroll
            for (int j = 0; j < 3; j++) {
                const float2 a = __half22float2(*(const __half2*)&w[2 * j]), b = __half22float2(*(const __half2*)&w[2 * j + 1]);
                d[j] = make_float4(a.x, a.y, b.x, b.y);
            }
        }
    }
};

template <bool F32, bool SIGNED, bool ROWS>
__global__ void __launch_bounds__(D7_NT, BC6D_OCC) decode_bc6h_lean_kernel(const DecodeParams P, const uint32_t nTiles, const uint32_t kStatic, const uint32_t rowMagic)
{
    constexpr int RB = F32 ? 48 : 24; // bytes of one block row
    __shared__ uint4 s_blk[2][D7_NT];
    extern __shared__ __align__(128) uint8_t s_stage[]; // 4 * D7_NT * RB bytes (dynamic: 48 KB for RGB32F)
    __shared__ lean7::Tables s_T;
    __shared__ __align__(16) uint32_t s_cnt[2][16];
    __shared__ uint8_t s_order[2][D7_NT];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    __shared__ lean6::FTab s_F;
    {
        const uint32_t* g = (const uint32_t*)&g_bc7dec_tables;
        uint32_t* d = (uint32_t*)&s_T;
        for (uint32_t i = tid; i < sizeof(lean7::Tables) / 4; i += D7_NT) d[i] = g[i];
        const uint32_t* gf = (const uint32_t*)&g_bc6h_ftab;
        uint32_t* df = (uint32_t*)&s_F;
        for (uint32_t i = tid; i < sizeof(lean6::FTab) / 4; i += D7_NT) df[i] = gf[i];
    }
    if (tid < 32) (&s_cnt[0][0])[tid] = 0;
    __shared__ __align__(8) uint64_t s_bar[2]; // as in the BC7 kernel: bulk-copied block tiles, tiles claimed from P.sched
    __shared__ uint32_t s_next[2];
    if (tid == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        fence_mbar_init();
        fetch_tile(P, blockIdx.x, &s_blk[0][0], &s_bar[0]); // grid <= nTiles
        s_next[0] = claim_value(claim_issue(P, 1, kStatic), 1, kStatic);
    }
    uint32_t tile = blockIdx.x, it = 0;
    __syncthreads();
    while (tile < nTiles) {
        const uint32_t par = it & 1u;
        const uint32_t t0 = tile * D7_NT;
        const bool valid = t0 + tid < P.nblocks;
        uint32_t claimed = 0;
        if (tid == 0) {
            const uint32_t nxt = s_next[par];
            if (nxt < nTiles) {
                fetch_tile(P, nxt, &s_blk[par ^ 1u][0], &s_bar[par ^ 1u]);
                claimed = claim_issue(P, it + 2, kStatic);
            }
        }
        mbar_wait(&s_bar[par], (it >> 1) & 1u);
        uint32_t mode = 15; // 15 = no block
        if (valid) mode = lean6::mode_of(s_blk[par][tid].x);
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, mode);
        const uint32_t leader = (uint32_t)__ffs(peers) - 1u, rank = (uint32_t)__popc(peers & ((1u << lane) - 1u));
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&s_cnt[par][mode], (uint32_t)__popc(peers));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (ROWS && tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();
        {
            // exclusive prefix of the sixteen counts, once per warp: lane m scans, every lane picks its mode's with a shuffle
            const uint32_t cnt = s_cnt[par][lane & 15u];
            uint32_t inc = cnt;
#pragma unroll
            for (int d = 1; d < 16; d <<= 1) {
                const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, inc, d, 16);
                if ((lane & 15u) >= (uint32_t)d) inc += up;
            }
            const uint32_t pre = __shfl_sync(0xFFFFFFFFu, inc - cnt, mode);
            if (valid) s_order[par][pre + base + rank] = (uint8_t)tid;
        }
        __syncthreads();
        if (tid < 16) s_cnt[par][tid] = 0;
        if (tid == 0) s_next[par ^ 1u] = s_next[par] < nTiles ? claim_value(claimed, it + 2, kStatic) : nTiles;
        const uint32_t nvalid = min((uint32_t)D7_NT, P.nblocks - t0);
        if (tid < nvalid) {
            const uint32_t k = s_order[par][tid];
            const uint4 blk = s_blk[par][k];
            Bc6hSink<F32> sink{s_stage + k * RB};
            lean6::decode_bc6h<SIGNED>(blk.x, blk.y, blk.z, blk.w, s_T, s_F, sink);
        }
        __syncthreads();
        if (ROWS) {
            if (tid == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                const uint32_t by0 = tile_row(tile, rowMagic), bx0 = t0 - by0 * P.nbx;
                uint8_t* d = P.dst + (size_t)by0 * 4 * P.dstRowStride + (size_t)bx0 * RB;
#pragma unroll
                for (int r = 0; r < 4; r++) bulk_s2g(d + r * P.dstRowStride, s_stage + r * (D7_NT * RB), nvalid * RB);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        } else if (valid) {
            const uint32_t b = t0 + tid, by = b / P.nbx, bx = b - by * P.nbx;
            uint8_t* d = P.dst + (size_t)by * 4 * P.dstRowStride + (size_t)bx * RB;
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int j = 0; j < RB / 8; j++)
                    __stcs((uint2*)(d + r * P.dstRowStride) + j, *(const uint2*)(s_stage + r * (D7_NT * RB) + tid * RB + j * 8));
        }
        tile = s_next[par];
        it++;
    }
    if (tid == 0) {
        if (ROWS) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        finish_claims(P);
    }
}

template <bool F32, bool SIGNED>
static cudaError_t launch_bc6h_lean(const DecodeParams& P, int sm_count, uint32_t nTiles, bool rows, cudaStream_t stream)
{
    constexpr int stage = 4 * D7_NT * (F32 ? 48 : 24);
    static std::atomic<uint64_t> attr_set[2]; // bit per device: function attributes live in the device's context
    static int occ[2] = {0, 0};
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    if (!(attr_set[rows].load(std::memory_order_acquire) & bit)) {
        cudaError_t e = rows ? cudaFuncSetAttribute(decode_bc6h_lean_kernel<F32, SIGNED, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, stage)
                             : cudaFuncSetAttribute(decode_bc6h_lean_kernel<F32, SIGNED, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, stage);
        if (e != cudaSuccess) return e;
        attr_set[rows].fetch_or(bit, std::memory_order_release);
    }
    if (rows) {
        const uint32_t grid = resident_grid(decode_bc6h_lean_kernel<F32, SIGNED, true>, stage, sm_count, nTiles, &occ[1]);
        decode_bc6h_lean_kernel<F32, SIGNED, true><<<grid, D7_NT, stage, stream>>>(P, nTiles, static_rounds(nTiles, grid), row_magic(P.nbx));
    } else {
        const uint32_t grid = resident_grid(decode_bc6h_lean_kernel<F32, SIGNED, false>, stage, sm_count, nTiles, &occ[0]);
        decode_bc6h_lean_kernel<F32, SIGNED, false><<<grid, D7_NT, stage, stream>>>(P, nTiles, static_rounds(nTiles, grid), row_magic(P.nbx));
    }
    return cudaGetLastError();
}

cudaError_t bc7dec_upload_tables()
{
    static lean7::Tables H;
    static const uint16_t shape2[64] = {
        0xCCCC, 0x8888, 0xEEEE, 0xECC8, 0xC880, 0xFEEC, 0xFEC8, 0xEC80, 0xC800, 0xFFEC, 0xFE80, 0xE800, 0xFFE8, 0xFF00, 0xFFF0, 0xF000,
        0xF710, 0x008E, 0x7100, 0x08CE, 0x008C, 0x7310, 0x3100, 0x8CCE, 0x088C, 0x3110, 0x6666, 0x366C, 0x17E8, 0x0FF0, 0x718E, 0x399C,
        0xAAAA, 0xF0F0, 0x5A5A, 0x33CC, 0x3C3C, 0x55AA, 0x9696, 0xA55A, 0x73CE, 0x13C8, 0x324C, 0x3BDC, 0x6996, 0xC33C, 0x9966, 0x0660,
        0x0272, 0x04E4, 0x4E40, 0x2720, 0xC936, 0x936C, 0x39C6, 0x639C, 0x9336, 0x9CC6, 0x817E, 0xE718, 0xCCF0, 0x0FCC, 0x7744, 0xEE22};
    static const uint8_t anchor2[64] = {15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 2, 8, 2, 2, 8, 8, 15, 2, 8, 2, 2, 8, 8, 2, 2,
                                        15, 15, 6, 8, 2, 8, 15, 15, 2, 8, 2, 2, 2, 15, 15, 6, 6, 2, 6, 8, 15, 15, 2, 2, 15, 15, 15, 15, 15, 2, 2, 15};
    lean7::fill_tables(&H, shape2, anchor2);
    static lean6::FTab F;
    lean6::fill_ftab(&F, shape2);
    cudaError_t e = cudaMemcpyToSymbol(g_bc6h_ftab, &F, sizeof(F));
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbol(g_bc7dec_tables, &H, sizeof(H));
}

template <int CODEC>
static cudaError_t launch_decode_t(const DecodeParams& P, int sm_count, cudaStream_t stream)
{
    // required alignment of a destination row segment for the vector path
    if constexpr (CODEC == C_BC1 || CODEC == C_BC2 || CODEC == C_BC3 || CODEC == C_BC4 || CODEC == C_BC5) {
        cudaError_t e;
        if (try_launch_lean<CODEC>(P, sm_count, stream, e)) return e;
    }
    if constexpr (CODEC == C_BC7) {
        if (!((size_t)P.dst & 15) && !(P.dstRowStride & 15) && !((size_t)P.src & 15)) {
            const uint32_t nTiles = (P.nblocks + D7_NT - 1) / D7_NT;
            static int occ[2] = {0, 0};
            if (P.nbx % D7_NT == 0 && row_magic_ok(P.nbx, nTiles)) {
                const uint32_t grid = resident_grid(decode_bc7_lean_kernel<true>, 0, sm_count, nTiles, &occ[1]);
                decode_bc7_lean_kernel<true><<<grid, D7_NT, 0, stream>>>(P, nTiles, static_rounds(nTiles, grid), row_magic(P.nbx));
            } else {
                const uint32_t grid = resident_grid(decode_bc7_lean_kernel<false>, 0, sm_count, nTiles, &occ[0]);
                decode_bc7_lean_kernel<false><<<grid, D7_NT, 0, stream>>>(P, nTiles, static_rounds(nTiles, grid), row_magic(P.nbx));
            }
            return cudaGetLastError();
        }
    }
    if constexpr (CODEC == C_BC6H) {
        if (!((size_t)P.dst & 7) && !(P.dstRowStride & 7) && !((size_t)P.src & 15)) {
            const uint32_t nTiles = (P.nblocks + D7_NT - 1) / D7_NT;
            const bool rows = (P.nbx % D7_NT == 0) && row_magic_ok(P.nbx, nTiles) && !((size_t)P.dst & 15) && !(P.dstRowStride & 15);
            if (P.f32out) return P.is_signed ? launch_bc6h_lean<true, true>(P, sm_count, nTiles, rows, stream) : launch_bc6h_lean<true, false>(P, sm_count, nTiles, rows, stream);
            return P.is_signed ? launch_bc6h_lean<false, true>(P, sm_count, nTiles, rows, stream) : launch_bc6h_lean<false, false>(P, sm_count, nTiles, rows, stream);
        }
    }
    size_t need = (CODEC == C_BC4) ? 4 : (CODEC == C_BC5) ? 8 : (CODEC == C_BC6H) ? (P.f32out ? 16 : 8) : 16;
    bool vec = (((size_t)P.dst % need) == 0) && ((P.dstRowStride % need) == 0);
    uint32_t ctas = (P.nblocks + 255) / 256;
    uint32_t cap = (uint32_t)sm_count * 8 * 4; // a few waves of 8 CTAs/SM; grid-stride beyond that
    if (ctas > cap) ctas = cap;
    if (ctas == 0) return cudaSuccess;
    if (vec) decode_kernel<CODEC, true><<<ctas, 256, 0, stream>>>(P);
    else decode_kernel<CODEC, false><<<ctas, 256, 0, stream>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_decode(const DecodeParams& P, int sm_count, cudaStream_t stream)
{
    switch (P.codec) {
    case C_BC1: return launch_decode_t<C_BC1>(P, sm_count, stream);
    case C_BC2: return launch_decode_t<C_BC2>(P, sm_count, stream);
    case C_BC3: return launch_decode_t<C_BC3>(P, sm_count, stream);
    case C_BC4: return launch_decode_t<C_BC4>(P, sm_count, stream);
    case C_BC5: return launch_decode_t<C_BC5>(P, sm_count, stream);
    case C_BC6H: return launch_decode_t<C_BC6H>(P, sm_count, stream);
    case C_BC7: return launch_decode_t<C_BC7>(P, sm_count, stream);
    }
    return cudaErrorInvalidValue;
}

} // namespace hkbcn
