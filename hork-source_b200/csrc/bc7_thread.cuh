// bc7_thread.cuh -- BC7 block encoder, device side (sm_100a): one thread per 4x4 block.
//
// Replaces bc7enc_compress_block with perceptual weights at uber level 4, the configuration
// Hk::TextureBlockCompression::CompressBC7 bakes in, and at levels 0..3 for Encode_BC7(.., Level) (Hork/Image/ImageEncoders.cpp:59-63, 207-212,
// 524-557 -> ThirdParty/bc7enc_rdo/bc7enc.cpp:2402-2438).  Same mode set (opaque: 6 then 1; any
// alpha < 255: 6, 5, 7), same candidate generators (principal axis, least-squares refits, selector
// perturbation and rescale sweep, single-colour tables), same integer YCbCr error metric, and every
// float expression evaluated in the reference's operand order with contraction off (--fmad=false),
// so the output is meant to be byte-identical to the reference's; the tests hold it to that.
//
// Why a thread per block (the first version used 16 lanes per block, profiles/r1_bc7_lane16_*):
// everything that is uniform over a block -- endpoint quantisation, palette construction, the 2x2
// least-squares solve -- was recomputed by all 16 lanes, and every per-block sum was a REDUX under a
// partial mask, which ptxas serialises per group behind WARPSYNC/ENDCOLLECTIVE.  Here the uniform
// work is done once, sums are plain loops over the 16 texels, and no cross-lane traffic exists.
// The two subsets of modes 1 / 7 are compressed concurrently inside the thread (NS = 2): one pass
// over the texels serves both, each texel picking its subset's palette.
//
// Texels live in shared memory, [16][NT] words, so lane t reads word i * NT + t (conflict-free).
#pragma once
#ifndef HKBCN_HOST_EMU // tools/bc7_host_check.cpp compiles this header for the CPU to debug against the reference
#include <cuda_runtime.h>
#endif
#include <stdint.h>
#include "bcn_tables.cuh"

namespace hkbcn {
namespace bc7 {

// error weights after the perceptual scaling of bc7enc.cpp:2412-2417 (checked on the host at init)
constexpr int WY = 512, WCR = 103, WCB = 18, WA = 128;

// lookup tables built on the host by the loops of bc7enc_compress_block_init (bc7enc.cpp:124-285)
// plus the least-squares weight tables (bc7enc.cpp:52-56: w*w, (1-w)*w, (1-w)*(1-w), w to 6 decimals)
struct Tables {
    uint32_t opt1[256][2];    // mode 1 single-colour: [c][pbit]      = err | lo << 16 | hi << 24
    uint32_t opt7[256][2][2]; // mode 7 single-colour: [c][hp][lp]
    float    mid1[64][2];     // endpoint quantisation midpoints, mode 1 [i][p]
    float    mid7[32][2];     // mode 7
    float    mid5[128];       // mode 5
    float    wx4[16][4];      // g_bc7_weights4x
    float    wx3[8][4];       // g_bc7_weights3x
    float    wx2[4][4];       // g_bc7_weights2x
};

// Partition scan order of estimate_partition (bc7enc.cpp:1767-1777) and its predictor masks (:1711-1748).
__device__ const uint8_t kPartOrder[64] = {
    0, 13, 1, 2, 15, 14, 10, 16, 3, 23, 26, 6, 7, 21, 19, 29, 8, 4, 9, 20, 5, 31, 22, 17, 18, 11, 12, 30, 24, 25, 28, 27,
    32, 33, 34, 45, 46, 51, 49, 50, 48, 38, 39, 37, 53, 52, 54, 36, 57, 58, 55, 41, 40, 42, 43, 59, 44, 56, 47, 35, 60, 63, 62, 61};
#define HKB(x) (1u << (x))
__device__ const uint32_t kPartPredictors[35] = {
    ~0u, ~0u, ~0u, ~0u, ~0u, HKB(1) | HKB(2) | HKB(8), HKB(1) | HKB(3) | HKB(7), ~0u, ~0u, HKB(2) | HKB(8) | HKB(16), HKB(7) | HKB(3) | HKB(15), ~0u,
    HKB(8) | HKB(14) | HKB(16), HKB(7) | HKB(14) | HKB(15), ~0u, ~0u, ~0u, ~0u, HKB(14) | HKB(15), HKB(16) | HKB(22) | HKB(14), HKB(17) | HKB(24) | HKB(14),
    HKB(2) | HKB(14) | HKB(15) | HKB(1), ~0u, HKB(1) | HKB(3) | HKB(14) | HKB(16) | HKB(22), ~0u, HKB(1) | HKB(2) | HKB(15) | HKB(17) | HKB(24),
    HKB(1) | HKB(3) | HKB(22), ~0u, ~0u, ~0u, HKB(14) | HKB(15) | HKB(16) | HKB(17), ~0u, ~0u, HKB(1) | HKB(2) | HKB(3) | HKB(27) | HKB(4) | HKB(24),
    HKB(14) | HKB(15) | HKB(16) | HKB(11) | HKB(17) | HKB(27)};
#undef HKB

__device__ __forceinline__ float sat(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// selector weights (bc7enc.cpp:48-50): {0,21,43,64}, {0,9,...,64}, {0,4,...,64}
template <int N> __device__ __forceinline__ constexpr int weight_of(int s)
{
    return N == 4 ? (s * 64 + 1) / 3 : N == 8 ? (s * 64 + 3) / 7 : (s * 64 + 7) / 15;
}
template <int N> __device__ __forceinline__ const float* wx_row(const Tables& T, int s)
{
    return N == 4 ? T.wx2[s] : N == 8 ? T.wx3[s] : T.wx4[s];
}

// per-mode constants (color_cell_compressor_params as the handlers fill it, bc7enc.cpp:2143-2153, 2315-2345, 2041-2049)
template <int MODE> struct Cfg;
template <> struct Cfg<6> { static constexpr int N = 16, CB = 7, NS = 1; static constexpr bool PB = true, SHARE = false; };
template <> struct Cfg<1> { static constexpr int N = 8, CB = 6, NS = 2; static constexpr bool PB = true, SHARE = true; };
template <> struct Cfg<5> { static constexpr int N = 4, CB = 7, NS = 1; static constexpr bool PB = false, SHARE = false; };
template <> struct Cfg<7> { static constexpr int N = 4, CB = 5, NS = 2; static constexpr bool PB = true, SHARE = false; };

// scale_color (bc7enc.cpp:487-503): n-bit component -> 8 bit
template <int NB> __device__ __forceinline__ int scale_comp(int v)
{
    v <<= (8 - NB);
    return v | (v >> NB);
}

__device__ __forceinline__ void unpack4(uint32_t v, int c[4])
{
    c[0] = v & 255;
    c[1] = (v >> 8) & 255;
    c[2] = (v >> 16) & 255;
    c[3] = v >> 24;
}
__device__ __forceinline__ uint32_t pack4(const int c[4]) { return (uint32_t)c[0] | ((uint32_t)c[1] << 8) | ((uint32_t)c[2] << 16) | ((uint32_t)c[3] << 24); }

// one subset's running best (color_cell_compressor_results, bc7enc.cpp:477-485); err 0xFFFFFFFF stands for UINT64_MAX
// (a real total is at most 16 texels x 173e6 < 2^32)
struct Sub {
    uint32_t err;
    uint32_t lo, hi; // quantised endpoints, 4 x 8 bit (r | g << 8 | b << 16 | a << 24), without p-bit
    uint32_t pb;     // bit 0 = pbits[0] (low endpoint), bit 1 = pbits[1]
};

// the thread's block: texel i at px[i * stride]
struct Block {
    const uint32_t* px;
    int             stride;
    __device__ __forceinline__ uint32_t operator[](int i) const { return px[i * stride]; }
};

// compute_color_distance_rgb[a] (bc7enc.cpp:505-535), perceptual branch, of colour (r,g,b,a) against a texel given as YCbCr + alpha
template <bool ALPHA> __device__ __forceinline__ uint32_t color_err(int r, int g, int b, int a, int l2, int cr2, int cb2, int a2)
{
    const int l1 = r * 109 + g * 366 + b * 37;
    const int dl = (l1 - l2) >> 8;
    const int dcr = ((r << 9) - l1 - cr2) >> 8;
    const int dcb = ((b << 9) - l1 - cb2) >> 8;
    uint32_t e = (uint32_t)(WY * (dl * dl)) + (uint32_t)(WCR * (dcr * dcr)) + (uint32_t)(WCB * (dcb * dcb));
    if (ALPHA) {
        const int da = a - a2;
        e += (uint32_t)(WA * (da * da));
    }
    return e;
}

// First half of find_optimal_solution (bc7enc.cpp:868-1085): saturate the float endpoints xl / xh and quantise
// them to the mode's grid, choosing p-bits by endpoint reconstruction error.  Returns packed endpoints + p-bits.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void quantize_endpoints(float xl[4], float xh[4], const Tables& T, uint32_t& lo, uint32_t& hi, uint32_t& pb)
{
    using C = Cfg<MODE>;
    constexpr int totalComps = ALPHA ? 4 : 3;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        xl[c] = sat(xl[c]);
        xh[c] = sat(xh[c]);
    }
    int bmin[4], bmax[4];
    pb = 0;
    if (C::PB) {
        constexpr int iscalep = (1 << (C::CB + 1)) - 1;
        constexpr float scalep = (float)iscalep;
        if (!C::SHARE) {
            float best0 = 1e+9f, best1 = 1e+9f;
#pragma unroll
            for (int pbit = 0; pbit < 2; pbit++) {
                int qmin[4], qmax[4];
                float err0 = 0.0f, err1 = 0.0f;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (C::CB == 5) {
                        int vl = (int)(xl[c] * 31.0f);
                        vl += (xl[c] > T.mid7[vl][pbit]);
                        qmin[c] = clampi(vl * 2 + pbit, pbit, 63 - 1 + pbit);
                        int vh = (int)(xh[c] * 31.0f);
                        vh += (xh[c] > T.mid7[vh][pbit]);
                        qmax[c] = clampi(vh * 2 + pbit, pbit, 63 - 1 + pbit);
                    } else {
                        qmin[c] = clampi(((int)((xl[c] * scalep - (float)pbit) / 2.0f + .5f)) * 2 + pbit, pbit, iscalep - 1 + pbit);
                        qmax[c] = clampi(((int)((xh[c] * scalep - (float)pbit) / 2.0f + .5f)) * 2 + pbit, pbit, iscalep - 1 + pbit);
                    }
                }
#pragma unroll
                for (int c = 0; c < totalComps; c++) {
                    const float dl = (float)scale_comp<C::CB + 1>(qmin[c]) - xl[c] * 255.0f;
                    const float dh = (float)scale_comp<C::CB + 1>(qmax[c]) - xh[c] * 255.0f;
                    err0 += dl * dl;
                    err1 += dh * dh;
                }
                if (err0 < best0) {
                    best0 = err0;
                    pb = (pb & 2u) | (uint32_t)pbit;
#pragma unroll
                    for (int c = 0; c < 4; c++) bmin[c] = qmin[c] >> 1;
                }
                if (err1 < best1) {
                    best1 = err1;
                    pb = (pb & 1u) | ((uint32_t)pbit << 1);
#pragma unroll
                    for (int c = 0; c < 4; c++) bmax[c] = qmax[c] >> 1;
                }
            }
        } else {
            // endpoints share the p-bit (mode 1, bc7enc.cpp:1021-1061)
            float best = 1e+9f;
#pragma unroll
            for (int pbit = 0; pbit < 2; pbit++) {
                int qmin[4], qmax[4];
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    int vl = (int)(xl[c] * 63.0f);
                    vl += (xl[c] > T.mid1[vl][pbit]);
                    qmin[c] = clampi(vl * 2 + pbit, pbit, 127 - 1 + pbit);
                    int vh = (int)(xh[c] * 63.0f);
                    vh += (xh[c] > T.mid1[vh][pbit]);
                    qmax[c] = clampi(vh * 2 + pbit, pbit, 127 - 1 + pbit);
                }
                float err = 0.0f;
#pragma unroll
                for (int c = 0; c < totalComps; c++) {
                    const float dl = (float)scale_comp<C::CB + 1>(qmin[c]) / 255.0f - xl[c];
                    const float dh = (float)scale_comp<C::CB + 1>(qmax[c]) / 255.0f - xh[c];
                    err += dl * dl + dh * dh;
                }
                if (err < best) {
                    best = err;
                    pb = (uint32_t)pbit * 3u;
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        bmin[c] = qmin[c] >> 1;
                        bmax[c] = qmax[c] >> 1;
                    }
                }
            }
            // fixDegenerateEndpoints (bc7enc.cpp:836-866), mode 1: iscale = 63
#pragma unroll
            for (int i = 0; i < 3; i++) {
                if (bmin[i] == bmax[i] && fabsf(xl[i] - xh[i]) > 0.0f) {
                    if (bmin[i] > 31) {
                        if (bmin[i] > 0) bmin[i]--;
                        else if (bmax[i] < 63) bmax[i]++;
                    } else {
                        if (bmax[i] < 63) bmax[i]++;
                        else if (bmin[i] > 0) bmin[i]--;
                    }
                }
            }
        }
    } else {
        // mode 5 colour: 7 bits, no p-bits (bc7enc.cpp:1066-1080)
#pragma unroll
        for (int c = 0; c < 4; c++) {
            int vl = (int)(xl[c] * 127.0f);
            vl += (xl[c] > T.mid5[vl]);
            bmin[c] = clampi(vl, 0, 127);
            int vh = (int)(xh[c] * 127.0f);
            vh += (xh[c] > T.mid5[vh]);
            bmax[c] = clampi(vh, 0, 127);
        }
    }
    lo = pack4(bmin);
    hi = pack4(bmax);
}

// actual 8-bit endpoint components of a quantised pair (evaluate_solution, bc7enc.cpp:648-677)
template <int MODE> __device__ __forceinline__ void expand_endpoints(uint32_t lo, uint32_t hi, uint32_t pb, int lo8[4], int hi8[4])
{
    using C = Cfg<MODE>;
    int l[4], h[4];
    unpack4(lo, l);
    unpack4(hi, h);
    const int p0 = pb & 1, p1 = C::SHARE ? p0 : (int)((pb >> 1) & 1);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (C::PB) {
            lo8[k] = scale_comp<C::CB + 1>((l[k] << 1) | p0);
            hi8[k] = scale_comp<C::CB + 1>((h[k] << 1) | p1);
        } else {
            lo8[k] = scale_comp<C::CB>(l[k]);
            hi8[k] = scale_comp<C::CB>(h[k]);
        }
    }
}

// pack_mode1_to_one_color / pack_mode7_to_one_color (bc7enc.cpp:536-643) for the subsets flagged in `want`:
// encode a subset as the single colour col[s] through the optimal-endpoint tables, every selector = 2 (mode 1)
// or 1 (mode 7); the total error is summed over the subset's texels.
template <int MODE>
__device__ __forceinline__ void pack_one_color(const Block& blk, uint32_t mask, const Tables& T, const bool want[2], const int col[2][4], Sub out[2])
{
    int pc[2][4];
#pragma unroll
    for (int s = 0; s < 2; s++) {
        out[s].err = 0;
        const int r = col[s][0], g = col[s][1], b = col[s][2], a = col[s][3];
        if (MODE == 1) {
            uint32_t best_err = 0xFFFFFFFFu, best_p = 0;
#pragma unroll
            for (uint32_t pp = 0; pp < 2; pp++) {
                const uint32_t e = (T.opt1[r][pp] & 0xFFFF) + (T.opt1[g][pp] & 0xFFFF) + (T.opt1[b][pp] & 0xFFFF);
                if (e < best_err) { best_err = e; best_p = pp; }
            }
            const uint32_t er = T.opt1[r][best_p], eg = T.opt1[g][best_p], eb = T.opt1[b][best_p];
            out[s].lo = ((er >> 16) & 255) | (((eg >> 16) & 255) << 8) | (((eb >> 16) & 255) << 16);
            out[s].hi = (er >> 24) | ((eg >> 24) << 8) | ((eb >> 24) << 16);
            out[s].pb = best_p; // pbits[0] = best_p, pbits[1] = 0 (bc7enc.cpp:560-561)
            int l[4], h[4];
            unpack4(out[s].lo, l);
            unpack4(out[s].hi, h);
#pragma unroll
            for (int i = 0; i < 3; i++) {
                const int low = scale_comp<7>((l[i] << 1) | (int)best_p);
                const int high = scale_comp<7>((h[i] << 1) | (int)best_p);
                pc[s][i] = (low * (64 - 18) + high * 18 + 32) >> 6; // g_bc7_weights3[2]
            }
            pc[s][3] = 255;
        } else {
            uint32_t best_err = 0xFFFFFFFFu, best_p = 0;
#pragma unroll
            for (uint32_t pp = 0; pp < 4; pp++) {
                const uint32_t hp = pp >> 1, lp = pp & 1;
                const uint32_t e = (T.opt7[r][hp][lp] & 0xFFFF) + (T.opt7[g][hp][lp] & 0xFFFF) + (T.opt7[b][hp][lp] & 0xFFFF) + (T.opt7[a][hp][lp] & 0xFFFF);
                if (e < best_err) { best_err = e; best_p = pp; }
            }
            const uint32_t hp = best_p >> 1, lp = best_p & 1;
            const uint32_t er = T.opt7[r][hp][lp], eg = T.opt7[g][hp][lp], eb = T.opt7[b][hp][lp], ea = T.opt7[a][hp][lp];
            out[s].lo = ((er >> 16) & 255) | (((eg >> 16) & 255) << 8) | (((eb >> 16) & 255) << 16) | (((ea >> 16) & 255) << 24);
            out[s].hi = (er >> 24) | ((eg >> 24) << 8) | ((eb >> 24) << 16) | ((ea >> 24) << 24);
            out[s].pb = lp | (hp << 1);
            int l[4], h[4];
            unpack4(out[s].lo, l);
            unpack4(out[s].hi, h);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                // the reference widens the 6-bit values as (v << 2) | (v >> 6), i.e. without replicating the top bits
                // (bc7enc.cpp:627-631); its error for this candidate is computed on that colour, so this must be too
                int low = (l[i] << 1) | (int)lp, high = (h[i] << 1) | (int)hp;
                low = (low << 2) | (low >> 6);
                high = (high << 2) | (high >> 6);
                pc[s][i] = (low * (64 - 21) + high * 21 + 32) >> 6; // g_bc7_weights2[1]
            }
        }
    }
    if (!want[0] && !want[1]) return;
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const int s = (mask >> i) & 1;
        const int r = v & 255, g = (v >> 8) & 255, b = (v >> 16) & 255, a = v >> 24;
        const int l2 = r * 109 + g * 366 + b * 37;
        const uint32_t e = color_err<MODE == 7>(s ? pc[1][0] : pc[0][0], s ? pc[1][1] : pc[0][1], s ? pc[1][2] : pc[0][2], s ? pc[1][3] : pc[0][3], l2,
                                                (r << 9) - l2, (b << 9) - l2, a);
        out[0].err += s ? 0u : e;
        out[1].err += s ? e : 0u;
    }
}

// min / max of component c over the texels of subset s (the rarely taken fallback of the least-squares solve,
// bc7enc.cpp:334-349)
__device__ __noinline__ void subset_minmax(const Block& blk, uint32_t mask, int ns, int s, int c, int& lo_v, int& hi_v)
{
    lo_v = 255;
    hi_v = 0;
    for (int i = 0; i < 16; i++) {
        if (ns == 2 && (int)((mask >> i) & 1) != s) continue;
        const int v = (blk[i] >> (8 * c)) & 255;
        lo_v = min(lo_v, v);
        hi_v = max(hi_v, v);
    }
}

// Initial endpoints of every subset from its principal axis (bc7enc.cpp:1143-1275).
// ms = subset mean (0..255 scale), mean = the same / 255 saturated.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void pca_endpoints(const Block& blk, uint32_t mask, const float ms[][4], const float mean[][4], float xl[][4], float xh[][4])
{
    constexpr int NS = Cfg<MODE>::NS;
    float axis[NS][4];
#pragma unroll
    for (int s = 0; s < NS; s++) axis[s][0] = axis[s][1] = axis[s][2] = axis[s][3] = 0.0f;
    if (ALPHA) {
        // incremental PCA over the subset's texels in index order (bc7enc.cpp:1160-1178)
        bool started[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) started[s] = false;
#pragma unroll 1
        for (int i = 0; i < 16; i++) {
            const uint32_t v = blk[i];
            const int sb = NS == 2 ? (int)((mask >> i) & 1) : 0;
#pragma unroll
            for (int s = 0; s < NS; s++) {
                if (sb != s) continue;
                float col[4];
                col[0] = (float)(v & 255) - ms[s][0];
                col[1] = (float)((v >> 8) & 255) - ms[s][1];
                col[2] = (float)((v >> 16) & 255) - ms[s][2];
                col[3] = (float)(v >> 24) - ms[s][3];
                float nv[4];
#pragma unroll
                for (int c = 0; c < 4; c++) nv[c] = started[s] ? axis[s][c] : col[c];
                started[s] = true;
                float len = nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2] + nv[3] * nv[3];
                if (len != 0.0f) {
                    len = 1.0f / sqrtf(len);
#pragma unroll
                    for (int c = 0; c < 4; c++) nv[c] *= len;
                }
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    // axis[k] += dot(color * color[k], n)
                    const float a0 = col[0] * col[k], a1 = col[1] * col[k], a2 = col[2] * col[k], a3 = col[3] * col[k];
                    axis[s][k] += a0 * nv[0] + a1 * nv[1] + a2 * nv[2] + a3 * nv[3];
                }
            }
        }
#pragma unroll
        for (int s = 0; s < NS; s++) {
            float len = axis[s][0] * axis[s][0] + axis[s][1] * axis[s][1] + axis[s][2] * axis[s][2] + axis[s][3] * axis[s][3];
            if (len != 0.0f) {
                len = 1.0f / sqrtf(len);
#pragma unroll
                for (int c = 0; c < 4; c++) axis[s][c] *= len;
            }
        }
    } else {
        // covariance + 3 power iterations (bc7enc.cpp:1181-1219)
        float cov[NS][6];
#pragma unroll
        for (int s = 0; s < NS; s++)
#pragma unroll
            for (int k = 0; k < 6; k++) cov[s][k] = 0.0f;
#pragma unroll 1
        for (int i = 0; i < 16; i++) {
            const uint32_t v = blk[i];
            const int sb = NS == 2 ? (int)((mask >> i) & 1) : 0;
#pragma unroll
            for (int s = 0; s < NS; s++) {
                if (sb != s) continue;
                const float r = (float)(v & 255) - ms[s][0];
                const float g = (float)((v >> 8) & 255) - ms[s][1];
                const float b = (float)((v >> 16) & 255) - ms[s][2];
                cov[s][0] += r * r; cov[s][1] += r * g; cov[s][2] += r * b;
                cov[s][3] += g * g; cov[s][4] += g * b; cov[s][5] += b * b;
            }
        }
#pragma unroll
        for (int s = 0; s < NS; s++) {
            float vfr = .9f, vfg = 1.0f, vfb = .7f;
#pragma unroll
            for (int it = 0; it < 3; it++) {
                float r = vfr * cov[s][0] + vfg * cov[s][1] + vfb * cov[s][2];
                float g = vfr * cov[s][1] + vfg * cov[s][3] + vfb * cov[s][4];
                float b = vfr * cov[s][2] + vfg * cov[s][4] + vfb * cov[s][5];
                float m = fmaxf(fmaxf(fabsf(r), fabsf(g)), fabsf(b));
                if (m > 1e-10f) {
                    m = 1.0f / m;
                    r *= m; g *= m; b *= m;
                }
                vfr = r; vfg = g; vfb = b;
            }
            float len = vfr * vfr + vfg * vfg + vfb * vfb;
            if (len >= 1e-10f) {
                len = 1.0f / sqrtf(len);
                axis[s][0] = vfr * len; axis[s][1] = vfg * len; axis[s][2] = vfb * len;
            }
        }
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
        if (axis[s][0] * axis[s][0] + axis[s][1] * axis[s][1] + axis[s][2] * axis[s][2] + axis[s][3] * axis[s][3] < .5f) {
            // perceptual fallback axis, normalised (bc7enc.cpp:1224-1231)
            axis[s][0] = .213f; axis[s][1] = .715f; axis[s][2] = .072f; axis[s][3] = ALPHA ? .715f : 0.0f;
            float len = axis[s][0] * axis[s][0] + axis[s][1] * axis[s][1] + axis[s][2] * axis[s][2] + axis[s][3] * axis[s][3];
            len = 1.0f / sqrtf(len);
#pragma unroll
            for (int c = 0; c < 4; c++) axis[s][c] *= len;
        }
    }
    // extent of the subset along the axis
    float l[NS], h[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) { l[s] = 1e+9f; h[s] = -1e+9f; }
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const int sb = NS == 2 ? (int)((mask >> i) & 1) : 0;
#pragma unroll
        for (int s = 0; s < NS; s++) {
            if (sb != s) continue;
            const float d = ((float)(v & 255) - ms[s][0]) * axis[s][0] + ((float)((v >> 8) & 255) - ms[s][1]) * axis[s][1] +
                            ((float)((v >> 16) & 255) - ms[s][2]) * axis[s][2] + ((float)(v >> 24) - ms[s][3]) * axis[s][3];
            l[s] = fminf(l[s], d);
            h[s] = fmaxf(h[s], d);
        }
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const float ll = l[s] * (1.0f / 255.0f), hh = h[s] * (1.0f / 255.0f);
        float sl = 0.0f, sh = 0.0f;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            xl[s][c] = sat(mean[s][c] + axis[s][c] * ll);
            xh[s][c] = sat(mean[s][c] + axis[s][c] * hh);
        }
        sl = xl[s][0] * 1.0f + xl[s][1] * 1.0f + xl[s][2] * 1.0f + xl[s][3] * 1.0f;
        sh = xh[s][0] * 1.0f + xh[s][1] * 1.0f + xh[s][2] * 1.0f + xh[s][3] * 1.0f;
        if (sl > sh) {
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const float tmp = xl[s][c];
                xl[s][c] = xh[s][c];
                xh[s][c] = tmp;
            }
        }
    }
}

// Selector the trial t feeds to the least-squares refit for a texel whose best selector is `cur` and whose
// snapshot selector (taken at t = 2) is `snap` (bc7enc.cpp:1281-1410).
// level: the uber level (bc7enc.cpp:1384): Q = level - 2 at level 4, else 1; the sweep runs ly in -Q..1, hy in N-2..N-1+Q.
template <int N> __device__ __forceinline__ int trial_selector(int t, int cur, int snap, int min_sel, int max_sel, int level)
{
    if (t == 1) return cur;
    if (t < 5) {
        const bool up = (t != 3) && snap == min_sel && snap < N - 1;
        const bool dn = (t != 2) && snap == max_sel && snap > 0;
        return up ? snap + 1 : (dn ? snap - 1 : snap);
    }
    int ly, hy;
    if (level >= 4) {
        const int q = (t - 5 < 9) ? t - 5 : t - 4; // skips (ly, hy) = (0, N - 1)
        ly = -2 + (q >> 2);
        hy = N - 2 + (q & 3);
    } else {
        const int q = (t - 5 < 4) ? t - 5 : t - 4; // 3 x 3 grid minus the identity
        ly = -1 + q / 3;
        hy = N - 2 + q % 3;
    }
    float f = floorf((float)(N - 1) * ((float)snap - (float)ly) / ((float)hy - (float)ly) + .5f);
    f = f < 0.0f ? 0.0f : (f > (float)(N - 1) ? (float)(N - 1) : f);
    return (int)f;
}

// compute_least_squares_endpoints_rgb / _rgba (bc7enc.cpp:287-407) for every subset, on the selectors of trial t.
// Output already multiplied by 1/255 as every call site does.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void least_squares(const Block& blk, uint32_t mask, const Tables& T, int t, uint64_t sels, uint64_t st, const int min_sel[],
                                              const int max_sel[], const int n[], const int sum[][4], float xl[][4], float xh[][4], int level)
{
    using C = Cfg<MODE>;
    constexpr int NS = C::NS, N = C::N, NC = ALPHA ? 4 : 3;
    float z00[NS], z10[NS], z11[NS], q00[NS][4];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        z00[s] = z10[s] = z11[s] = 0.0f;
        q00[s][0] = q00[s][1] = q00[s][2] = q00[s][3] = 0.0f;
    }
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const int sb = NS == 2 ? (int)((mask >> i) & 1) : 0;
        const int mn = (NS == 2 && sb) ? min_sel[NS - 1] : min_sel[0], mx = (NS == 2 && sb) ? max_sel[NS - 1] : max_sel[0];
        const int sel = trial_selector<N>(t, (int)(sels >> (4 * i)) & 15, (int)(st >> (4 * i)) & 15, mn, mx, level);
        const float4 wr = *(const float4*)wx_row<N>(T, sel);
        const float c[4] = {(float)(v & 255), (float)((v >> 8) & 255), (float)((v >> 16) & 255), (float)(v >> 24)};
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const bool in = sb == s;
            z00[s] += in ? wr.x : 0.0f;
            z10[s] += in ? wr.y : 0.0f;
            z11[s] += in ? wr.z : 0.0f;
#pragma unroll
            for (int k = 0; k < NC; k++) q00[s][k] += in ? wr.w * c[k] : 0.0f;
        }
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const float z01 = z10[s];
        float det = z00[s] * z11[s] - z01 * z10[s];
        if (det != 0.0f) det = 1.0f / det;
        const float iz00 = z11[s] * det, iz01 = -z01 * det, iz10 = -z10[s] * det, iz11 = z00[s] * det;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < NC) {
                const float q10 = (float)sum[s][k] - q00[s][k];
                float l = iz00 * q00[s][k] + iz01 * q10;
                float h = iz10 * q00[s][k] + iz11 * q10;
                if (l < 0.0f || h > 255.0f) {
                    int lo_v, hi_v;
                    subset_minmax(blk, mask, NS, s, k, lo_v, hi_v);
                    if (lo_v == hi_v) {
                        l = (float)lo_v;
                        h = (float)hi_v;
                    }
                }
                xl[s][k] = l * (1.0f / 255.0f);
                xh[s][k] = h * (1.0f / 255.0f);
            } else {
                xl[s][k] = 255.0f * (1.0f / 255.0f);
                xh[s][k] = 255.0f * (1.0f / 255.0f);
            }
        }
    }
}

// evaluate_solution (bc7enc.cpp:645-834) for the subsets flagged in `need`: palette from (lo, hi, pbits), per-texel
// exhaustive selector search (first minimum, perceptual branch :778-815), keep a subset's trial if strictly better.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void evaluate_solution(const Block& blk, uint32_t mask, const bool need[], const uint32_t lo[], const uint32_t hi[],
                                                  const uint32_t pb[], Sub res[], uint64_t& sels)
{
    using C = Cfg<MODE>;
    constexpr int NS = C::NS, N = C::N;
    int pl[NS][N], pcr[NS][N], pcb[NS][N], pa[NS][ALPHA ? N : 1];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        int lo8[4], hi8[4];
        expand_endpoints<MODE>(lo[s], hi[s], pb[s], lo8, hi8);
#pragma unroll
        for (int j = 0; j < N; j++) {
            constexpr int dummy = 0;
            (void)dummy;
            const int w = weight_of<N>(j);
            int c[4];
#pragma unroll
            for (int k = 0; k < (ALPHA ? 4 : 3); k++) c[k] = (j == 0) ? lo8[k] : (j == N - 1) ? hi8[k] : ((lo8[k] * (64 - w) + hi8[k] * w + 32) >> 6);
            pl[s][j] = c[0] * 109 + c[1] * 366 + c[2] * 37;
            pcr[s][j] = (c[0] << 9) - pl[s][j];
            pcb[s][j] = (c[2] << 9) - pl[s][j];
            if (ALPHA) pa[s][j] = c[3];
        }
    }
    uint32_t tot[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) tot[s] = 0;
    uint64_t trial = 0;
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const bool sb = NS == 2 && ((mask >> i) & 1);
        const int r = v & 255, g = (v >> 8) & 255, b = (v >> 16) & 255, a = v >> 24;
        const int l2 = r * 109 + g * 366 + b * 37;
        const int cr2 = (r << 9) - l2, cb2 = (b << 9) - l2;
        // The first entry of least error, as `if (e < best)` over ascending j finds it, taken as the minimum of 16 e + j: the
        // error of a texel is below 2^28 (|dl| <= 510, |dcr| <= 803, |dcb| <= 947, |da| <= 255: at most 224 M), the entry index
        // rides in the low four bits through the same multiply-adds, and one unsigned min replaces a compare and two selects.
        uint32_t bestKey = 0xFFFFFFFFu;
#pragma unroll
        for (int j = 0; j < N; j++) {
            const int dl = ((sb ? pl[NS - 1][j] : pl[0][j]) - l2) >> 8;
            const int dcr = ((sb ? pcr[NS - 1][j] : pcr[0][j]) - cr2) >> 8;
            const int dcb = ((sb ? pcb[NS - 1][j] : pcb[0][j]) - cb2) >> 8;
            uint32_t key = (uint32_t)j + (uint32_t)(16 * WY * (dl * dl)) + (uint32_t)(16 * WCR * (dcr * dcr)) + (uint32_t)(16 * WCB * (dcb * dcb));
            if (ALPHA) {
                const int da = (sb ? pa[NS - 1][ALPHA ? j : 0] : pa[0][ALPHA ? j : 0]) - a;
                key += (uint32_t)(16 * WA * (da * da));
            }
            bestKey = min(bestKey, key);
        }
        const uint32_t best = bestKey >> 4;
        const int bs = (int)(bestKey & 15u);
        tot[0] += sb ? 0u : best;
        if (NS == 2) tot[NS - 1] += sb ? best : 0u;
        trial |= (uint64_t)bs << (4 * i);
    }
    // nibble mask of subset 1's texels
    uint64_t m1 = 0;
    if (NS == 2) {
#pragma unroll
        for (int i = 0; i < 16; i++) m1 |= (uint64_t)((mask >> i) & 1) << (4 * i);
        m1 *= 15;
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
        if (need[s] && tot[s] < res[s].err) {
            res[s].err = tot[s];
            res[s].lo = lo[s];
            res[s].hi = hi[s];
            res[s].pb = pb[s];
            const uint64_t m = NS == 1 ? ~0ull : (s ? m1 : ~m1);
            sels = (sels & ~m) | (trial & m);
        }
    }
}

// color_cell_compression (bc7enc.cpp:1101-1441) at uber level `level` (4 = the Compress* preset), for all NS subsets of the mode
// at once.  The level only shortens the trial schedule: level 0 stops after t = 1 (:1300), level 1 after t = 4 (:1382), levels
// 2 and 3 sweep a 3 x 3 grid (t = 5..12), level 4 the 4 x 4 one (t = 5..19) (:1384).
// mask: bit i = subset of texel i (0 for one-subset modes).  res[s]: best per subset; sels: 4 bits per texel.
//
// Trial schedule (one loop, so that the quantise + evaluate body exists once in the instruction stream):
//   t = 0      endpoints from the principal axis                                   (bc7enc.cpp:1143-1279)
//   t = 1      least-squares refit of the best selectors                           (:1281-1298)
//   t = 2..4   refits with the minimum selectors raised / maximum lowered / both   (:1300-1376)
//   t = 5..19  refits of the t = 2 snapshot rescaled by (ly, hy), ly in -2..1, hy in N-2..N+1 minus the identity,
//              if the error after t = 4 exceeds (n * 56) >> 4                      (:1378-1410)
// then, modes 1 / 7, the single-colour encoding of the subset mean (:1413-1438).  A subset whose error reaches 0
// stops (the reference returns), as does one whose texels are all equal (:1108-1141).
template <int MODE, bool ALPHA>
__device__ __forceinline__ void color_cell_compression(const Block& blk, uint32_t mask, const Tables& T, Sub res[Cfg<MODE>::NS], uint64_t& sels,
                                                       const int level = 4)
{
    const int tmax = level >= 4 ? 20 : level >= 2 ? 13 : level == 1 ? 5 : 2;
    using C = Cfg<MODE>;
    constexpr int NS = C::NS, N = C::N;
    constexpr bool ONECOL = (MODE == 1 || MODE == 7);

    // ---- subset statistics
    int n[NS], sum[NS][4];
    uint32_t first[NS];
    bool same[NS], fin[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        n[s] = 0;
        sum[s][0] = sum[s][1] = sum[s][2] = sum[s][3] = 0;
        first[s] = 0;
        same[s] = true;
        fin[s] = false;
        res[s].err = 0xFFFFFFFFu;
        res[s].lo = res[s].hi = res[s].pb = 0;
    }
    sels = 0;
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const int sb = NS == 2 ? (int)((mask >> i) & 1) : 0;
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const bool in = sb == s;
            if (in && n[s] == 0) first[s] = v;
            if (in && ((v ^ first[s]) & (MODE == 1 ? 0x00FFFFFFu : 0xFFFFFFFFu))) same[s] = false;
            n[s] += in ? 1 : 0;
            sum[s][0] += in ? (int)(v & 255) : 0;
            sum[s][1] += in ? (int)((v >> 8) & 255) : 0;
            sum[s][2] += in ? (int)((v >> 16) & 255) : 0;
            sum[s][3] += in ? (int)(v >> 24) : 0;
        }
    }
    if constexpr (ONECOL) {
        // a subset whose texels are all equal is packed through the single-colour tables and is done (bc7enc.cpp:1108-1141)
        if (same[0] || same[1]) {
            const bool want[2] = {same[0], same[1]};
            int col[2][4];
            Sub pk[2];
            unpack4(first[0], col[0]);
            unpack4(first[1], col[1]);
            pack_one_color<MODE>(blk, mask, T, want, col, pk);
#pragma unroll
            for (int s = 0; s < 2; s++)
                if (same[s]) {
                    res[s] = pk[s];
                    fin[s] = true;
                }
#pragma unroll 1
            for (int i = 0; i < 16; i++)
                if (((mask >> i) & 1) ? same[1] : same[0]) sels |= (uint64_t)(MODE == 1 ? 2 : 1) << (4 * i);
        }
    }

    float ms[NS][4], mean[NS][4];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const float inv_n = 1.0f / (float)n[s];
        const float inv_n255 = 1.0f / ((float)n[s] * 255.0f);
#pragma unroll
        for (int c = 0; c < 4; c++) {
            ms[s][c] = (float)sum[s][c] * inv_n;
            mean[s][c] = sat((float)sum[s][c] * inv_n255);
        }
    }

    uint64_t st = 0;
    int min_sel[NS], max_sel[NS];
    bool sweep[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        min_sel[s] = max_sel[s] = 0;
        sweep[s] = true;
    }

#pragma unroll 1
    for (int t = 0; t < tmax; t++) {
        bool live = false;
#pragma unroll
        for (int s = 0; s < NS; s++) live |= !fin[s];
        if (!live) break;
        float xl[NS][4], xh[NS][4];
        if (t == 0) {
            pca_endpoints<MODE, ALPHA>(blk, mask, ms, mean, xl, xh);
        } else {
            if (t == 2) {
                st = sels;
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    min_sel[s] = 16;
                    max_sel[s] = 0;
                }
#pragma unroll 1
                for (int i = 0; i < 16; i++) {
                    const int sb = NS == 2 ? (int)((mask >> i) & 1) : 0;
                    const int sl = (int)(st >> (4 * i)) & 15;
#pragma unroll
                    for (int s = 0; s < NS; s++)
                        if (sb == s) {
                            min_sel[s] = min(min_sel[s], sl);
                            max_sel[s] = max(max_sel[s], sl);
                        }
                }
            }
            if (t == 5) {
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    sweep[s] = res[s].err > (uint32_t)((n[s] * 56) >> 4);
                    if (!sweep[s]) fin[s] = true; // nothing further is tried for this subset
                }
                bool left = false;
#pragma unroll
                for (int s = 0; s < NS; s++) left |= !fin[s];
                if (!left) break;
            }
            least_squares<MODE, ALPHA>(blk, mask, T, t, sels, st, min_sel, max_sel, n, sum, xl, xh, level);
        }
        // find_optimal_solution: quantise, then evaluate unless these are the current best endpoints (bc7enc.cpp:1083-1084, 1094-1095)
        uint32_t lo[NS], hi[NS], pb[NS];
        bool need[NS], any = false;
#pragma unroll
        for (int s = 0; s < NS; s++) {
            quantize_endpoints<MODE, ALPHA>(xl[s], xh[s], T, lo[s], hi[s], pb[s]);
            need[s] = !fin[s] && (res[s].err == 0xFFFFFFFFu || lo[s] != res[s].lo || hi[s] != res[s].hi || (C::PB && pb[s] != res[s].pb));
            any |= need[s];
        }
        if (any) evaluate_solution<MODE, ALPHA>(blk, mask, need, lo, hi, pb, res, sels);
#pragma unroll
        for (int s = 0; s < NS; s++)
            if (res[s].err == 0) fin[s] = true; // `if (!find_optimal_solution(...)) return 0`
    }

    if constexpr (ONECOL) {
        // single colour at the subset mean (bc7enc.cpp:1413-1438); all-equal subsets and subsets at zero error returned earlier
        const bool want[2] = {!same[0] && res[0].err != 0, !same[1] && res[1].err != 0};
        if (want[0] || want[1]) {
            int col[2][4];
            Sub pk[2];
#pragma unroll
            for (int s = 0; s < 2; s++)
#pragma unroll
                for (int c = 0; c < 4; c++) col[s][c] = (int)(.5f + mean[s][c] * 255.0f);
            pack_one_color<MODE>(blk, mask, T, want, col, pk);
            bool take[2];
#pragma unroll
            for (int s = 0; s < 2; s++) {
                take[s] = want[s] && pk[s].err < res[s].err;
                if (take[s]) res[s] = pk[s];
            }
            if (take[0] || take[1]) {
#pragma unroll 1
                for (int i = 0; i < 16; i++)
                    if (((mask >> i) & 1) ? take[1] : take[0]) sels = (sels & ~(15ull << (4 * i))) | ((uint64_t)(MODE == 1 ? 2 : 1) << (4 * i));
            }
        }
    }
}

// color_cell_compression_est_mode1 / _mode7 (bc7enc.cpp:1443-1709) for both subsets of partition `mask`:
// endpoints at each subset's bounding-box corners, selector by thresholds along the diagonal; returns the
// summed error of the block.  (The reference's early-outs against best_err only save time: a partition that
// trips them is already not better.)
template <bool M7> __device__ __forceinline__ uint32_t estimate_partition_error(const Block& blk, uint32_t mask)
{
    constexpr int N = M7 ? 4 : 8;
    constexpr int NC = M7 ? 4 : 3;
    int lo[2][4], hi[2][4];
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
            lo[s][c] = 255;
            hi[s][c] = 0;
        }
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const bool sb = (mask >> i) & 1;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const int x = (v >> (8 * c)) & 255;
            lo[0][c] = min(lo[0][c], sb ? 255 : x);
            hi[0][c] = max(hi[0][c], sb ? 0 : x);
            lo[1][c] = min(lo[1][c], sb ? x : 255);
            hi[1][c] = max(hi[1][c], sb ? x : 0);
        }
    }
    int ax[2][4], thr[2][N - 1];
#pragma unroll
    for (int s = 0; s < 2; s++) {
#pragma unroll
        for (int c = 0; c < NC; c++) ax[s][c] = hi[s][c] - lo[s][c];
        int prev = 0;
#pragma unroll
        for (int j = 0; j < N; j++) {
            const int w = weight_of<N>(j);
            int dot = 0;
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const int e = (j == 0) ? lo[s][c] : (j == N - 1) ? hi[s][c] : ((lo[s][c] * (64 - w) + hi[s][c] * w + 32) >> 6);
                dot += e * ax[s][c];
            }
            if (j > 0) thr[s][j - 1] = (prev + dot + 1) >> 1;
            prev = dot;
        }
    }
    uint32_t total = 0;
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const uint32_t v = blk[i];
        const bool sb = (mask >> i) & 1;
        int x[4], l[4], h[4];
        unpack4(v, x);
        int d = 0;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            l[c] = sb ? lo[1][c] : lo[0][c];
            h[c] = sb ? hi[1][c] : hi[0][c];
            d += (h[c] - l[c]) * x[c];
        }
        // thresholds are non-decreasing, so the reference's if-chain (bc7enc.cpp:1511-1524) counts those <= d
        int s = 0;
#pragma unroll
        for (int j = 0; j < N - 1; j++) s += (d >= (sb ? thr[1][j] : thr[0][j])) ? 1 : 0;
        const int w = weight_of<N>(s);
        int e[4];
#pragma unroll
        for (int c = 0; c < NC; c++) e[c] = (s == 0) ? l[c] : (s == N - 1) ? h[c] : ((l[c] * (64 - w) + h[c] * w + 32) >> 6);
        const int l2 = x[0] * 109 + x[1] * 366 + x[2] * 37;
        total += color_err<M7>(e[0], e[1], e[2], M7 ? e[3] : 255, l2, (x[0] << 9) - l2, (x[2] << 9) - l2, x[3]);
    }
    return total;
}

// estimate_partition (bc7enc.cpp:1754-1835): scan in the usage-sorted order, keep the first minimum; slots 14..34
// are evaluated only when the best of the first 14 predicts them (the mode-1/7 filterbank), and the scan stops
// after slot 34 unless the checkerboard (partition 34) ranks best.
template <bool M7> __device__ __forceinline__ uint32_t estimate_partition(const Block& blk)
{
    uint32_t best_err = 0xFFFFFFFFu, best_part = 0;
    int key = 0;
#pragma unroll 1
    for (int it = 0; it < 64 && best_err > 0; it++) {
        const uint32_t part = kPartOrder[it];
        if (it >= 14 && it <= 34) {
            if ((kPartPredictors[part] & (1u << (key + 1))) == 0) {
                if (it == 34) break;
                continue;
            }
        }
        const uint32_t e = estimate_partition_error<M7>(blk, kShape2[part]);
        if (e < best_err) {
            best_err = e;
            best_part = part;
        }
        if (part == 34 && best_part != 34) break;
        if (it == 13) key = (int)best_part;
    }
    return best_part;
}

// ---- bit packing (encode_bc7_block, bc7enc.cpp:1867-2037) ---------------------------------------
struct Bits {
    uint64_t lo = 0, hi = 0;
    int      pos = 0;
    __device__ __forceinline__ void put(uint64_t v, int n)
    {
        if (pos < 64) {
            lo |= v << pos;
            if (pos + n > 64) hi |= v >> (64 - pos);
        } else {
            hi |= v << (pos - 64);
        }
        pos += n;
    }
    __device__ __forceinline__ uint4 get() const { return make_uint4((uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32)); }
};

// 4-bit-per-texel selectors -> BITS per texel, complementing the texels of subsets whose anchor has its top bit set
// (inv bit s = subset s is flipped); the anchors' top bits are then dropped by the caller
template <int BITS> __device__ __forceinline__ uint64_t repack_selectors(uint64_t sels, uint32_t mask, uint32_t inv)
{
    uint64_t out = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t s = (uint32_t)(sels >> (4 * i)) & 15u;
        if ((inv >> ((mask >> i) & 1)) & 1) s = ((1u << BITS) - 1) - s;
        out |= (uint64_t)s << (BITS * i);
    }
    return out;
}
__device__ __forceinline__ uint64_t remove_bit(uint64_t x, int pos) { return (x & ((1ull << pos) - 1)) | ((x >> (pos + 1)) << pos); }

// mode 6: 7-bit RGBA endpoints + 2 p-bits, 4-bit indices
__device__ __forceinline__ uint4 pack_mode6(const Sub& r, uint64_t sels)
{
    const bool inv = (sels & 8) != 0;
    const uint64_t idx = repack_selectors<4>(sels, 0, inv ? 1 : 0);
    int l[4], h[4];
    unpack4(inv ? r.hi : r.lo, l);
    unpack4(inv ? r.lo : r.hi, h);
    const uint32_t p0 = inv ? (r.pb >> 1) & 1 : r.pb & 1, p1 = inv ? r.pb & 1 : (r.pb >> 1) & 1;
    Bits B;
    B.put(1u << 6, 7);
#pragma unroll
    for (int c = 0; c < 4; c++) {
        B.put(l[c], 7);
        B.put(h[c], 7);
    }
    B.put(p0, 1);
    B.put(p1, 1);
    B.put((idx & 7) | ((idx >> 4) << 3), 63);
    return B.get();
}

// modes 1 / 7: two subsets
template <int MODE> __device__ __forceinline__ uint4 pack_mode17(uint32_t part, const Sub r[2], uint64_t sels)
{
    constexpr int IB = (MODE == 1) ? 3 : 2; // index bits
    constexpr int EB = (MODE == 1) ? 6 : 5; // endpoint bits
    constexpr int NC = (MODE == 1) ? 3 : 4;
    const uint32_t mask = kShape2[part];
    const int a1 = kAnchor2[part];
    const uint32_t inv = ((uint32_t)(sels >> (IB - 1)) & 1u) | (((uint32_t)(sels >> (4 * a1 + IB - 1)) & 1u) << 1);
    uint64_t idx = repack_selectors<IB>(sels, mask, inv);
    Bits B;
    B.put(1u << MODE, MODE + 1);
    B.put(part, 6);
    int l[2][4], h[2][4];
    uint32_t pb[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const bool iv = (inv >> k) & 1;
        unpack4(iv ? r[k].hi : r[k].lo, l[k]);
        unpack4(iv ? r[k].lo : r[k].hi, h[k]);
        pb[k] = (MODE == 7 && iv) ? ((r[k].pb >> 1) | ((r[k].pb & 1) << 1)) : r[k].pb;
    }
#pragma unroll
    for (int c = 0; c < NC; c++)
#pragma unroll
        for (int k = 0; k < 2; k++) {
            B.put(l[k][c], EB);
            B.put(h[k][c], EB);
        }
#pragma unroll
    for (int k = 0; k < 2; k++) {
        B.put(pb[k] & 1, 1);
        if (MODE == 7) B.put((pb[k] >> 1) & 1, 1);
    }
    idx = remove_bit(idx, IB * a1 + IB - 1);
    idx = remove_bit(idx, IB - 1);
    B.put(idx, 16 * IB - 2);
    return B.get();
}

// mode 5: 7-bit RGB + 8-bit alpha endpoints, separate 2-bit colour and alpha indices, rotation 0
__device__ __forceinline__ uint4 pack_mode5(const Sub& r, uint64_t csels, int alo, int ahi, uint64_t asels)
{
    const bool cinv = (csels & 2) != 0, ainv = (asels & 2) != 0;
    const uint64_t cidx = repack_selectors<2>(csels, 0, cinv ? 1 : 0);
    const uint64_t aidx = repack_selectors<2>(asels, 0, ainv ? 1 : 0);
    int l[4], h[4];
    unpack4(cinv ? r.hi : r.lo, l);
    unpack4(cinv ? r.lo : r.hi, h);
    Bits B;
    B.put(1u << 5, 6);
    B.put(0, 2);
#pragma unroll
    for (int c = 0; c < 3; c++) {
        B.put(l[c], 7);
        B.put(h[c], 7);
    }
    B.put(ainv ? ahi : alo, 8);
    B.put(ainv ? alo : ahi, 8);
    B.put((cidx & 1) | ((cidx >> 2) << 1), 31);
    B.put((aidx & 1) | ((aidx >> 2) << 1), 31);
    return B.get();
}

// error totals pass through float when modes are compared: (uint64_t)(err * m_modeN_error_weight + .5f) with
// weight 1.0f (bc7enc.cpp:2168, 2184, 2237, 2330, 2381)
__device__ __forceinline__ uint64_t weighted(uint64_t err) { return (uint64_t)(__ull2float_rn(err) * 1.0f + .5f); }

// handle_opaque_block (bc7enc.cpp:2293-2400): mode 6, then mode 1 on the estimated partition.
__device__ __forceinline__ uint4 encode_opaque(const Block& blk, const Tables& T, const int level = 4)
{
    Sub r6[1];
    uint64_t s6;
    color_cell_compression<6, false>(blk, 0, T, r6, s6, level);
    uint4 out = pack_mode6(r6[0], s6);
    const uint64_t best = weighted(r6[0].err);
    if (best > 0) {
        const uint32_t part = estimate_partition<false>(blk);
        Sub r1[2];
        uint64_t s1;
        color_cell_compression<1, false>(blk, kShape2[part], T, r1, s1, level);
        // a first subset already above the mode-6 error ends the trial (bc7enc.cpp:2380-2382); the sum can only be larger
        if (weighted((uint64_t)r1[0].err + r1[1].err) < best) out = pack_mode17<1>(part, r1, s1);
    }
    return out;
}

// Alpha half of mode 5 (handle_alpha_block_mode5, bc7enc.cpp:2078-2136): up to 3 passes of nearest-of-4
// selection and least-squares endpoint refit on the alpha channel alone.
__device__ __forceinline__ uint32_t mode5_alpha(const Block& blk, const Tables& T, int& out_lo, int& out_hi, uint64_t& out_sels, const int level = 4)
{
    const int total_passes = level >= 1 ? 3 : 2; // bc7enc.cpp:2078
    int lo_a = 255, hi_a = 0, ta = 0;
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const int a = blk[i] >> 24;
        lo_a = min(lo_a, a);
        hi_a = max(hi_a, a);
        ta += a;
    }
    out_lo = lo_a;
    out_hi = hi_a;
    out_sels = 0;
    if (lo_a == hi_a) return 0;
    uint32_t best = 0xFFFFFFFFu;
#pragma unroll 1
    for (int pass = 0; pass < total_passes; pass++) {
        int vals[4];
        vals[0] = lo_a;
        vals[3] = hi_a;
        vals[1] = (vals[0] * (64 - 21) + vals[3] * 21 + 32) >> 6;
        vals[2] = (vals[0] * (64 - 43) + vals[3] * 43 + 32) >> 6;
        uint64_t tsel = 0;
        uint32_t terr = 0;
        float z00 = 0.0f, z10 = 0.0f, z11 = 0.0f, q00 = 0.0f;
#pragma unroll 1
        for (int i = 0; i < 16; i++) {
            const int a = blk[i] >> 24;
            int s = 0, be = abs(a - vals[0]);
#pragma unroll
            for (int j = 1; j < 4; j++) {
                const int e = abs(a - vals[j]);
                if (e < be) { be = e; s = j; }
            }
            tsel |= (uint64_t)s << (4 * i);
            terr += (uint32_t)(be * be) * (uint32_t)WA;
            const float4 wr = *(const float4*)T.wx2[s];
            z00 += wr.x;
            z10 += wr.y;
            z11 += wr.z;
            q00 += wr.w * (float)a;
        }
        if (terr < best) {
            best = terr;
            out_lo = lo_a;
            out_hi = hi_a;
            out_sels = tsel;
        }
        if (pass != total_passes - 1) {
            // compute_least_squares_endpoints_a (bc7enc.cpp:410-459); its min == max fallback cannot trigger here (lo_a != hi_a at entry)
            const float q10 = (float)ta - q00;
            const float z01 = z10;
            float det = z00 * z11 - z01 * z10;
            if (det != 0.0f) det = 1.0f / det;
            const float iz00 = z11 * det, iz01 = -z01 * det, iz10 = -z10 * det, iz11 = z00 * det;
            const float xl = iz00 * q00 + iz01 * q10, xh = iz10 * q00 + iz11 * q10;
            int nlo = clampi((int)floorf(xl + .5f), 0, 255), nhi = clampi((int)floorf(xh + .5f), 0, 255);
            if (nlo > nhi) { const int tmp = nlo; nlo = nhi; nhi = tmp; }
            if (nlo == lo_a && nhi == hi_a) break;
            lo_a = nlo;
            hi_a = nhi;
        }
    }
    return best;
}

// handle_alpha_block (bc7enc.cpp:2139-2291): mode 6 (RGBA), mode 5, mode 7 on the estimated partition.
__device__ __forceinline__ uint4 encode_alpha(const Block& blk, const Tables& T, const int level = 4)
{
    Sub r6[1];
    uint64_t s6;
    color_cell_compression<6, true>(blk, 0, T, r6, s6, level);
    uint4 out = pack_mode6(r6[0], s6);
    uint64_t best = weighted(r6[0].err);
    if (best > 0) {
        Sub r5[1];
        uint64_t s5, as5;
        color_cell_compression<5, false>(blk, 0, T, r5, s5, level);
        int alo, ahi;
        const uint32_t aerr = mode5_alpha(blk, T, alo, ahi, as5, level);
        const uint64_t e5 = weighted((uint64_t)r5[0].err + aerr);
        if (e5 < best) {
            best = e5;
            out = pack_mode5(r5[0], s5, alo, ahi, as5);
        }
    }
    if (best > 0) {
        const uint32_t part = estimate_partition<true>(blk);
        Sub r7[2];
        uint64_t s7;
        color_cell_compression<7, true>(blk, kShape2[part], T, r7, s7, level);
        if (weighted((uint64_t)r7[0].err + r7[1].err) < best) out = pack_mode17<7>(part, r7, s7);
    }
    return out;
}

// encode_opaque / encode_alpha with the CTA's warps re-aligned between the stages of the search.  The encoder is ~200 KB of
// SASS against a 32 KB L1.5 instruction cache, so warps at different stages each stream their own instructions from L2;
// warps inside the same stage share the fetches.  cls: 0 opaque, 1 alpha, 2 no block (takes part in the barriers only).
// Every thread of the CTA must call this the same number of times.
#ifdef __CUDACC__
__device__ __forceinline__ uint4 encode_phased(const Block& blk, const Tables& T, int cls, const int level = 4)
{
    uint4 out = make_uint4(0, 0, 0, 0);
    uint64_t best = 0;
    if (cls < 2) {
        Sub r6[1];
        uint64_t s6;
        if (cls == 0)
            color_cell_compression<6, false>(blk, 0, T, r6, s6, level);
        else
            color_cell_compression<6, true>(blk, 0, T, r6, s6, level);
        out = pack_mode6(r6[0], s6);
        best = weighted(r6[0].err);
    }
    __syncthreads();
    uint32_t part = 0;
    if (best > 0) {
        if (cls == 1) {
            Sub r5[1];
            uint64_t s5, as5;
            color_cell_compression<5, false>(blk, 0, T, r5, s5, level);
            int alo, ahi;
            const uint32_t aerr = mode5_alpha(blk, T, alo, ahi, as5, level);
            const uint64_t e5 = weighted((uint64_t)r5[0].err + aerr);
            if (e5 < best) {
                best = e5;
                out = pack_mode5(r5[0], s5, alo, ahi, as5);
            }
            if (best > 0) part = estimate_partition<true>(blk);
        } else {
            part = estimate_partition<false>(blk);
        }
    }
    __syncthreads();
    if (best > 0) {
        Sub r[2];
        uint64_t s;
        if (cls == 1) {
            color_cell_compression<7, true>(blk, kShape2[part], T, r, s, level);
            if (weighted((uint64_t)r[0].err + r[1].err) < best) out = pack_mode17<7>(part, r, s);
        } else {
            color_cell_compression<1, false>(blk, kShape2[part], T, r, s, level);
            if (weighted((uint64_t)r[0].err + r[1].err) < best) out = pack_mode17<1>(part, r, s);
        }
    }
    return out;
}
#endif

} // namespace bc7
} // namespace hkbcn
