// bc7_block.cuh -- BC7 block encoder, device side (sm_100a).
//
// Replaces bc7enc_compress_block at uber level 4 with perceptual weights, the configuration
// Hk::TextureBlockCompression::CompressBC7 bakes in (Hork/Image/ImageEncoders.cpp:59-63, 207-212,
// 524-557 -> ThirdParty/bc7enc_rdo/bc7enc.cpp:2402-2438).  Same mode set (opaque: 6 then 1;
// any alpha < 255: 6, 5, 7), same candidate generators (PCA, least-squares refits, selector
// perturbation and rescale sweep, single-colour tables), same integer YCbCr error metric.
// Contract: PSNR >= reference - 0.05 dB and mean block error <= reference (not bit-exact).
//
// Mapping: 16 lanes per 4x4 block, lane = texel; a warp holds two blocks.  Everything that the
// reference loops over pixels for becomes a REDUX (redux.sync) over the lanes of a "group":
// the whole block for modes 5/6, one partition subset for modes 1/7 -- both subsets of a
// two-subset mode run at the same time, each lane carrying the state of its own subset.
// All floating-point statistics are formed from exact integer sums, so a group's lanes agree
// bit-for-bit on every derived endpoint.
#pragma once
#include "hkbcn_common.cuh"
#include "bcn_tables.cuh"

namespace hkbcn {
namespace bc7 {

constexpr uint32_t FULL = 0xFFFFFFFFu;

// error weights after the perceptual scaling of bc7enc.cpp:2412-2417 (checked on the host at init)
constexpr int WY = 512, WCR = 103, WCB = 18, WA = 128;

// lookup tables built on the host by the loops of bc7enc_compress_block_init (bc7enc.cpp:124-285)
struct Tables {
    uint32_t opt1[256][2];    // mode 1 single-colour: [c][pbit]      = err | lo << 16 | hi << 24
    uint32_t opt7[256][2][2]; // mode 7 single-colour: [c][hp][lp]
    float    mid1[64][2];     // endpoint quantisation midpoints, mode 1 [i][p]
    float    mid7[32][2];     // mode 7
    float    mid5[128];       // mode 5
};

// Partition scan order of estimate_partition (bc7enc.cpp:1767-1777).
__device__ const uint8_t kPartOrder[64] = {
    0, 13, 1, 2, 15, 14, 10, 16, 3, 23, 26, 6, 7, 21, 19, 29, 8, 4, 9, 20, 5, 31, 22, 17, 18, 11, 12, 30, 24, 25, 28, 27,
    32, 33, 34, 45, 46, 51, 49, 50, 48, 38, 39, 37, 53, 52, 54, 36, 57, 58, 55, 41, 40, 42, 43, 59, 44, 56, 47, 35, 60, 63, 62, 61};

struct Px {
    int r, g, b, a;   // texel
    int l, cr, cb;    // its integer YCbCr (compute_color_distance_rgb, bc7enc.cpp:505-529)
};

// Result of one colour-cell compression (color_cell_compressor_results, bc7enc.cpp:477-485);
// uniform over the lanes of a group except `sel`.
struct Res {
    uint32_t err;
    uint32_t lo, hi; // quantised endpoints, 4 x 8 bit (r | g << 8 | b << 16 | a << 24), without p-bit
    uint32_t pb;     // bit 0 = pbits[0] (low endpoint), bit 1 = pbits[1]
    int      sel;    // this lane's selector
};

__device__ __forceinline__ int gsum(uint32_t m, int v) { return __reduce_add_sync(m, v); }
__device__ __forceinline__ int gmin(uint32_t m, int v) { return __reduce_min_sync(m, v); }
__device__ __forceinline__ int gmax(uint32_t m, int v) { return __reduce_max_sync(m, v); }
__device__ __forceinline__ uint32_t gor(uint32_t m, uint32_t v) { return __reduce_or_sync(m, v); }

// order-preserving float <-> int map so that float min/max can ride on REDUX
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i ^ ((i >> 31) & 0x7FFFFFFF); }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7FFFFFFF)); }

__device__ __forceinline__ float sat(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// selector weights (bc7enc.cpp:48-50): {0,21,43,64}, {0,9,...,64}, {0,4,...,64}
template <int N> __device__ __forceinline__ int weight_of(int s)
{
    if (N == 4) return (s * 64 + 1) / 3;
    if (N == 8) return (s * 64 + 3) / 7;
    return (s * 64 + 7) / 15;
}

// per-mode constants (color_cell_compressor_params as the handlers fill it, bc7enc.cpp:2143-2153, 2315-2345, 2041-2049)
template <int MODE> struct Cfg;
template <> struct Cfg<6> { static constexpr int N = 16, CB = 7; static constexpr bool PB = true, SHARE = false; };
template <> struct Cfg<1> { static constexpr int N = 8, CB = 6; static constexpr bool PB = true, SHARE = true; };
template <> struct Cfg<5> { static constexpr int N = 4, CB = 7; static constexpr bool PB = false, SHARE = false; };
template <> struct Cfg<7> { static constexpr int N = 4, CB = 5; static constexpr bool PB = true, SHARE = false; };

// scale_color (bc7enc.cpp:487-503): n-bit component -> 8 bit
template <int NB> __device__ __forceinline__ int scale_comp(int v)
{
    v <<= (8 - NB);
    return v | (v >> NB);
}

// Per-pixel exhaustive selector search of evaluate_solution's perceptual branch (bc7enc.cpp:778-815):
// first minimum over the N interpolated colours.  lo8 / hi8: actual 8-bit endpoint components.
template <int N, bool ALPHA>
__device__ __forceinline__ uint32_t best_selector(const int lo8[4], const int hi8[4], const Px& p, int& best_sel)
{
    uint32_t best = 0xFFFFFFFFu;
    best_sel = 0;
#pragma unroll
    for (int j = 0; j < N; j++) {
        const int w = weight_of<N>(j);
        int c[4];
#pragma unroll
        for (int k = 0; k < (ALPHA ? 4 : 3); k++) {
            if (j == 0) c[k] = lo8[k];
            else if (j == N - 1) c[k] = hi8[k];
            else c[k] = (lo8[k] * (64 - w) + hi8[k] * w + 32) >> 6;
        }
        const int l1 = c[0] * 109 + c[1] * 366 + c[2] * 37;
        const int dl = (l1 - p.l) >> 8;
        const int dcr = ((c[0] << 9) - l1 - p.cr) >> 8;
        const int dcb = ((c[2] << 9) - l1 - p.cb) >> 8;
        uint32_t e = (uint32_t)(WY * (dl * dl)) + (uint32_t)(WCR * (dcr * dcr)) + (uint32_t)(WCB * (dcb * dcb));
        if (ALPHA) {
            const int da = c[3] - p.a;
            e += (uint32_t)(WA * (da * da));
        }
        if (e < best) {
            best = e;
            best_sel = j;
        }
    }
    return best;
}

// error of one fixed colour against this lane's texel
template <bool ALPHA> __device__ __forceinline__ uint32_t color_err(int r, int g, int b, int a, const Px& p)
{
    const int l1 = r * 109 + g * 366 + b * 37;
    const int dl = (l1 - p.l) >> 8;
    const int dcr = ((r << 9) - l1 - p.cr) >> 8;
    const int dcb = ((b << 9) - l1 - p.cb) >> 8;
    uint32_t e = (uint32_t)(WY * (dl * dl)) + (uint32_t)(WCR * (dcr * dcr)) + (uint32_t)(WCB * (dcb * dcb));
    if (ALPHA) {
        const int da = a - p.a;
        e += (uint32_t)(WA * (da * da));
    }
    return e;
}

__device__ __forceinline__ void unpack4(uint32_t v, int c[4])
{
    c[0] = v & 255;
    c[1] = (v >> 8) & 255;
    c[2] = (v >> 16) & 255;
    c[3] = v >> 24;
}
__device__ __forceinline__ uint32_t pack4(const int c[4]) { return (uint32_t)c[0] | ((uint32_t)c[1] << 8) | ((uint32_t)c[2] << 16) | ((uint32_t)c[3] << 24); }

// evaluate_solution (bc7enc.cpp:645-834): palette from (lo, hi, pbits), selector search, keep if better.
// `doit` is uniform over the group; groups that skip still take part in the warp-wide instruction stream.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void evaluate_solution(uint32_t gm, bool doit, uint32_t lo, uint32_t hi, uint32_t pb, const Px& p, Res& res)
{
    using C = Cfg<MODE>;
    int l[4], h[4], lo8[4], hi8[4];
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
    int sel;
    const uint32_t e = best_selector<C::N, ALPHA>(lo8, hi8, p, sel);
    const uint32_t total = (uint32_t)gsum(gm, (int)e);
    if (doit && total < res.err) {
        res.err = total;
        res.lo = lo;
        res.hi = hi;
        res.pb = pb;
        res.sel = sel;
    }
}

// fixDegenerateEndpoints (bc7enc.cpp:836-866), mode 1 only: iscale = 63
__device__ __forceinline__ void fix_degenerate_mode1(int mn[4], int mx[4], const float xl[4], const float xh[4])
{
#pragma unroll
    for (int i = 0; i < 3; i++) {
        if (mn[i] == mx[i] && fabsf(xl[i] - xh[i]) > 0.0f) {
            if (mn[i] > 31) {
                if (mn[i] > 0) mn[i]--;
                else if (mx[i] < 63) mx[i]++;
            } else {
                if (mx[i] < 63) mx[i]++;
                else if (mn[i] > 0) mn[i]--;
            }
        }
    }
}

// find_optimal_solution (bc7enc.cpp:868-1098): quantise float endpoints xl/xh (0..1) to the mode's
// grid, choosing p-bits by endpoint reconstruction error, then evaluate unless it is the current best.
// `active` (uniform over the group) gates the evaluation; the warp skips it when no group needs it.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void find_optimal_solution(uint32_t gm, bool active, float xl[4], float xh[4], const Px& p, Res& res,
                                                      const Tables& T)
{
    using C = Cfg<MODE>;
    constexpr int totalComps = ALPHA ? 4 : 3;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        xl[c] = sat(xl[c]);
        xh[c] = sat(xh[c]);
    }
    int bmin[4], bmax[4];
    uint32_t pb = 0;
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
                        int vl = __float2int_rz(xl[c] * 31.0f);
                        vl += (xl[c] > T.mid7[vl][pbit]);
                        qmin[c] = clampi(vl * 2 + pbit, pbit, 63 - 1 + pbit);
                        int vh = __float2int_rz(xh[c] * 31.0f);
                        vh += (xh[c] > T.mid7[vh][pbit]);
                        qmax[c] = clampi(vh * 2 + pbit, pbit, 63 - 1 + pbit);
                    } else {
                        qmin[c] = clampi(__float2int_rz((xl[c] * scalep - (float)pbit) / 2.0f + .5f) * 2 + pbit, pbit, iscalep - 1 + pbit);
                        qmax[c] = clampi(__float2int_rz((xh[c] * scalep - (float)pbit) / 2.0f + .5f) * 2 + pbit, pbit, iscalep - 1 + pbit);
                    }
                    if (c < totalComps) {
                        const float dl = (float)scale_comp<C::CB + 1>(qmin[c]) - xl[c] * 255.0f;
                        const float dh = (float)scale_comp<C::CB + 1>(qmax[c]) - xh[c] * 255.0f;
                        err0 += dl * dl;
                        err1 += dh * dh;
                    }
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
            // endpoints share the p-bit (mode 1)
            float best = 1e+9f;
#pragma unroll
            for (int pbit = 0; pbit < 2; pbit++) {
                int qmin[4], qmax[4];
                float err = 0.0f;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    int vl = __float2int_rz(xl[c] * 63.0f);
                    vl += (xl[c] > T.mid1[vl][pbit]);
                    qmin[c] = clampi(vl * 2 + pbit, pbit, 127 - 1 + pbit);
                    int vh = __float2int_rz(xh[c] * 63.0f);
                    vh += (xh[c] > T.mid1[vh][pbit]);
                    qmax[c] = clampi(vh * 2 + pbit, pbit, 127 - 1 + pbit);
                    if (c < totalComps) {
                        const float dl = (float)scale_comp<C::CB + 1>(qmin[c]) * (1.0f / 255.0f) - xl[c];
                        const float dh = (float)scale_comp<C::CB + 1>(qmax[c]) * (1.0f / 255.0f) - xh[c];
                        err += dl * dl + dh * dh;
                    }
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
            fix_degenerate_mode1(bmin, bmax, xl, xh);
        }
    } else {
        // mode 5 colour: 7 bits, no p-bits
#pragma unroll
        for (int c = 0; c < 4; c++) {
            int vl = __float2int_rz(xl[c] * 127.0f);
            vl += (xl[c] > T.mid5[vl]);
            bmin[c] = clampi(vl, 0, 127);
            int vh = __float2int_rz(xh[c] * 127.0f);
            vh += (xh[c] > T.mid5[vh]);
            bmax[c] = clampi(vh, 0, 127);
        }
    }
    const uint32_t lo = pack4(bmin), hi = pack4(bmax);
    const bool need = active && (res.err == 0xFFFFFFFFu || lo != res.lo || hi != res.hi || (C::PB && pb != res.pb));
    if (__any_sync(FULL, need)) evaluate_solution<MODE, ALPHA>(gm, need, lo, hi, pb, p, res);
}

// Group statistics that stay fixed over all trials of one colour-cell compression.
struct Stats {
    int   n;          // texels in the group
    int   tr, tg, tb, ta; // component sums
    uint32_t mn, mx;  // component-wise bounding box, packed 4 x 8
};

template <bool ALPHA> __device__ __forceinline__ void group_stats(uint32_t gm, const Px& p, Stats& S)
{
    S.n = __popc(gm);
    const uint32_t s0 = (uint32_t)gsum(gm, p.r | (p.g << 16));
    const uint32_t s1 = (uint32_t)gsum(gm, p.b | (p.a << 16));
    S.tr = s0 & 0xFFFF;
    S.tg = s0 >> 16;
    S.tb = s1 & 0xFFFF;
    S.ta = s1 >> 16;
    int mn[4], mx[4];
    mn[0] = gmin(gm, p.r); mx[0] = gmax(gm, p.r);
    mn[1] = gmin(gm, p.g); mx[1] = gmax(gm, p.g);
    mn[2] = gmin(gm, p.b); mx[2] = gmax(gm, p.b);
    if (ALPHA) { mn[3] = gmin(gm, p.a); mx[3] = gmax(gm, p.a); }
    else { mn[3] = 255; mx[3] = 255; }
    S.mn = pack4(mn);
    S.mx = pack4(mx);
}

// compute_least_squares_endpoints_rgb / _rgba (bc7enc.cpp:287-407) for the selectors `s`, from exact
// integer sums: with w = k/64, z00 = sum w^2, z10 = sum (1-w)w, z11 = sum (1-w)^2, q00_c = sum w c.
// Output already scaled by 1/255 as every call site does.
template <int N, bool ALPHA>
__device__ __forceinline__ void least_squares(uint32_t gm, int s, const Px& p, const Stats& S, float xl[4], float xh[4])
{
    const int k = weight_of<N>(s);
    const uint32_t ks = (uint32_t)gsum(gm, (k * k) | (k << 17));
    const int S2 = ks & 0x1FFFF, S1 = ks >> 17;
    int q[4];
    q[0] = gsum(gm, k * p.r);
    q[1] = gsum(gm, k * p.g);
    q[2] = gsum(gm, k * p.b);
    q[3] = ALPHA ? gsum(gm, k * p.a) : 0;
    const float z00 = (float)S2 * (1.0f / 4096.0f);
    const float z10 = (float)(64 * S1 - S2) * (1.0f / 4096.0f);
    const float z11 = (float)(4096 * S.n - 128 * S1 + S2) * (1.0f / 4096.0f);
    float det = z00 * z11 - z10 * z10;
    if (det != 0.0f) det = 1.0f / det;
    const float iz00 = z11 * det, iz01 = -z10 * det, iz11 = z00 * det;
    const int t[4] = {S.tr, S.tg, S.tb, S.ta};
    int mn[4], mx[4];
    unpack4(S.mn, mn);
    unpack4(S.mx, mx);
#pragma unroll
    for (int c = 0; c < 4; c++) {
        if (c == 3 && !ALPHA) {
            xl[c] = 255.0f * (1.0f / 255.0f);
            xh[c] = 255.0f * (1.0f / 255.0f);
        } else {
            const float q00 = (float)q[c] * (1.0f / 64.0f);
            const float q10 = (float)t[c] - q00;
            float l = iz00 * q00 + iz01 * q10;
            float h = iz01 * q00 + iz11 * q10;
            if ((l < 0.0f || h > 255.0f) && mn[c] == mx[c]) {
                l = (float)mn[c];
                h = (float)mx[c];
            }
            xl[c] = l * (1.0f / 255.0f);
            xh[c] = h * (1.0f / 255.0f);
        }
    }
}

// Initial endpoints from the group's principal axis (bc7enc.cpp:1143-1275).  `hm`/`base`: the lanes of
// the 4x4 block this lane belongs to (the RGBA incremental PCA walks the block's texels in order).
template <bool ALPHA>
__device__ __forceinline__ void pca_endpoints(uint32_t gm, int base, const Px& p, uint32_t packed, const Stats& S, float mean[4],
                                              float xl[4], float xh[4])
{
    const float inv_n = 1.0f / (float)S.n;
    const float ms[4] = {(float)S.tr * inv_n, (float)S.tg * inv_n, (float)S.tb * inv_n, (float)S.ta * inv_n};
    const float inv_n255 = 1.0f / ((float)S.n * 255.0f);
    mean[0] = sat((float)S.tr * inv_n255);
    mean[1] = sat((float)S.tg * inv_n255);
    mean[2] = sat((float)S.tb * inv_n255);
    mean[3] = sat((float)S.ta * inv_n255);
    float axis[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (ALPHA) {
        // incremental PCA over the group's texels in index order (bc7enc.cpp:1160-1178)
        bool first = true;
#pragma unroll 1
        for (int j = 0; j < 16; j++) {
            const uint32_t cj = __shfl_sync(FULL, packed, base + j);
            if ((gm >> (base + j)) & 1u) {
                float col[4];
                col[0] = (float)(cj & 255) - ms[0];
                col[1] = (float)((cj >> 8) & 255) - ms[1];
                col[2] = (float)((cj >> 16) & 255) - ms[2];
                col[3] = (float)(cj >> 24) - ms[3];
                float nv[4];
#pragma unroll
                for (int c = 0; c < 4; c++) nv[c] = first ? col[c] : axis[c];
                first = false;
                const float s = nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2] + nv[3] * nv[3];
                float d = col[0] * nv[0] + col[1] * nv[1] + col[2] * nv[2] + col[3] * nv[3];
                if (s != 0.0f) d *= rsqrtf(s);
#pragma unroll
                for (int c = 0; c < 4; c++) axis[c] += col[c] * d;
            }
        }
        const float s = axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2] + axis[3] * axis[3];
        if (s != 0.0f) {
            const float is = rsqrtf(s);
#pragma unroll
            for (int c = 0; c < 4; c++) axis[c] *= is;
        }
    } else {
        // covariance + 3 power iterations (bc7enc.cpp:1181-1219); cov = (n * sum xy - sum x * sum y) / n, exact numerators
        const int n = S.n;
        const float c0 = (float)(n * gsum(gm, p.r * p.r) - S.tr * S.tr) * inv_n;
        const float c1 = (float)(n * gsum(gm, p.r * p.g) - S.tr * S.tg) * inv_n;
        const float c2 = (float)(n * gsum(gm, p.r * p.b) - S.tr * S.tb) * inv_n;
        const float c3 = (float)(n * gsum(gm, p.g * p.g) - S.tg * S.tg) * inv_n;
        const float c4 = (float)(n * gsum(gm, p.g * p.b) - S.tg * S.tb) * inv_n;
        const float c5 = (float)(n * gsum(gm, p.b * p.b) - S.tb * S.tb) * inv_n;
        float vr = .9f, vg = 1.0f, vb = .7f;
#pragma unroll
        for (int it = 0; it < 3; it++) {
            float r = vr * c0 + vg * c1 + vb * c2;
            float g = vr * c1 + vg * c3 + vb * c4;
            float b = vr * c2 + vg * c4 + vb * c5;
            float m = fmaxf(fmaxf(fabsf(r), fabsf(g)), fabsf(b));
            if (m > 1e-10f) {
                m = 1.0f / m;
                r *= m; g *= m; b *= m;
            }
            vr = r; vg = g; vb = b;
        }
        const float len = vr * vr + vg * vg + vb * vb;
        if (len >= 1e-10f) {
            const float il = rsqrtf(len);
            axis[0] = vr * il; axis[1] = vg * il; axis[2] = vb * il;
        }
    }
    if (axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2] + axis[3] * axis[3] < .5f) {
        // perceptual fallback axis, normalised (bc7enc.cpp:1224-1231)
        axis[0] = .213f; axis[1] = .715f; axis[2] = .072f; axis[3] = ALPHA ? .715f : 0.0f;
        const float is = rsqrtf(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2] + axis[3] * axis[3]);
#pragma unroll
        for (int c = 0; c < 4; c++) axis[c] *= is;
    }
    const float d = ((float)p.r - ms[0]) * axis[0] + ((float)p.g - ms[1]) * axis[1] + ((float)p.b - ms[2]) * axis[2] +
                    ((float)p.a - ms[3]) * axis[3];
    const float l = ord2f(gmin(gm, f2ord(d))) * (1.0f / 255.0f);
    const float h = ord2f(gmax(gm, f2ord(d))) * (1.0f / 255.0f);
    float sl = 0.0f, sh = 0.0f;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        xl[c] = sat(mean[c] + axis[c] * l);
        xh[c] = sat(mean[c] + axis[c] * h);
        sl += xl[c];
        sh += xh[c];
    }
    if (sl > sh) {
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const float t = xl[c];
            xl[c] = xh[c];
            xh[c] = t;
        }
    }
}

// pack_mode1_to_one_color / pack_mode7_to_one_color (bc7enc.cpp:536-643): encode the group as the single
// colour (r,g,b[,a]) through the optimal-endpoint tables; every selector = 2 (mode 1) / 1 (mode 7).
template <int MODE>
__device__ __forceinline__ void pack_one_color(uint32_t gm, int r, int g, int b, int a, const Px& p, const Tables& T, Res& res)
{
    uint32_t lo, hi, pb;
    int col[4];
    if (MODE == 1) {
        uint32_t best_err = 0xFFFFFFFFu, best_p = 0;
#pragma unroll
        for (uint32_t pp = 0; pp < 2; pp++) {
            const uint32_t e = (T.opt1[r][pp] & 0xFFFF) + (T.opt1[g][pp] & 0xFFFF) + (T.opt1[b][pp] & 0xFFFF);
            if (e < best_err) { best_err = e; best_p = pp; }
        }
        const uint32_t er = T.opt1[r][best_p], eg = T.opt1[g][best_p], eb = T.opt1[b][best_p];
        lo = ((er >> 16) & 255) | (((eg >> 16) & 255) << 8) | (((eb >> 16) & 255) << 16);
        hi = (er >> 24) | ((eg >> 24) << 8) | ((eb >> 24) << 16);
        pb = best_p * 3u;
        int l[4], h[4];
        unpack4(lo, l);
        unpack4(hi, h);
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const int low = scale_comp<7>((l[i] << 1) | (int)best_p);
            const int high = scale_comp<7>((h[i] << 1) | (int)best_p);
            col[i] = (low * (64 - 18) + high * 18 + 32) >> 6; // g_bc7_weights3[2]
        }
        col[3] = 255;
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
        lo = ((er >> 16) & 255) | (((eg >> 16) & 255) << 8) | (((eb >> 16) & 255) << 16) | (((ea >> 16) & 255) << 24);
        hi = (er >> 24) | ((eg >> 24) << 8) | ((eb >> 24) << 16) | ((ea >> 24) << 24);
        pb = lp | (hp << 1);
        int l[4], h[4];
        unpack4(lo, l);
        unpack4(hi, h);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int low = scale_comp<6>((l[i] << 1) | (int)lp);
            const int high = scale_comp<6>((h[i] << 1) | (int)hp);
            col[i] = (low * (64 - 21) + high * 21 + 32) >> 6; // g_bc7_weights2[1]
        }
    }
    const uint32_t e = (MODE == 1) ? color_err<false>(col[0], col[1], col[2], 255, p) : color_err<true>(col[0], col[1], col[2], col[3], p);
    const uint32_t total = (uint32_t)gsum(gm, (int)e);
    if (total < res.err) {
        res.err = total;
        res.lo = lo;
        res.hi = hi;
        res.pb = pb;
        res.sel = (MODE == 1) ? 2 : 1;
    }
}

// color_cell_compression (bc7enc.cpp:1101-1441) at uber level 4 for the group `gm` of this lane.
// Trial schedule, one loop so the quantise+evaluate body exists once in the instruction stream:
//   t = 0      endpoints from the principal axis
//   t = 1      least-squares refit of the best selectors
//   t = 2..4   refits of the selectors with min raised / max lowered / both (snapshot taken at t = 2)
//   t = 5..19  refits of the snapshot rescaled by (ly, hy), ly in -2..1, hy in N-2..N+1 minus the identity,
//              only when the error after t = 4 exceeds (n * 56) >> 4
// then for modes 1 / 7 the single-colour encoding of the group mean.  The reference's early returns at
// zero error and its "same endpoints as the best" skips only save time (a later trial must be strictly
// better to replace the best), so running every trial gives the same result.
template <int MODE, bool ALPHA>
__device__ __forceinline__ void color_cell_compression(uint32_t gm, int base, const Px& p, uint32_t packed, const Tables& T, Res& res)
{
    using C = Cfg<MODE>;
    constexpr int N = C::N;
    Stats S;
    group_stats<ALPHA>(gm, p, S);
    res.err = 0xFFFFFFFFu;
    res.lo = res.hi = res.pb = 0;
    res.sel = 0;
    float mean[4];
    int st = 0, min_sel = 0, max_sel = 0;
    bool sweep = true;
#pragma unroll 1
    for (int t = 0; t < 20; t++) {
        float xl[4], xh[4];
        bool active = true;
        if (t == 0) {
            pca_endpoints<ALPHA>(gm, base, p, packed, S, mean, xl, xh);
        } else {
            int s;
            if (t == 1) {
                s = res.sel;
            } else {
                if (t == 2) {
                    st = res.sel;
                    min_sel = gmin(gm, st);
                    max_sel = gmax(gm, st);
                }
                if (t == 5) {
                    sweep = res.err > (uint32_t)((S.n * 56) >> 4);
                    if (!__any_sync(FULL, sweep)) break;
                }
                if (t < 5) {
                    const bool up = (t != 3) && st == min_sel && st < N - 1;
                    const bool dn = (t != 2) && st == max_sel && st > 0;
                    s = up ? st + 1 : (dn ? st - 1 : st);
                } else {
                    active = sweep;
                    const int q = (t - 5 < 9) ? t - 5 : t - 4;
                    const int ly = -2 + (q >> 2), hy = N - 2 + (q & 3);
                    const float f = floorf((float)(N - 1) * ((float)st - (float)ly) / ((float)hy - (float)ly) + .5f);
                    s = (int)fminf(fmaxf(f, 0.0f), (float)(N - 1));
                }
            }
            least_squares<N, ALPHA>(gm, s, p, S, xl, xh);
        }
        find_optimal_solution<MODE, ALPHA>(gm, active, xl, xh, p, res, T);
    }
    if (MODE == 1 || MODE == 7) {
        const int r = __float2int_rz(.5f + mean[0] * 255.0f), g = __float2int_rz(.5f + mean[1] * 255.0f);
        const int b = __float2int_rz(.5f + mean[2] * 255.0f), a = __float2int_rz(.5f + mean[3] * 255.0f);
        pack_one_color<MODE>(gm, r, g, b, a, p, T, res);
    }
}

// color_cell_compression_est_mode1 / _mode7 (bc7enc.cpp:1443-1709) for this lane's subset `gm`:
// endpoints at the subset's bounding-box corners, selector by thresholds along the diagonal,
// returns this texel's error (the caller sums over the block).
template <bool M7> __device__ __forceinline__ uint32_t estimate_texel_err(uint32_t gm, const Px& p)
{
    constexpr int N = M7 ? 4 : 8;
    constexpr int NC = M7 ? 4 : 3;
    const int v[4] = {p.r, p.g, p.b, p.a};
    int lo[4], hi[4], ax[4];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        lo[c] = gmin(gm, v[c]);
        hi[c] = gmax(gm, v[c]);
        ax[c] = hi[c] - lo[c];
    }
    int d = 0;
#pragma unroll
    for (int c = 0; c < NC; c++) d += ax[c] * v[c];
    // dots of the N interpolated colours along the diagonal, thresholds between neighbours
    int dots[N];
#pragma unroll
    for (int j = 0; j < N; j++) {
        const int w = weight_of<N>(j);
        int acc = 0;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const int e = (j == 0) ? lo[c] : (j == N - 1) ? hi[c] : ((lo[c] * (64 - w) + hi[c] * w + 32) >> 6);
            acc += e * ax[c];
        }
        dots[j] = acc;
    }
    int s = 0;
#pragma unroll
    for (int j = 0; j < N - 1; j++) s += (d >= ((dots[j] + dots[j + 1] + 1) >> 1)) ? 1 : 0; // thresholds are non-decreasing
    const int w = weight_of<N>(s);
    int e[4];
#pragma unroll
    for (int c = 0; c < NC; c++) e[c] = (s == 0) ? lo[c] : (s == N - 1) ? hi[c] : ((lo[c] * (64 - w) + hi[c] * w + 32) >> 6);
    return M7 ? color_err<true>(e[0], e[1], e[2], e[3], p) : color_err<false>(e[0], e[1], e[2], 255, p);
}

// estimate_partition (bc7enc.cpp:1754-1835).  Scans the partitions in the reference's usage-sorted
// order, keeps the first minimum of the estimated error.  Differences from the reference, both of
// which only widen the candidate set: the mode-1/7 predictor filterbank is not applied (the reference
// comments put its cost at ~0.05 dB), and the early-outs inside one estimate are dropped (they cannot
// change which partition wins).  The reference's rule "stop after slot 34 unless the checkerboard
// (partition 34) is the best so far" is kept.
template <bool M7> __device__ __forceinline__ uint32_t estimate_partition(uint32_t hm, int base, int lane16, const Px& p)
{
    uint32_t best_err = 0xFFFFFFFFu, best_part = 0;
    bool live = true;
#pragma unroll 1
    for (int it = 0; it < 64; it++) {
        if (it == 35) {
            live = best_part == 34;
            if (!__any_sync(FULL, live)) break;
        }
        const uint32_t part = kPartOrder[it];
        const uint32_t m16 = kShape2[part];
        const uint32_t gm = (((m16 >> lane16) & 1u) ? m16 : (~m16 & 0xFFFFu)) << base;
        const uint32_t e = estimate_texel_err<M7>(gm, p);
        const uint32_t tot = (uint32_t)gsum(hm, (int)e);
        if (live && tot < best_err) {
            best_err = tot;
            best_part = part;
        }
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
__device__ __forceinline__ uint64_t remove_bit(uint64_t x, int pos) { return (x & ((1ull << pos) - 1)) | ((x >> (pos + 1)) << pos); }

// all 16 selectors of the block as one word, BITS bits per texel in texel order
template <int BITS> __device__ __forceinline__ uint64_t gather_selectors(uint32_t hm, int lane16, int sel)
{
    constexpr int PER = 32 / BITS; // texels whose bits fit the low word (BITS = 2: 16, 3: 10, 4: 8)
    const uint32_t w0 = gor(hm, lane16 < PER ? (uint32_t)sel << (BITS * lane16) : 0u);
    if (BITS == 2) return w0;
    const uint32_t w1 = gor(hm, lane16 >= PER ? (uint32_t)sel << (BITS * (lane16 - PER)) : 0u);
    return (uint64_t)w0 | ((uint64_t)w1 << (BITS * PER));
}

// mode 6: 7-bit RGBA endpoints + 2 p-bits, 4-bit indices
__device__ __forceinline__ uint4 pack_mode6(uint32_t hm, int base, int lane16, const Res& r)
{
    const int s0 = __shfl_sync(FULL, r.sel, base);
    const bool inv = (s0 & 8) != 0;
    const uint64_t idx = gather_selectors<4>(hm, lane16, inv ? 15 - r.sel : r.sel);
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

// modes 1 / 7: two subsets; every lane carries its own subset's result in `r`
template <int MODE>
__device__ __forceinline__ uint4 pack_mode17(uint32_t hm, int base, int lane16, uint32_t part, const Res& r)
{
    constexpr int IB = (MODE == 1) ? 3 : 2;       // index bits
    constexpr int EB = (MODE == 1) ? 6 : 5;       // endpoint bits
    constexpr int NC = (MODE == 1) ? 3 : 4;
    constexpr int MSB = 1 << (IB - 1), MAXS = (1 << IB) - 1;
    const int a1 = kAnchor2[part];
    const int mysub = (kShape2[part] >> lane16) & 1;
    const int asel = __shfl_sync(FULL, r.sel, base + (mysub ? a1 : 0));
    const bool inv = (asel & MSB) != 0;
    uint64_t idx = gather_selectors<IB>(hm, lane16, inv ? MAXS - r.sel : r.sel);
    const uint32_t mylo = inv ? r.hi : r.lo, myhi = inv ? r.lo : r.hi;
    const uint32_t mypb = (MODE == 7 && inv) ? ((r.pb >> 1) | ((r.pb & 1) << 1)) : r.pb;
    int l[2][4], h[2][4];
    uint32_t pb[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int src = base + (k ? a1 : 0);
        unpack4(__shfl_sync(FULL, mylo, src), l[k]);
        unpack4(__shfl_sync(FULL, myhi, src), h[k]);
        pb[k] = __shfl_sync(FULL, mypb, src);
    }
    Bits B;
    B.put(1u << MODE, MODE + 1);
    B.put(part, 6);
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
__device__ __forceinline__ uint4 pack_mode5(uint32_t hm, int base, int lane16, const Res& r, int alo, int ahi, int asel)
{
    const int cs0 = __shfl_sync(FULL, r.sel, base), as0 = __shfl_sync(FULL, asel, base);
    const bool cinv = (cs0 & 2) != 0, ainv = (as0 & 2) != 0;
    const uint64_t cidx = gather_selectors<2>(hm, lane16, cinv ? 3 - r.sel : r.sel);
    const uint64_t aidx = gather_selectors<2>(hm, lane16, ainv ? 3 - asel : asel);
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

// ---- block handlers -------------------------------------------------------------------------------
// hm = mask of the 16 lanes of this block, base = its first lane, lane16 = texel index.

// handle_opaque_block (bc7enc.cpp:2293-2400): mode 6, then mode 1 on the estimated partition.
__device__ __forceinline__ uint4 encode_opaque(uint32_t hm, int base, int lane16, const Px& p, uint32_t packed, const Tables& T)
{
    Res r6;
    color_cell_compression<6, false>(hm, base, p, packed, T, r6);
    uint4 out = pack_mode6(hm, base, lane16, r6);
    if (__any_sync(FULL, r6.err > 0)) {
        const uint32_t part = estimate_partition<false>(hm, base, lane16, p);
        const uint32_t m16 = kShape2[part];
        const uint32_t gm = (((m16 >> lane16) & 1u) ? m16 : (~m16 & 0xFFFFu)) << base;
        Res r1;
        color_cell_compression<1, false>(gm, base, p, packed, T, r1);
        const uint64_t e1 = (uint64_t)__shfl_sync(FULL, r1.err, base) + (uint64_t)__shfl_sync(FULL, r1.err, base + kAnchor2[part]);
        const uint4 o1 = pack_mode17<1>(hm, base, lane16, part, r1);
        if (e1 < (uint64_t)r6.err) out = o1;
    }
    return out;
}

// Alpha half of mode 5 (handle_alpha_block_mode5, bc7enc.cpp:2039-2137): up to 3 passes of
// nearest-of-4 selection and least-squares endpoint refit on the alpha channel alone.
__device__ __forceinline__ uint32_t mode5_alpha(uint32_t hm, const Px& p, int& out_lo, int& out_hi, int& out_sel)
{
    int lo_a = gmin(hm, p.a), hi_a = gmax(hm, p.a);
    out_lo = lo_a;
    out_hi = hi_a;
    out_sel = 0;
    if (!__any_sync(FULL, lo_a != hi_a)) return 0;
    const bool flat = lo_a == hi_a;
    uint32_t best = flat ? 0u : 0xFFFFFFFFu;
    bool live = !flat;
    const int ta = gsum(hm, p.a);
#pragma unroll 1
    for (int pass = 0; pass < 3; pass++) {
        int vals[4];
        vals[0] = lo_a;
        vals[3] = hi_a;
        vals[1] = (lo_a * (64 - 21) + hi_a * 21 + 32) >> 6;
        vals[2] = (lo_a * (64 - 43) + hi_a * 43 + 32) >> 6;
        int s = 0, be = abs(p.a - vals[0]);
#pragma unroll
        for (int j = 1; j < 4; j++) {
            const int e = abs(p.a - vals[j]);
            if (e < be) { be = e; s = j; }
        }
        const uint32_t terr = (uint32_t)gsum(hm, be * be * WA);
        if (live && terr < best) {
            best = terr;
            out_lo = lo_a;
            out_hi = hi_a;
            out_sel = s;
        }
        if (pass != 2) {
            // compute_least_squares_endpoints_a (bc7enc.cpp:410-459) with the 2-bit weights
            const int k = weight_of<4>(s);
            const uint32_t ks = (uint32_t)gsum(hm, (k * k) | (k << 17));
            const int S2 = ks & 0x1FFFF, S1 = ks >> 17;
            const float q00 = (float)gsum(hm, k * p.a) * (1.0f / 64.0f);
            const float q10 = (float)ta - q00;
            const float z00 = (float)S2 * (1.0f / 4096.0f);
            const float z10 = (float)(64 * S1 - S2) * (1.0f / 4096.0f);
            const float z11 = (float)(4096 * 16 - 128 * S1 + S2) * (1.0f / 4096.0f);
            float det = z00 * z11 - z10 * z10;
            if (det != 0.0f) det = 1.0f / det;
            const float xl = (z11 * det) * q00 + (-z10 * det) * q10;
            const float xh = (-z10 * det) * q00 + (z00 * det) * q10;
            int nlo = clampi((int)floorf(xl + .5f), 0, 255), nhi = clampi((int)floorf(xh + .5f), 0, 255);
            if (nlo > nhi) { const int t = nlo; nlo = nhi; nhi = t; }
            if (nlo == lo_a && nhi == hi_a) live = false;
            if (live) { lo_a = nlo; hi_a = nhi; }
            if (!__any_sync(FULL, live)) break;
        }
    }
    return best;
}

// handle_alpha_block (bc7enc.cpp:2139-2291): mode 6 (RGBA), mode 5, mode 7 on the estimated partition.
__device__ __forceinline__ uint4 encode_alpha(uint32_t hm, int base, int lane16, const Px& p, uint32_t packed, const Tables& T)
{
    Res r6;
    color_cell_compression<6, true>(hm, base, p, packed, T, r6);
    uint4 out = pack_mode6(hm, base, lane16, r6);
    uint64_t best = r6.err;
    if (__any_sync(FULL, best > 0)) {
        Res r5;
        color_cell_compression<5, false>(hm, base, p, packed, T, r5);
        int alo, ahi, asel;
        const uint32_t aerr = mode5_alpha(hm, p, alo, ahi, asel);
        const uint64_t e5 = (uint64_t)r5.err + aerr;
        const uint4 o5 = pack_mode5(hm, base, lane16, r5, alo, ahi, asel);
        if (e5 < best) { best = e5; out = o5; }
    }
    if (__any_sync(FULL, best > 0)) {
        const uint32_t part = estimate_partition<true>(hm, base, lane16, p);
        const uint32_t m16 = kShape2[part];
        const uint32_t gm = (((m16 >> lane16) & 1u) ? m16 : (~m16 & 0xFFFFu)) << base;
        Res r7;
        color_cell_compression<7, true>(gm, base, p, packed, T, r7);
        const uint64_t e7 = (uint64_t)__shfl_sync(FULL, r7.err, base) + (uint64_t)__shfl_sync(FULL, r7.err, base + kAnchor2[part]);
        const uint4 o7 = pack_mode17<7>(hm, base, lane16, part, r7);
        if (e7 < best) { best = e7; out = o7; }
    }
    return out;
}

} // namespace bc7
} // namespace hkbcn
