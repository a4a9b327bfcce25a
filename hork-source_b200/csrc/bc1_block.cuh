// bc1_block.cuh -- BC1 colour-block encoder for one 4x4 RGBA8 block (one thread).
//
// Behaviour contract: byte-identical to rgbcx::encode_bc1(level, dst, px, false, false) for the levels whose flag sets
// take the encoder's `!needs_block_error` branch (ThirdParty/bc7enc_rdo/rgbcx.cpp:1692-1726 -> :2263-2604, branch
// :2316-2371) -- the call behind Encode_BC1/2/3 in Hork/Image/ImageEncoders.cpp:138-169:
//   level 0  cEncodeBC1BoundingBoxInt: inset bounding box as initial endpoints (:2030-2085), one least-squares pass
//   level 1  cEncodeBC1Use2DLS: 2-D least squares on the widest channel (:1837-1986), one pass
//   level 2  principal axis (:2086-2174), one pass
//   level 3  principal axis, two passes            level 5 (the Compress* preset): the same -- its extra flag,
//            cEncodeBC1UseFasterMSEEval, is only read by the other branch
//   level 4  cEncodeBC1UseFullMSEEval | cEncodeBC1Use6PowerIters: the error-tracking branch (:2371-2470) in its simplest form --
//            six power iterations, selectors by exhaustive search with the block error (bc1_find_sels4_fullerr, :1047-1081), a
//            least-squares pass is kept only if it lowers the error
// Levels 6..18 and the 3-colour flags add the total-orderings search and the endpoint search rounds of that branch: bc1_hq.cuh.
//
// FP32 discipline: the reference is built without FMA contraction (baseline x86-64,
// SURVEY.md 7), so every float op here is an explicitly rounded __fmul_rn/__fadd_rn/
// __fdiv_rn in the reference's evaluation order; float->int casts truncate (__float2int_rz).
#pragma once
#include <stdint.h>

namespace hkbcn {

// a12 tables, filled by bc1_upload_tables(): optimal single-colour endpoints
// (g_bc1_match5_equals_1 / g_bc1_match6_equals_1, rgbcx.cpp:485-527, cBC1Ideal) packed as
// hi | lo << 8, and the 565 rounding midpoints (g_midpoint5/6, rgbcx.cpp:693-694).
static __device__ uint16_t g_bc1_match5[256];
static __device__ uint16_t g_bc1_match6[256];
static __device__ float g_bc1_mid5[32];
static __device__ float g_bc1_mid6[64];

__device__ __forceinline__ int to_5(uint32_t v) { v = v * 31 + 128; return (int)(((v + (v >> 8)) >> 8) & 0xFF); }
__device__ __forceinline__ int to_6(uint32_t v) { v = v * 63 + 128; return (int)(((v + (v >> 8)) >> 8) & 0xFF); }

struct Bc1Ep { int lr, lg, lb, hr, hg, hb; };

// rgbcx.cpp:915-950 (bc1_find_sels4_noerr) with bc1_get_block_colors4 in cBC1Ideal mode (:869-879).
// Returns the 16 selectors packed 2 bits each (reference selector numbering 0..3 = l, 2/3l, 2/3h, h).
__device__ __forceinline__ uint32_t bc1_find_sels4(const uint32_t px[16], const Bc1Ep& e)
{
    int r0 = (e.lr << 3) | (e.lr >> 2), g0 = (e.lg << 2) | (e.lg >> 4), b0 = (e.lb << 3) | (e.lb >> 2);
    int r3 = (e.hr << 3) | (e.hr >> 2), g3 = (e.hg << 2) | (e.hg >> 4), b3 = (e.hb << 3) | (e.hb >> 2);
    int r1 = (r0 * 2 + r3) / 3, g1 = (g0 * 2 + g3) / 3, b1 = (b0 * 2 + b3) / 3;
    int r2 = (r3 * 2 + r0) / 3, g2 = (g3 * 2 + g0) / 3, b2 = (b3 * 2 + b0) / 3;
    int ar = r3 - r0, ag = g3 - g0, ab = b3 - b0;
    int d0 = r0 * ar + g0 * ag + b0 * ab;
    int d1 = r1 * ar + g1 * ag + b1 * ab;
    int d2 = r2 * ar + g2 * ag + b2 * ab;
    int d3 = r3 * ar + g3 * ag + b3 * ab;
    int t0 = d0 + d1, t1 = d1 + d2, t2 = d2 + d3;
    ar *= 2; ag *= 2; ab *= 2;
    uint32_t sels = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        int r = px[i] & 0xFF, g = (px[i] >> 8) & 0xFF, b = (px[i] >> 16) & 0xFF;
        int d = r * ar + g * ag + b * ab;
        int c = (d <= t0) + (d < t1) + (d < t2); // s_sels = {3,2,1,0}
        sels |= (uint32_t)(3 - c) << (2 * i);
    }
    return sels;
}

// rgbcx.cpp:1047-1081 (bc1_find_sels4_fullerr): the nearest of the four colours per texel (ties: the later one only for
// j = 3), the block's squared error; stops at the texel where the total reaches cur_err (the caller then drops the result).
__device__ __forceinline__ uint32_t bc1_find_sels4_fullerr(const uint32_t px[16], const Bc1Ep& e, uint32_t& sels_out, uint32_t cur_err)
{
    int pr[4], pg[4], pb[4];
    pr[0] = (e.lr << 3) | (e.lr >> 2); pg[0] = (e.lg << 2) | (e.lg >> 4); pb[0] = (e.lb << 3) | (e.lb >> 2);
    pr[3] = (e.hr << 3) | (e.hr >> 2); pg[3] = (e.hg << 2) | (e.hg >> 4); pb[3] = (e.hb << 3) | (e.hb >> 2);
    pr[1] = (pr[0] * 2 + pr[3]) / 3; pg[1] = (pg[0] * 2 + pg[3]) / 3; pb[1] = (pb[0] * 2 + pb[3]) / 3;
    pr[2] = (pr[3] * 2 + pr[0]) / 3; pg[2] = (pg[3] * 2 + pg[0]) / 3; pb[2] = (pb[3] * 2 + pb[0]) / 3;
    uint32_t total = 0, sels = 0;
    bool open = true; // the reference breaks out of the texel loop; here the remaining texels are predicated off
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int r = px[i] & 0xFF, g = (px[i] >> 8) & 0xFF, b = (px[i] >> 16) & 0xFF;
        uint32_t best = (uint32_t)((pr[0] - r) * (pr[0] - r) + (pg[0] - g) * (pg[0] - g) + (pb[0] - b) * (pb[0] - b));
        uint32_t bs = 0;
#pragma unroll
        for (int j = 1; j < 4; j++) {
            const uint32_t err = (uint32_t)((pr[j] - r) * (pr[j] - r) + (pg[j] - g) * (pg[j] - g) + (pb[j] - b) * (pb[j] - b));
            // `(j < 4) && best_err` in the loop header: an exact match ends the scan (later entries cannot be strictly better,
            // and the j == 3 tie rule does not apply once the loop has ended)
            if (best != 0 && (err < best || (err == best && j == 3))) {
                best = err;
                bs = (uint32_t)j;
            }
        }
        if (open) {
            total += best;
            if (total >= cur_err) open = false;
            else sels |= bs << (2 * i);
        }
    }
    sels_out = sels;
    return total;
}

// rgbcx.cpp:1194-1228 (precise_round_565) for one endpoint.
__device__ __forceinline__ void round_565(float x0, float x1, float x2, int& r, int& g, int& b)
{
    r = __float2int_rz(__fmul_rn(x0, 31.0f));
    g = __float2int_rz(__fmul_rn(x1, 63.0f));
    b = __float2int_rz(__fmul_rn(x2, 31.0f));
    if ((uint32_t)r > 31u) r = (~r >> 31) & 31;
    if ((uint32_t)b > 31u) b = (~b >> 31) & 31;
    if ((uint32_t)g > 63u) g = (~g >> 31) & 63;
    r = (r + (x0 > g_bc1_mid5[r])) & 31;
    g = (g + (x1 > g_bc1_mid6[g])) & 63;
    b = (b + (x2 > g_bc1_mid5[b])) & 31;
}

// rgbcx.cpp:629-691 (encode_bc1_solid_block, allow_3color = false)
__device__ __forceinline__ uint2 bc1_solid(uint32_t fr, uint32_t fg, uint32_t fb)
{
    uint32_t mr = g_bc1_match5[fr], mg = g_bc1_match6[fg], mb = g_bc1_match5[fb];
    int max16 = (int)(((mr & 0xFF) << 11) | ((mg & 0xFF) << 5) | (mb & 0xFF));
    int min16 = (int)(((mr >> 8) << 11) | ((mg >> 8) << 5) | (mb >> 8));
    uint32_t mask = 0xAA;
    if (min16 == max16) {
        mask = 0;
        if (min16 > 0) min16--;
        else { max16 = 1; min16 = 0; mask = 0x55; }
    }
    if (max16 < min16) { int t = max16; max16 = min16; min16 = t; mask ^= 0x55; }
    return make_uint2((uint32_t)max16 | ((uint32_t)min16 << 16), mask * 0x01010101u);
}

// rgbcx.cpp:1245-1300 (bc1_encode4): never emits a 3-colour block.
__device__ __forceinline__ uint2 bc1_pack4(const Bc1Ep& e, uint32_t sels)
{
    uint32_t lc16 = ((uint32_t)e.lr << 11) | ((uint32_t)e.lg << 5) | (uint32_t)e.lb;
    uint32_t hc16 = ((uint32_t)e.hr << 11) | ((uint32_t)e.hg << 5) | (uint32_t)e.hb;
    if (lc16 == hc16) {
        uint32_t mask = 0;
        if (hc16 > 0) hc16--;
        else { hc16 = 0; lc16 = 1; mask = 0x55; }
        return make_uint2(lc16 | (hc16 << 16), mask * 0x01010101u);
    }
    uint32_t inv = 0;
    if (lc16 < hc16) { uint32_t t = lc16; lc16 = hc16; hc16 = t; inv = 0x55555555u; }
    // s_sel_trans = {0,2,3,1}: reference selector -> BC1 hardware selector
    uint32_t packed = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t s = (sels >> (2 * i)) & 3;
        packed |= ((0x78u >> (2 * s)) & 3u) << (2 * i); // 0b01'11'10'00
    }
    return make_uint2(lc16 | (hc16 << 16), packed ^ inv);
}

// precise_round_565_noscale (rgbcx.cpp:1230-1243) for one endpoint
__device__ __forceinline__ void round_565_noscale(float x0, float x1, float x2, int& r, int& g, int& b)
{
    round_565(__fmul_rn(x0, 1.0f / 255.0f), __fmul_rn(x1, 1.0f / 255.0f), __fmul_rn(x2, 1.0f / 255.0f), r, g, b);
}

// cEncodeBC1Use2DLS (rgbcx.cpp:1837-1986): regress the two narrower channels on the widest one (x), endpoints at x's range
// inset by 1/16.  c[0] = x, c[1], c[2] = the other two in R, G, B order; the three branches of the reference differ only in
// which channel plays x.
__device__ __forceinline__ void bc1_initial_2dls(const uint32_t px[16], int min_v[3], int max_v[3], const int total[3], Bc1Ep& e)
{
    int big = 0, min_c = min_v[0], max_c = max_v[0];
    if ((max_v[1] - min_v[1]) > (max_c - min_c)) { big = 1; min_c = min_v[1]; max_c = max_v[1]; }
    if ((max_v[2] - min_v[2]) > (max_c - min_c)) { big = 2; min_c = min_v[2]; max_c = max_v[2]; }
    const int o1 = big == 0 ? 1 : 0, o2 = big == 2 ? 1 : 2; // the other two channels, ascending
    int sum_x2 = 0, sum_xy1 = 0, sum_xy2 = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int c0 = px[i] & 0xFF, c1 = (px[i] >> 8) & 0xFF, c2 = (px[i] >> 16) & 0xFF;
        const int x = big == 0 ? c0 : big == 1 ? c1 : c2;
        const int y1 = o1 == 0 ? c0 : c1, y2 = o2 == 1 ? c1 : c2;
        sum_x2 += x * x; sum_xy1 += x * y1; sum_xy2 += x * y2;
    }
    const int sum_x = total[big], t1 = total[o1], t2 = total[o2];
    float div = (float)(16 * sum_x2 - sum_x * sum_x);
    float b1 = 0.0f, b2 = 0.0f;
    if (fabsf(div) > 1e-8f) {
        div = __fdiv_rn(1.0f, div);
        b1 = __fmul_rn((float)(16 * sum_xy1 - sum_x * t1), div);
        b2 = __fmul_rn((float)(16 * sum_xy2 - sum_x * t2), div);
    }
    const float a1 = __fdiv_rn(__fadd_rn((float)t1, -__fmul_rn(b1, (float)sum_x)), 16.0f);
    const float a2 = __fdiv_rn(__fadd_rn((float)t2, -__fmul_rn(b2, (float)sum_x)), 16.0f);
    float l1 = __fadd_rn(a1, __fmul_rn(b1, (float)min_c)), l2 = __fadd_rn(a2, __fmul_rn(b2, (float)min_c));
    float h1 = __fadd_rn(a1, __fmul_rn(b1, (float)max_c)), h2 = __fadd_rn(a2, __fmul_rn(b2, (float)max_c));
    const float d1 = __fadd_rn(h1, -l1), d2 = __fadd_rn(h2, -l2);
    h1 = __fadd_rn(l1, __fmul_rn(d1, 15.0f / 16.0f));
    h2 = __fadd_rn(l2, __fmul_rn(d2, 15.0f / 16.0f));
    l1 = __fadd_rn(l1, __fmul_rn(d1, 1.0f / 16.0f));
    l2 = __fadd_rn(l2, __fmul_rn(d2, 1.0f / 16.0f));
    const float d = (float)(max_c - min_c);
    const float l0 = __fadd_rn((float)min_c, __fmul_rn(d, 1.0f / 16.0f)), h0 = __fadd_rn((float)min_c, __fmul_rn(d, 15.0f / 16.0f));
    float l[3], h[3];
    l[big] = l0; h[big] = h0;
    l[o1] = l1; h[o1] = h1;
    l[o2] = l2; h[o2] = h2;
    round_565_noscale(l[0], l[1], l[2], e.lr, e.lg, e.lb);
    round_565_noscale(h[0], h[1], h[2], e.hr, e.hg, e.hb);
}

// cEncodeBC1BoundingBoxInt (rgbcx.cpp:2030-2085)
__device__ __forceinline__ void bc1_initial_bbox_int(const uint32_t px[16], int min_r, int min_g, int min_b, int max_r, int max_g, int max_b, int avg_r,
                                                     int avg_g, int avg_b, Bc1Ep& e)
{
    const int inset_r = (max_r - min_r - 8) >> 4, inset_g = (max_g - min_g - 8) >> 4, inset_b = (max_b - min_b - 8) >> 4;
    min_r += inset_r; min_g += inset_g; min_b += inset_b;
    if ((uint32_t)(min_r | min_g | min_b) > 255u) {
        min_r = min(max(min_r, 0), 255); min_g = min(max(min_g, 0), 255); min_b = min(max(min_b, 0), 255);
    }
    max_r -= inset_r; max_g -= inset_g; max_b -= inset_b;
    if ((uint32_t)(max_r | max_g | max_b) > 255u) {
        max_r = min(max(max_r, 0), 255); max_g = min(max(max_g, 0), 255); max_b = min(max(max_b, 0), 255);
    }
    int icov_xz = 0, icov_yz = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int r = (int)(px[i] & 0xFF) - avg_r, g = (int)((px[i] >> 8) & 0xFF) - avg_g, b = (int)((px[i] >> 16) & 0xFF) - avg_b;
        icov_xz += r * b;
        icov_yz += g * b;
    }
    int x0 = min_r, y0 = min_g, x1 = max_r, y1 = max_g;
    if (icov_xz < 0) { const int t = x0; x0 = x1; x1 = t; }
    if (icov_yz < 0) { const int t = y0; y0 = y1; y1 = t; }
    e.lr = to_5(x0); e.lg = to_6(y0); e.lb = to_5(min_b);
    e.hr = to_5(x1); e.hg = to_6(y1); e.hb = to_5(max_b);
}

// The colour path of levels 0, 1, 2, 3 and 5 (see the header).  px[i] = R | G<<8 | B<<16 | A<<24 (alpha ignored).
__device__ __forceinline__ uint2 bc1_encode_block(const uint32_t px[16], const int level = 5)
{
    const uint32_t rgb0 = px[0] & 0xFFFFFFu;
    const int fr = rgb0 & 0xFF, fg = (rgb0 >> 8) & 0xFF, fb = rgb0 >> 16;
    bool solid = true;
#pragma unroll
    for (int i = 1; i < 16; i++) solid &= ((px[i] & 0xFFFFFFu) == rgb0);
    if (solid) return bc1_solid(fr, fg, fb);

    int min_r = fr, min_g = fg, min_b = fb, max_r = fr, max_g = fg, max_b = fb;
    int total_r = fr, total_g = fg, total_b = fb;
    bool gray = (fr == fg) && (fr == fb);
#pragma unroll
    for (int i = 1; i < 16; i++) {
        int r = px[i] & 0xFF, g = (px[i] >> 8) & 0xFF, b = (px[i] >> 16) & 0xFF;
        gray &= (r == g) && (r == b);
        max_r = max(max_r, r); max_g = max(max_g, g); max_b = max(max_b, b);
        min_r = min(min_r, r); min_g = min(min_g, g); min_b = min(min_b, b);
        total_r += r; total_g += g; total_b += b;
    }
    const int avg_r = (total_r + 8) >> 4, avg_g = (total_g + 8) >> 4, avg_b = (total_b + 8) >> 4;

    // ---- encode_bc1_pick_initial (rgbcx.cpp:1813-2175)
    Bc1Ep e;
    if (gray) {
        if ((max_r - min_r) < 2) {
            e.lr = e.lb = e.hr = e.hb = to_5(fr);
            e.lg = e.hg = to_6(fr);
        } else {
            e.lr = e.lb = to_5(min_r); e.lg = to_6(min_r);
            e.hr = e.hb = to_5(max_r); e.hg = to_6(max_r);
        }
    } else if (level == 1) {
        int mn[3] = {min_r, min_g, min_b}, mx[3] = {max_r, max_g, max_b};
        const int tot[3] = {total_r, total_g, total_b};
        bc1_initial_2dls(px, mn, mx, tot, e);
    } else if (level == 0) {
        bc1_initial_bbox_int(px, min_r, min_g, min_b, max_r, max_g, max_b, avg_r, avg_g, avg_b, e);
    } else {
        int ic0 = 0, ic1 = 0, ic2 = 0, ic3 = 0, ic4 = 0, ic5 = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            int r = (int)(px[i] & 0xFF) - avg_r, g = (int)((px[i] >> 8) & 0xFF) - avg_g, b = (int)((px[i] >> 16) & 0xFF) - avg_b;
            ic0 += r * r; ic1 += r * g; ic2 += r * b; ic3 += g * g; ic4 += g * b; ic5 += b * b;
        }
        int saxis_r = 306, saxis_g = 601, saxis_b = 117;
        float xr = (float)(max_r - min_r), xg = (float)(max_g - min_g), xb = (float)(max_b - min_b);
        if (ic2 < 0) xr = -xr;
        if (ic4 < 0) xg = -xg;
        const float inv255 = 1.0f / 255.0f;
        const float c0 = __fmul_rn((float)ic0, inv255), c1 = __fmul_rn((float)ic1, inv255), c2 = __fmul_rn((float)ic2, inv255);
        const float c3 = __fmul_rn((float)ic3, inv255), c4 = __fmul_rn((float)ic4, inv255), c5 = __fmul_rn((float)ic5, inv255);
        const int power_iters = level == 4 ? 6 : 4; // cEncodeBC1Use6PowerIters (rgbcx.cpp:2121)
#pragma unroll
        for (int it = 0; it < 6; it++) {
            if (it < power_iters) {
                float r = __fadd_rn(__fadd_rn(__fmul_rn(xr, c0), __fmul_rn(xg, c1)), __fmul_rn(xb, c2));
                float g = __fadd_rn(__fadd_rn(__fmul_rn(xr, c1), __fmul_rn(xg, c3)), __fmul_rn(xb, c4));
                float b = __fadd_rn(__fadd_rn(__fmul_rn(xr, c2), __fmul_rn(xg, c4)), __fmul_rn(xb, c5));
                xr = r; xg = g; xb = b;
            }
        }
        float k = fmaxf(fmaxf(fabsf(xr), fabsf(xg)), fabsf(xb));
        if (k >= 2.0f) {
            float m = __fdiv_rn(2048.0f, k);
            saxis_r = __float2int_rz(__fmul_rn(xr, m));
            saxis_g = __float2int_rz(__fmul_rn(xg, m));
            saxis_b = __float2int_rz(__fmul_rn(xb, m));
        }
        saxis_r = (int)((uint32_t)saxis_r << 4);
        saxis_g = (int)((uint32_t)saxis_g << 4);
        saxis_b = (int)((uint32_t)saxis_b << 4);
        int low_dot = 0x7FFFFFFF, high_dot = (int)0x80000000;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            int r = px[i] & 0xFF, g = (px[i] >> 8) & 0xFF, b = (px[i] >> 16) & 0xFF;
            int dot = ((r * saxis_r + g * saxis_g + b * saxis_b) & ~0xF) + i;
            low_dot = min(low_dot, dot);
            high_dot = max(high_dot, dot);
        }
        uint32_t lc = 0, hc = 0;
        const int low_c = low_dot & 15, high_c = high_dot & 15;
#pragma unroll
        for (int i = 0; i < 16; i++) { // register-array pick without dynamic indexing
            if (i == low_c) lc = px[i];
            if (i == high_c) hc = px[i];
        }
        e.lr = to_5(lc & 0xFF); e.lg = to_6((lc >> 8) & 0xFF); e.lb = to_5((lc >> 16) & 0xFF);
        e.hr = to_5(hc & 0xFF); e.hg = to_6((hc >> 8) & 0xFF); e.hb = to_5((hc >> 16) & 0xFF);
    }

    // level 4: the error-tracking form of the same loop (rgbcx.cpp:2388-2440)
    const bool track = level == 4;
    uint32_t sels, cur_err = 0xFFFFFFFFu;
    if (track) cur_err = bc1_find_sels4_fullerr(px, e, sels, 0xFFFFFFFFu);
    else sels = bc1_find_sels4(px, e);

    // ---- one (levels 0..2) or two least-squares passes (rgbcx.cpp:2333-2370; LS solve :727-773)
    const int passes = level >= 3 ? 2 : 1;
#pragma unroll 1
    for (int pass = 0; pass < passes; pass++) {
        uint32_t wacc = 0, uq_r = 0, uq_g = 0, uq_b = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            uint32_t s = (sels >> (2 * i)) & 3;
            // g_weight_vals4 = {0x000009, 0x010204, 0x040201, 0x090000}
            wacc += s == 0 ? 0x000009u : s == 1 ? 0x010204u : s == 2 ? 0x040201u : 0x090000u;
            uq_r += s * (px[i] & 0xFF);
            uq_g += s * ((px[i] >> 8) & 0xFF);
            uq_b += s * ((px[i] >> 16) & 0xFF);
        }
        const int q_r = total_r * 3 - (int)uq_r, q_g = total_g * 3 - (int)uq_g, q_b = total_b * 3 - (int)uq_b;
        const float z00 = (float)((wacc >> 16) & 0xFF), z10 = (float)((wacc >> 8) & 0xFF), z11 = (float)(wacc & 0xFF);
        float det = __fadd_rn(__fmul_rn(z00, z11), -__fmul_rn(z10, z10));
        Bc1Ep t;
        if (fabsf(det) < 1e-8f) {
            // all selectors equal: treat as a solid block of the average colour (:2341-2350)
            uint32_t mr = g_bc1_match5[avg_r], mg = g_bc1_match6[avg_g], mb = g_bc1_match5[avg_b];
            t.lr = mr & 0xFF; t.lg = mg & 0xFF; t.lb = mb & 0xFF;
            t.hr = mr >> 8; t.hg = mg >> 8; t.hb = mb >> 8;
        } else {
            det = __fdiv_rn(3.0f / 255.0f, det);
            const float iz00 = __fmul_rn(z11, det), iz01 = __fmul_rn(-z10, det), iz10 = iz01, iz11 = __fmul_rn(z00, det);
            const float fr_ = (float)uq_r, fg_ = (float)uq_g, fb_ = (float)uq_b;
            const float qr = (float)q_r, qg = (float)q_g, qb = (float)q_b;
            const float xl0 = __fadd_rn(__fmul_rn(iz00, fr_), __fmul_rn(iz01, qr));
            const float xh0 = __fadd_rn(__fmul_rn(iz10, fr_), __fmul_rn(iz11, qr));
            const float xl1 = __fadd_rn(__fmul_rn(iz00, fg_), __fmul_rn(iz01, qg));
            const float xh1 = __fadd_rn(__fmul_rn(iz10, fg_), __fmul_rn(iz11, qg));
            const float xl2 = __fadd_rn(__fmul_rn(iz00, fb_), __fmul_rn(iz01, qb));
            const float xh2 = __fadd_rn(__fmul_rn(iz10, fb_), __fmul_rn(iz11, qb));
            // call site passes (trial_h*, trial_l*) for (xl, xh): xl rounds into the HIGH endpoint (:2355)
            round_565(xl0, xl1, xl2, t.hr, t.hg, t.hb);
            round_565(xh0, xh1, xh2, t.lr, t.lg, t.lb);
        }
        if (e.lr == t.lr && e.lg == t.lg && e.lb == t.lb && e.hr == t.hr && e.hg == t.hg && e.hb == t.hb) break;
        if (track) {
            uint32_t tsels;
            const uint32_t terr = bc1_find_sels4_fullerr(px, t, tsels, cur_err);
            if (!(terr < cur_err)) break; // the pass did not help: keep what there is
            cur_err = terr;
            sels = tsels;
            e = t;
            continue;
        }
        sels = bc1_find_sels4(px, t);
        e = t;
    }
    return bc1_pack4(e, sels);
}

} // namespace hkbcn
