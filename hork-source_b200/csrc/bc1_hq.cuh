// bc1_hq.cuh -- the error-tracking branch of the BC1 colour-block encoder, for one 4x4 RGBA8 block (one thread).
//
// Behaviour contract: byte-identical to rgbcx::encode_bc1(level, dst, px, allow_3color, allow_transparent_texels_for_black)
// (ThirdParty/bc7enc_rdo/rgbcx.cpp:1692-1811 -> :2263-2604) for every flag set that makes `needs_block_error` true -- levels
// 6..18 of Encode_BC1 / Encode_BC2 / Encode_BC3 and level 5 with either 3-colour flag (Hork/Image/ImageEncoders.cpp:138-169):
//   rounds over the initial endpoints (:2371-2466; principal axis, then the float bounding box when cEncodeBC1TryAllInitialEndponts),
//   selectors with the block error by one of three evaluators (:952-1118), error-checked least-squares passes,
//   the likely-total-orderings search (:2468-2573): sort the texels along the endpoint axis, and for each histogram the table
//     proposes solve the least-squares endpoints from prefix sums (:698-725), keep what lowers the error; twice when cEncodeBC1Iterative,
//   3-colour blocks (:1522-1690) with their own orderings search, 3-colour blocks with transparent black (:1348-1520),
//   the endpoint search rounds (:2177-2261), bc1_encode3 (:1302-1338), the 3-colour solid block (:629-691).
// The fast branch (levels 0..5 without flags) stays in bc1_block.cuh; this one is plain scalar code with its arrays in local
// memory -- these levels trade time for quality in the reference too (level 18 runs 2 x 128 orderings and 256 search rounds).
//
// The orderings tables are the reference library's data; build/bc1_orderings.inc (tools/gen_bc1_orderings.py) carries them
// from the engine tree into the build, bc1_upload_tables() derives the hash and the factors as rgbcx::init does (:591-625).
#pragma once
#include "bc1_block.cuh"

namespace hkbcn {

// rgbcx.h:160-213
enum : uint32_t {
    RF_LIKELY = 2, RF_TWO_LS = 4, RF_BLACK3 = 8, RF_USE3 = 16, RF_ITERATIVE = 32, RF_BBOX = 64, RF_FASTER = 128, RF_FULL = 256,
    RF_2DLS = 512, RF_6POWER = 2048, RF_TRYALL = 16384, RF_BBOX_INT = 32768, RF_ROUNDS_SHIFT = 22, RF_ROUNDS_MASK = 1023u << 22
};
constexpr int HQ_N4 = 969, HQ_N3 = 153, HQ_MAX4 = 128, HQ_MAX3 = 32;

static __device__ uint8_t  g_hq_orders4[HQ_N4 * 4];
static __device__ uint16_t g_hq_best4[HQ_N4 * HQ_MAX4];
static __device__ uint16_t g_hq_hash4[4096];
static __device__ float    g_hq_fac4[HQ_N4 * 3];
static __device__ uint8_t  g_hq_orders3[HQ_N3 * 3];
static __device__ uint8_t  g_hq_best3[HQ_N3 * HQ_MAX3];
static __device__ uint16_t g_hq_hash3[256];
static __device__ float    g_hq_fac3[HQ_N3 * 3];
// single-colour tables with their error: hi | lo << 8 | e << 16 (g_bc1_match5/6_equals_1, g_bc1_match5/6_half, :370-371)
static __device__ uint32_t g_hq_eq5[256], g_hq_eq6[256], g_hq_half5[256], g_hq_half6[256];

// rgbcx::encode_bc1(level, ...) :1692-1811: level -> flags and the number of orderings to try
__host__ __device__ inline uint32_t bc1_level_flags(int level, bool allow3, bool black, uint32_t& to4, uint32_t& to3)
{
    const uint32_t opt = (allow3 ? RF_USE3 : 0u) | (black ? RF_BLACK3 : 0u);
    const uint32_t hi = RF_TWO_LS | RF_FULL | RF_LIKELY | RF_6POWER | RF_TRYALL;
    to4 = 1; to3 = 1;
    switch (level) {
    case 0: return RF_BBOX_INT;
    case 1: return RF_2DLS;
    case 2: return 0;
    case 3: return RF_TWO_LS;
    case 4: return RF_TWO_LS | RF_FULL | RF_6POWER;
    default:
    case 5: return RF_TWO_LS | RF_FASTER | opt;
    case 6: return RF_TWO_LS | RF_FASTER | RF_LIKELY | opt;
    case 7: to4 = 4; return RF_TWO_LS | RF_FASTER | RF_LIKELY | opt;
    case 8: to4 = 8; return RF_TWO_LS | RF_FASTER | RF_LIKELY | opt;
    case 9: to4 = 11; to3 = 3; return RF_TWO_LS | RF_LIKELY | opt;
    case 10: to4 = 20; to3 = 8; return RF_TWO_LS | RF_LIKELY | opt;
    case 11: to4 = 28; to3 = 16; return RF_TWO_LS | RF_LIKELY | opt;
    case 12: to4 = 32; to3 = 32; return RF_TWO_LS | RF_LIKELY | opt;
    case 13: to4 = 32; to3 = 32; return hi | (20u << RF_ROUNDS_SHIFT) | opt;
    case 14: to4 = 32; to3 = 32; return hi | (32u << RF_ROUNDS_SHIFT) | opt;
    case 15: to4 = (((32 + HQ_MAX4) / 2) + 32) / 2; to3 = 32; return hi | (32u << RF_ROUNDS_SHIFT) | opt;
    case 16: to4 = (32 + HQ_MAX4) / 2; to3 = 32; return hi | (256u << RF_ROUNDS_SHIFT) | opt;
    case 17: to4 = HQ_MAX4; to3 = 32; return hi | (256u << RF_ROUNDS_SHIFT) | opt;
    case 18: to4 = HQ_MAX4; to3 = 32; return hi | RF_ITERATIVE | (256u << RF_ROUNDS_SHIFT) | opt;
    }
}

struct HqPx { int r[16], g[16], b[16]; };

__device__ __forceinline__ int hq_sq(int a) { return a * a; }
__device__ __forceinline__ uint32_t hq_sel(uint32_t sels, int i) { return (sels >> (2 * i)) & 3u; }

// bc1_get_block_colors4 / 3, cBC1Ideal (:869-913)
__device__ inline void hq_colors4(const Bc1Ep& e, int cr[4], int cg[4], int cb[4])
{
    cr[0] = (e.lr << 3) | (e.lr >> 2); cg[0] = (e.lg << 2) | (e.lg >> 4); cb[0] = (e.lb << 3) | (e.lb >> 2);
    cr[3] = (e.hr << 3) | (e.hr >> 2); cg[3] = (e.hg << 2) | (e.hg >> 4); cb[3] = (e.hb << 3) | (e.hb >> 2);
    cr[1] = (cr[0] * 2 + cr[3]) / 3; cg[1] = (cg[0] * 2 + cg[3]) / 3; cb[1] = (cb[0] * 2 + cb[3]) / 3;
    cr[2] = (cr[3] * 2 + cr[0]) / 3; cg[2] = (cg[3] * 2 + cg[0]) / 3; cb[2] = (cb[3] * 2 + cb[0]) / 3;
}
__device__ inline void hq_colors3(const Bc1Ep& e, int cr[3], int cg[3], int cb[3])
{
    cr[0] = (e.lr << 3) | (e.lr >> 2); cg[0] = (e.lg << 2) | (e.lg >> 4); cb[0] = (e.lb << 3) | (e.lb >> 2);
    cr[1] = (e.hr << 3) | (e.hr >> 2); cg[1] = (e.hg << 2) | (e.hg >> 4); cb[1] = (e.hb << 3) | (e.hb >> 2);
    cr[2] = (cr[0] + cr[1]) / 2; cg[2] = (cg[0] + cg[1]) / 2; cb[2] = (cb[0] + cb[1]) / 2;
}

// ---- the three selector evaluators (:952-1118).  Each returns the block error and stops adding once it reaches cur_err
// (the caller then drops the selectors, so what they hold after an early stop does not matter).
__device__ inline uint32_t hq_find4_fasterr(const HqPx& p, const Bc1Ep& e, uint32_t& sels_out, uint32_t cur_err)
{
    int cr[4], cg[4], cb[4];
    hq_colors4(e, cr, cg, cb);
    int ar = cr[3] - cr[0], ag = cg[3] - cg[0], ab = cb[3] - cb[0];
    int dots[4];
    for (int i = 0; i < 4; i++) dots[i] = cr[i] * ar + cg[i] * ag + cb[i] * ab;
    const int t0 = dots[0] + dots[1], t1 = dots[1] + dots[2], t2 = dots[2] + dots[3];
    ar *= 2; ag *= 2; ab *= 2;
    uint32_t total = 0, sels = 0;
    for (int i = 0; i < 16; i += 4) {
        for (int k = i; k < i + 4; k++) {
            const int d = p.r[k] * ar + p.g[k] * ag + p.b[k] * ab;
            const int s = 3 - ((d <= t0) + (d < t1) + (d < t2));
            sels |= (uint32_t)s << (2 * k);
            total += hq_sq(p.r[k] - cr[s]) + hq_sq(p.g[k] - cg[s]) + hq_sq(p.b[k] - cb[s]);
        }
        if (total >= cur_err) break; // checked once per row of four
    }
    sels_out = sels;
    return total;
}

__device__ inline uint32_t hq_find4_check2(const HqPx& p, const Bc1Ep& e, uint32_t& sels_out, uint32_t cur_err)
{
    int cr[4], cg[4], cb[4];
    hq_colors4(e, cr, cg, cb);
    const int dr = cr[3] - cr[0], dg = cg[3] - cg[0], db = cb[3] - cb[0];
    const float f = 4.0f / ((float)(hq_sq(dr) + hq_sq(dg) + hq_sq(db)) + .00000125f);
    uint32_t total = 0, sels = 0;
    for (int i = 0; i < 16; i++) {
        const int r = p.r[i], g = p.g[i], b = p.b[i];
        int sel = (int)((float)((r - cr[0]) * dr + (g - cg[0]) * dg + (b - cb[0]) * db) * f + .5f);
        sel = sel < 1 ? 1 : sel > 3 ? 3 : sel;
        const uint32_t err0 = hq_sq(cr[sel - 1] - r) + hq_sq(cg[sel - 1] - g) + hq_sq(cb[sel - 1] - b);
        const uint32_t err1 = hq_sq(cr[sel] - r) + hq_sq(cg[sel] - g) + hq_sq(cb[sel] - b);
        int best_sel = sel;
        uint32_t best_err = err1;
        if (err0 == err1) {
            if (best_sel - 1 == 0) best_sel = 0; // prefer the endpoint to the interpolant
        } else if (err0 < best_err) {
            best_sel = sel - 1;
            best_err = err0;
        }
        total += best_err;
        if (total >= cur_err) break;
        sels |= (uint32_t)best_sel << (2 * i);
    }
    sels_out = sels;
    return total;
}

__device__ inline uint32_t hq_find4_fullerr(const HqPx& p, const Bc1Ep& e, uint32_t& sels_out, uint32_t cur_err)
{
    int cr[4], cg[4], cb[4];
    hq_colors4(e, cr, cg, cb);
    uint32_t total = 0, sels = 0;
    for (int i = 0; i < 16; i++) {
        const int r = p.r[i], g = p.g[i], b = p.b[i];
        uint32_t best = hq_sq(cr[0] - r) + hq_sq(cg[0] - g) + hq_sq(cb[0] - b);
        uint32_t bs = 0;
        for (int j = 1; j < 4 && best; j++) {
            const uint32_t err = hq_sq(cr[j] - r) + hq_sq(cg[j] - g) + hq_sq(cb[j] - b);
            if (err < best || (err == best && j == 3)) { best = err; bs = j; }
        }
        total += best;
        if (total >= cur_err) break;
        sels |= bs << (2 * i);
    }
    sels_out = sels;
    return total;
}

__device__ inline uint32_t hq_find4(uint32_t flags, const HqPx& p, const Bc1Ep& e, uint32_t& sels, uint32_t cur_err)
{
    if (flags & RF_FASTER) return hq_find4_fasterr(p, e, sels, cur_err);
    if (flags & RF_FULL) return hq_find4_fullerr(p, e, sels, cur_err);
    return hq_find4_check2(p, e, sels, cur_err);
}

// bc1_find_sels3_fullerr (:1120-1192): selectors in BC1's own 3-colour numbering (0 = l, 1 = h, 2 = their mean, 3 = black)
__device__ inline uint32_t hq_find3(bool use_black, const HqPx& p, const Bc1Ep& e, uint32_t& sels_out, uint32_t cur_err)
{
    int cr[3], cg[3], cb[3];
    hq_colors3(e, cr, cg, cb);
    uint32_t total = 0, sels = 0;
    for (int i = 0; i < 16; i++) {
        const int r = p.r[i], g = p.g[i], b = p.b[i];
        uint32_t best = hq_sq(cr[0] - r) + hq_sq(cg[0] - g) + hq_sq(cb[0] - b), bs = 0;
        const uint32_t e1 = hq_sq(cr[1] - r) + hq_sq(cg[1] - g) + hq_sq(cb[1] - b);
        if (e1 < best) { best = e1; bs = 1; }
        const uint32_t e2 = hq_sq(cr[2] - r) + hq_sq(cg[2] - g) + hq_sq(cb[2] - b);
        if (e2 < best) { best = e2; bs = 2; }
        if (use_black) {
            const uint32_t e3 = hq_sq(r) + hq_sq(g) + hq_sq(b);
            if (e3 < best) { best = e3; bs = 3; }
        }
        total += best;
        if (total >= cur_err) break;
        sels |= bs << (2 * i);
    }
    sels_out = sels;
    return total;
}

// precise_round_565(xl, xh, ...) for both endpoints.  Every call site of this branch passes (trial_h*, trial_l*) for the
// (l, h) outputs, so xl rounds into the HIGH endpoint and xh into the LOW one (:2407, :2547, :1407, :1559, :1649).
__device__ inline void hq_round_swapped(const float xl[3], const float xh[3], Bc1Ep& t)
{
    round_565(xl[0], xl[1], xl[2], t.hr, t.hg, t.hb);
    round_565(xh[0], xh[1], xh[2], t.lr, t.lg, t.lb);
}
__device__ inline void hq_single4(int avg_r, int avg_g, int avg_b, Bc1Ep& t)
{ // trial_l* = g_bc1_match*_equals_1[avg].m_hi, trial_h* = .m_lo (:2394-2401)
    const uint32_t mr = g_hq_eq5[avg_r], mg = g_hq_eq6[avg_g], mb = g_hq_eq5[avg_b];
    t.lr = mr & 0xFF; t.lg = mg & 0xFF; t.lb = mb & 0xFF;
    t.hr = (mr >> 8) & 0xFF; t.hg = (mg >> 8) & 0xFF; t.hb = (mb >> 8) & 0xFF;
}
__device__ inline void hq_single3(int avg_r, int avg_g, int avg_b, Bc1Ep& t)
{
    const uint32_t mr = g_hq_half5[avg_r], mg = g_hq_half6[avg_g], mb = g_hq_half5[avg_b];
    t.lr = mr & 0xFF; t.lg = mg & 0xFF; t.lb = mb & 0xFF;
    t.hr = (mr >> 8) & 0xFF; t.hg = (mg >> 8) & 0xFF; t.hb = (mb >> 8) & 0xFF;
}
__device__ __forceinline__ bool hq_same(const Bc1Ep& a, const Bc1Ep& b)
{
    return a.lr == b.lr && a.lg == b.lg && a.lb == b.lb && a.hr == b.hr && a.hg == b.hg && a.hb == b.hb;
}

// compute_least_squares_endpoints4_rgb from selectors (:727-773); false when all selectors are equal
__device__ inline bool hq_ls4(const HqPx& p, uint32_t sels, int total_r, int total_g, int total_b, float xl[3], float xh[3])
{
    uint32_t uq_r = 0, uq_g = 0, uq_b = 0, wacc = 0;
    for (int i = 0; i < 16; i++) {
        const uint32_t s = hq_sel(sels, i);
        wacc += s == 0 ? 0x000009u : s == 1 ? 0x010204u : s == 2 ? 0x040201u : 0x090000u;
        uq_r += s * (uint32_t)p.r[i]; uq_g += s * (uint32_t)p.g[i]; uq_b += s * (uint32_t)p.b[i];
    }
    const int q_r = total_r * 3 - (int)uq_r, q_g = total_g * 3 - (int)uq_g, q_b = total_b * 3 - (int)uq_b;
    const float z00 = (float)((wacc >> 16) & 0xFF), z10 = (float)((wacc >> 8) & 0xFF), z11 = (float)(wacc & 0xFF), z01 = z10;
    float det = z00 * z11 - z01 * z10;
    if (fabsf(det) < 1e-8f) return false;
    det = (3.0f / 255.0f) / det;
    const float iz00 = z11 * det, iz01 = -z01 * det, iz10 = -z10 * det, iz11 = z00 * det;
    xl[0] = iz00 * (float)uq_r + iz01 * (float)q_r; xh[0] = iz10 * (float)uq_r + iz11 * (float)q_r;
    xl[1] = iz00 * (float)uq_g + iz01 * (float)q_g; xh[1] = iz10 * (float)uq_g + iz11 * (float)q_g;
    xl[2] = iz00 * (float)uq_b + iz01 * (float)q_b; xh[2] = iz10 * (float)uq_b + iz11 * (float)q_b;
    return true;
}

// ---- the orderings search's shared pieces
// texels sorted along the axis between the two endpoint colours, then prefix sums of the sorted colours (:2487-2522, :1595-1623)
struct HqSums { uint32_t r[17], g[17], b[17]; };
__device__ inline void hq_sorted_sums(const HqPx& p, const Bc1Ep& e, int total_r, int total_g, int total_b, HqSums& S)
{
    const int r0 = (e.lr << 3) | (e.lr >> 2), g0 = (e.lg << 2) | (e.lg >> 4), b0 = (e.lb << 3) | (e.lb >> 2);
    const int r3 = (e.hr << 3) | (e.hr >> 2), g3 = (e.hg << 2) | (e.hg >> 4), b3 = (e.hb << 3) | (e.hb >> 2);
    const int ar = r3 - r0, ag = g3 - g0, ab = b3 - b0;
    int dots[16];
    for (int i = 0; i < 16; i++) {
        const int d = 0x1000000 + (p.r[i] * ar + p.g[i] * ag + p.b[i] * ab);
        dots[i] = (int)((uint32_t)d << 4) + i;
    }
    for (int i = 1; i < 16; i++) { // the keys are distinct (texel index in the low bits): any sort gives std::sort's order
        const int v = dots[i];
        int j = i - 1;
        while (j >= 0 && dots[j] > v) { dots[j + 1] = dots[j]; j--; }
        dots[j + 1] = v;
    }
    uint32_t r = 0, g = 0, b = 0;
    for (int i = 0; i < 16; i++) {
        const int q = dots[i] & 15;
        S.r[i] = r; S.g[i] = g; S.b[i] = b;
        r += (uint32_t)p.r[q]; g += (uint32_t)p.g[q]; b += (uint32_t)p.b[q];
    }
    S.r[16] = (uint32_t)total_r; S.g[16] = (uint32_t)total_g; S.b[16] = (uint32_t)total_b;
}

// compute_least_squares_endpoints4_rgb from a histogram and the prefix sums (:698-725)
__device__ inline void hq_ls4_sums(uint32_t s, const HqSums& S, int total_r, int total_g, int total_b, float xl[3], float xh[3])
{
    const float iz00 = g_hq_fac4[3 * s + 0], iz10 = g_hq_fac4[3 * s + 1], iz11 = g_hq_fac4[3 * s + 2], iz01 = iz10;
    const uint32_t f1 = g_hq_orders4[4 * s + 0], f2 = f1 + g_hq_orders4[4 * s + 1], f3 = f2 + g_hq_orders4[4 * s + 2];
    const uint32_t uq_r = (S.r[f2] - S.r[f1]) + (S.r[f3] - S.r[f2]) * 2 + (S.r[16] - S.r[f3]) * 3;
    const uint32_t uq_g = (S.g[f2] - S.g[f1]) + (S.g[f3] - S.g[f2]) * 2 + (S.g[16] - S.g[f3]) * 3;
    const uint32_t uq_b = (S.b[f2] - S.b[f1]) + (S.b[f3] - S.b[f2]) * 2 + (S.b[16] - S.b[f3]) * 3;
    const float q_r = (float)((uint32_t)(total_r * 3) - uq_r), q_g = (float)((uint32_t)(total_g * 3) - uq_g), q_b = (float)((uint32_t)(total_b * 3) - uq_b);
    xl[0] = iz00 * (float)uq_r + iz01 * q_r; xh[0] = iz10 * (float)uq_r + iz11 * q_r;
    xl[1] = iz00 * (float)uq_g + iz01 * q_g; xh[1] = iz10 * (float)uq_g + iz11 * q_g;
    xl[2] = iz00 * (float)uq_b + iz01 * q_b; xh[2] = iz10 * (float)uq_b + iz11 * q_b;
}
// compute_least_squares_endpoints3_rgb from a histogram (:775-802): BC1's 3-colour order is l, h, mean
__device__ inline void hq_ls3_sums(uint32_t s, const HqSums& S, int total_r, int total_g, int total_b, float xl[3], float xh[3])
{
    const float iz00 = g_hq_fac3[3 * s + 0], iz10 = g_hq_fac3[3 * s + 1], iz11 = g_hq_fac3[3 * s + 2], iz01 = iz10;
    const uint32_t f1 = g_hq_orders3[3 * s + 0], f2 = f1 + g_hq_orders3[3 * s + 2];
    const uint32_t uq_r = (S.r[16] - S.r[f2]) * 2 + (S.r[f2] - S.r[f1]);
    const uint32_t uq_g = (S.g[16] - S.g[f2]) * 2 + (S.g[f2] - S.g[f1]);
    const uint32_t uq_b = (S.b[16] - S.b[f2]) * 2 + (S.b[f2] - S.b[f1]);
    const float q_r = (float)((uint32_t)(total_r * 2) - uq_r), q_g = (float)((uint32_t)(total_g * 2) - uq_g), q_b = (float)((uint32_t)(total_b * 2) - uq_b);
    xl[0] = iz00 * (float)uq_r + iz01 * q_r; xh[0] = iz10 * (float)uq_r + iz11 * q_r;
    xl[1] = iz00 * (float)uq_g + iz01 * q_g; xh[1] = iz10 * (float)uq_g + iz11 * q_g;
    xl[2] = iz00 * (float)uq_b + iz01 * q_b; xh[2] = iz10 * (float)uq_b + iz11 * q_b;
}
// compute_least_squares_endpoints3_rgb from selectors (:804-867); with use_black the near-black texels and selector 3 stay out
__device__ inline bool hq_ls3(bool use_black, const HqPx& p, uint32_t sels, float xl[3], float xh[3])
{
    int uq_r = 0, uq_g = 0, uq_b = 0, total_r = 0, total_g = 0, total_b = 0;
    uint32_t wacc = 0;
    for (int i = 0; i < 16; i++) {
        const int r = p.r[i], g = p.g[i], b = p.b[i];
        if (use_black && (r | g | b) < 4) continue;
        const uint32_t sel = hq_sel(sels, i);
        if (sel == 3) continue;
        wacc += sel == 0 ? 0x000004u : sel == 1 ? 0x040000u : 0x010101u;
        const int tsel = sel == 0 ? 0 : sel == 1 ? 2 : 1;
        uq_r += tsel * r; uq_g += tsel * g; uq_b += tsel * b;
        total_r += r; total_g += g; total_b += b;
    }
    const int q_r = total_r * 2 - uq_r, q_g = total_g * 2 - uq_g, q_b = total_b * 2 - uq_b;
    const float z00 = (float)((wacc >> 16) & 0xFF), z10 = (float)((wacc >> 8) & 0xFF), z11 = (float)(wacc & 0xFF), z01 = z10;
    float det = z00 * z11 - z01 * z10;
    if (fabsf(det) < 1e-8f) return false;
    det = (2.0f / 255.0f) / det;
    const float iz00 = z11 * det, iz01 = -z01 * det, iz10 = -z10 * det, iz11 = z00 * det;
    xl[0] = iz00 * (float)uq_r + iz01 * (float)q_r; xh[0] = iz10 * (float)uq_r + iz11 * (float)q_r;
    xl[1] = iz00 * (float)uq_g + iz01 * (float)q_g; xh[1] = iz10 * (float)uq_g + iz11 * (float)q_g;
    xl[2] = iz00 * (float)uq_b + iz01 * (float)q_b; xh[2] = iz10 * (float)uq_b + iz11 * (float)q_b;
    return true;
}

// ---- encode_bc1_pick_initial (:1813-2175) for the flag sets of this branch
struct HqStats { int min_r, min_g, min_b, max_r, max_g, max_b, avg_r, avg_g, avg_b, total_r, total_g, total_b; bool gray; };

__device__ inline void hq_pick_initial(const uint32_t px[16], const HqPx& p, uint32_t flags, const HqStats& S, Bc1Ep& e)
{
    if (S.gray) {
        const int fr = p.r[0];
        if ((S.max_r - S.min_r) < 2) {
            e.lr = e.lb = e.hr = e.hb = to_5(fr);
            e.lg = e.hg = to_6(fr);
        } else {
            e.lr = e.lb = to_5(S.min_r); e.lg = to_6(S.min_r);
            e.hr = e.hb = to_5(S.max_r); e.hg = to_6(S.max_r);
        }
    } else if (flags & RF_2DLS) {
        int mn[3] = {S.min_r, S.min_g, S.min_b}, mx[3] = {S.max_r, S.max_g, S.max_b};
        const int tot[3] = {S.total_r, S.total_g, S.total_b};
        bc1_initial_2dls(px, mn, mx, tot, e);
    } else if (flags & RF_BBOX) {
        // the float bounding box (:1987-2029, icbc's compress_dxt1_fast)
        float l[3] = {(float)S.min_r * (1.0f / 255.0f), (float)S.min_g * (1.0f / 255.0f), (float)S.min_b * (1.0f / 255.0f)};
        float h[3] = {(float)S.max_r * (1.0f / 255.0f), (float)S.max_g * (1.0f / 255.0f), (float)S.max_b * (1.0f / 255.0f)};
        const float bias = 8.0f / 255.0f;
        for (int c = 0; c < 3; c++) {
            const float inset = (h[c] - l[c] - bias) * (1.0f / 16.0f);
            float lo = l[c] + inset, hi = h[c] - inset;
            lo = lo < 0.0f ? 0.0f : lo > 1.0f ? 1.0f : lo;
            hi = hi < 0.0f ? 0.0f : hi > 1.0f ? 1.0f : hi;
            l[c] = lo; h[c] = hi;
        }
        int icov_xz = 0, icov_yz = 0;
        for (int i = 0; i < 16; i++) {
            const int r = p.r[i] - S.avg_r, g = p.g[i] - S.avg_g, b = p.b[i] - S.avg_b;
            icov_xz += r * b;
            icov_yz += g * b;
        }
        if (icov_xz < 0) { const float t = l[0]; l[0] = h[0]; h[0] = t; }
        if (icov_yz < 0) { const float t = l[1]; l[1] = h[1]; h[1] = t; }
        round_565(l[0], l[1], l[2], e.lr, e.lg, e.lb); // here the reference passes (l, h) in order
        round_565(h[0], h[1], h[2], e.hr, e.hg, e.hb);
    } else if (flags & RF_BBOX_INT) {
        bc1_initial_bbox_int(px, S.min_r, S.min_g, S.min_b, S.max_r, S.max_g, S.max_b, S.avg_r, S.avg_g, S.avg_b, e);
    } else {
        // two texels along the principal axis (:2086-2174)
        int ic[6] = {0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 16; i++) {
            const int r = p.r[i] - S.avg_r, g = p.g[i] - S.avg_g, b = p.b[i] - S.avg_b;
            ic[0] += r * r; ic[1] += r * g; ic[2] += r * b; ic[3] += g * g; ic[4] += g * b; ic[5] += b * b;
        }
        int saxis_r = 306, saxis_g = 601, saxis_b = 117;
        float xr = (float)(S.max_r - S.min_r), xg = (float)(S.max_g - S.min_g), xb = (float)(S.max_b - S.min_b);
        if (ic[2] < 0) xr = -xr;
        if (ic[4] < 0) xg = -xg;
        float cov[6];
        for (int i = 0; i < 6; i++) cov[i] = (float)ic[i] * (1.0f / 255.0f);
        const int iters = (flags & RF_6POWER) ? 6 : 4;
        for (int it = 0; it < iters; it++) {
            const float r = xr * cov[0] + xg * cov[1] + xb * cov[2];
            const float g = xr * cov[1] + xg * cov[3] + xb * cov[4];
            const float b = xr * cov[2] + xg * cov[4] + xb * cov[5];
            xr = r; xg = g; xb = b;
        }
        const float k = fmaxf(fmaxf(fabsf(xr), fabsf(xg)), fabsf(xb));
        if (k >= 2.0f) {
            const float m = 2048.0f / k;
            saxis_r = (int)(xr * m); saxis_g = (int)(xg * m); saxis_b = (int)(xb * m);
        }
        saxis_r = (int)((uint32_t)saxis_r << 4); saxis_g = (int)((uint32_t)saxis_g << 4); saxis_b = (int)((uint32_t)saxis_b << 4);
        int low_dot = 0x7FFFFFFF, high_dot = (int)0x80000000;
        for (int i = 0; i < 16; i++) {
            const int dot = ((p.r[i] * saxis_r + p.g[i] * saxis_g + p.b[i] * saxis_b) & ~0xF) + i;
            low_dot = min(low_dot, dot);
            high_dot = max(high_dot, dot);
        }
        const int lc = low_dot & 15, hc = high_dot & 15;
        e.lr = to_5(p.r[lc]); e.lg = to_6(p.g[lc]); e.lb = to_5(p.b[lc]);
        e.hr = to_5(p.r[hc]); e.hg = to_6(p.g[hc]); e.hb = to_5(p.b[hc]);
    }
}

struct HqResult { Bc1Ep e; uint32_t sels; bool three; };

// histogram of the selectors -> index of that total ordering (hist4 / hist3::lookup_total_ordering_index, :303-316, :348-359)
__device__ inline uint32_t hq_order_index4(uint32_t sels)
{
    int h[4] = {0, 0, 0, 0};
    for (int i = 0; i < 16; i++) h[hq_sel(sels, i)]++;
    if (h[0] == 16) return 15;
    if (h[1] == 16) return 700;
    if (h[2] == 16) return 753;
    if (h[3] == 16) return 515;
    return g_hq_hash4[h[0] | (h[1] << 4) | (h[2] << 8)];
}
__device__ inline uint32_t hq_order_index3(uint32_t sels)
{
    int h[4] = {0, 0, 0, 0};
    for (int i = 0; i < 16; i++) h[hq_sel(sels, i)]++;
    if (h[0] == 16) return 12;
    if (h[1] == 16) return 15;
    if (h[2] == 16) return 89;
    return g_hq_hash3[h[0] | (h[1] << 4)];
}

// the error-checked least-squares passes of a 3-colour trial (:1396-1433, :1547-1584): shared by both 3-colour searches
__device__ inline void hq_ls3_passes(bool use_black, const HqPx& p, uint32_t flags, int avg_r, int avg_g, int avg_b, Bc1Ep& e, uint32_t& sels,
                                     uint32_t& err)
{
    const int passes = (flags & RF_TWO_LS) ? 2 : 1;
    for (int t = 0; t < passes; t++) {
        float xl[3], xh[3];
        Bc1Ep e2;
        if (!hq_ls3(use_black, p, sels, xl, xh)) hq_single3(avg_r, avg_g, avg_b, e2);
        else hq_round_swapped(xl, xh, e2);
        if (hq_same(e, e2)) break;
        uint32_t sels2;
        const uint32_t err2 = hq_find3(use_black, p, e2, sels2, err);
        if (err2 < err) { err = err2; e = e2; sels = sels2; }
        else break;
    }
}

// try_3color_block_useblack (:1348-1520): near-black texels (r | g | b < 4) take the transparent-black selector
__device__ inline void hq_try3_black(const HqPx& p, uint32_t flags, uint32_t& cur_err, HqResult& R)
{
    int total_r = 0, total_g = 0, total_b = 0, max_r = 0, max_g = 0, max_b = 0, min_r = 255, min_g = 255, min_b = 255, n = 0;
    for (int i = 0; i < 16; i++) {
        const int r = p.r[i], g = p.g[i], b = p.b[i];
        if ((r | g | b) < 4) continue;
        max_r = max(max_r, r); max_g = max(max_g, g); max_b = max(max_b, b);
        min_r = min(min_r, r); min_g = min(min_g, g); min_b = min(min_b, b);
        total_r += r; total_g += g; total_b += b;
        n++;
    }
    if (!n) return;
    const int half = n >> 1;
    const int avg_r = (total_r + half) / n, avg_g = (total_g + half) / n, avg_b = (total_b + half) / n;
    int ic[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 16; i++) {
        int r = p.r[i], g = p.g[i], b = p.b[i];
        if ((r | g | b) < 4) continue;
        r -= avg_r; g -= avg_g; b -= avg_b;
        ic[0] += r * r; ic[1] += r * g; ic[2] += r * b; ic[3] += g * g; ic[4] += g * b; ic[5] += b * b;
    }
    float cov[6];
    for (int i = 0; i < 6; i++) cov[i] = (float)ic[i] * (1.0f / 255.0f);
    float xr = (float)(max_r - min_r), xg = (float)(max_g - min_g), xb = (float)(max_b - min_b);
    if (ic[2] < 0) xr = -xr;
    if (ic[4] < 0) xg = -xg;
    for (int it = 0; it < 4; it++) {
        const float r = xr * cov[0] + xg * cov[1] + xb * cov[2];
        const float g = xr * cov[1] + xg * cov[3] + xb * cov[4];
        const float b = xr * cov[2] + xg * cov[4] + xb * cov[5];
        xr = r; xg = g; xb = b;
    }
    const float k = fmaxf(fmaxf(fabsf(xr), fabsf(xg)), fabsf(xb));
    int saxis_r = 306, saxis_g = 601, saxis_b = 117;
    if (k >= 2.0f) {
        const float m = 1024.0f / k;
        saxis_r = (int)(xr * m); saxis_g = (int)(xg * m); saxis_b = (int)(xb * m);
    }
    int low_dot = 0x7FFFFFFF, high_dot = (int)0x80000000, low_c = 0, high_c = 0;
    for (int i = 0; i < 16; i++) {
        const int r = p.r[i], g = p.g[i], b = p.b[i];
        if ((r | g | b) < 4) continue;
        const int dot = r * saxis_r + g * saxis_g + b * saxis_b;
        if (dot < low_dot) { low_dot = dot; low_c = i; }
        if (dot > high_dot) { high_dot = dot; high_c = i; }
    }
    Bc1Ep e;
    e.lr = to_5(p.r[low_c]); e.lg = to_6(p.g[low_c]); e.lb = to_5(p.b[low_c]);
    e.hr = to_5(p.r[high_c]); e.hg = to_6(p.g[high_c]); e.hb = to_5(p.b[high_c]);
    uint32_t sels, err = hq_find3(true, p, e, sels, 0xFFFFFFFFu);
    if (err) hq_ls3_passes(true, p, flags, avg_r, avg_g, avg_b, e, sels, err);
    if (err < cur_err) {
        R.three = true; R.e = e; R.sels = sels;
        cur_err = err;
    }
}

// try_3color_block (:1522-1690): starts from the initial endpoints of the 4-colour search
__device__ inline void hq_try3(const HqPx& p, uint32_t flags, uint32_t& cur_err, const HqStats& S, Bc1Ep e, uint32_t to3, HqResult& R)
{
    uint32_t sels, err = hq_find3(false, p, e, sels, 0xFFFFFFFFu);
    if (err) hq_ls3_passes(false, p, flags, S.avg_r, S.avg_g, S.avg_b, e, sels, err);
    if (err && (flags & RF_LIKELY) && to3) {
        const uint32_t base = hq_order_index3(sels);
        HqSums sums;
        hq_sorted_sums(p, e, S.total_r, S.total_g, S.total_b, sums);
        const uint32_t q_total = min(to3, (uint32_t)HQ_MAX3);
        for (uint32_t q = 0; q < q_total; q++) {
            const uint32_t s = g_hq_best3[base * HQ_MAX3 + q];
            Bc1Ep t;
            if (s == 12 || s == 15 || s == 89) hq_single3(S.avg_r, S.avg_g, S.avg_b, t);
            else {
                float xl[3], xh[3];
                hq_ls3_sums(s, sums, S.total_r, S.total_g, S.total_b, xl, xh);
                hq_round_swapped(xl, xh, t);
            }
            uint32_t sels2;
            const uint32_t err2 = hq_find3(false, p, t, sels2, 0xFFFFFFFFu);
            if (err2 < err) { err = err2; e = t; sels = sels2; }
        }
    }
    if (err < cur_err) {
        R.three = true; R.e = e; R.sels = sels;
        cur_err = err;
    }
}

// encode_bc1_endpoint_search (:2198-2261, icbc's high-quality mode): step one endpoint to a neighbouring 565 voxel
__device__ inline void hq_endpoint_search(const HqPx& p, bool any_black, uint32_t flags, HqResult& R, uint32_t cur_err)
{
    // s_adjacent_voxels (:2177-2195): dr, dg, db, index of the opposite step
    const int8_t adj[16][4] = {{1, 0, 0, 3}, {0, 1, 0, 4}, {0, 0, 1, 5}, {-1, 0, 0, 0}, {0, -1, 0, 1}, {0, 0, -1, 2}, {1, 1, 0, 9}, {1, 0, 1, 10},
                               {0, 1, 1, 11}, {-1, -1, 0, 6}, {-1, 0, -1, 7}, {0, -1, -1, 8}, {-1, 1, 0, 13}, {1, -1, 0, 12}, {0, -1, 1, 15}, {0, 1, -1, 14}};
    int prev_improvement = 0, forbidden = -1;
    const int rounds = (int)((flags & RF_ROUNDS_MASK) >> RF_ROUNDS_SHIFT);
    for (int i = 0; i < rounds; i++) {
        if (forbidden == (i & 31)) continue;
        const int d0 = adj[i & 15][0], d1 = adj[i & 15][1], d2 = adj[i & 15][2];
        Bc1Ep t = R.e;
        if ((i >> 4) & 1) {
            t.lr = min(max(t.lr + d0, 0), 31); t.lg = min(max(t.lg + d1, 0), 63); t.lb = min(max(t.lb + d2, 0), 31);
        } else {
            t.hr = min(max(t.hr + d0, 0), 31); t.hg = min(max(t.hg + d1, 0), 63); t.hb = min(max(t.hb + d2, 0), 31);
        }
        uint32_t tsels, terr;
        if (R.three) terr = hq_find3(any_black && (flags & RF_BLACK3) != 0, p, t, tsels, cur_err);
        else terr = hq_find4(flags, p, t, tsels, cur_err);
        if (terr < cur_err) {
            cur_err = terr;
            forbidden = adj[i & 15][3] | (i & 16);
            R.e = t;
            R.sels = tsels;
            prev_improvement = i;
        }
        if (i - prev_improvement > 32) break;
    }
}

// encode_bc1_solid_block with allow_3color (:629-691)
__device__ inline uint2 hq_solid(uint32_t fr, uint32_t fg, uint32_t fb, bool allow3)
{
    if (allow3) {
        const uint32_t er = g_hq_eq5[fr], eg = g_hq_eq6[fg], eb = g_hq_eq5[fb];
        const uint32_t hr = g_hq_half5[fr], hg = g_hq_half6[fg], hb = g_hq_half5[fb];
        const uint32_t err4 = (er >> 16) + (eg >> 16) + (eb >> 16), err3 = (hr >> 16) + (hg >> 16) + (hb >> 16);
        if (err3 < err4) {
            int max16 = (int)(((hr & 0xFF) << 11) | ((hg & 0xFF) << 5) | (hb & 0xFF));
            int min16 = (int)((((hr >> 8) & 0xFF) << 11) | (((hg >> 8) & 0xFF) << 5) | ((hb >> 8) & 0xFF));
            if (max16 > min16) { const int t = max16; max16 = min16; min16 = t; }
            return make_uint2((uint32_t)max16 | ((uint32_t)min16 << 16), 0xAAAAAAAAu);
        }
    }
    return bc1_solid(fr, fg, fb);
}

// bc1_encode3 (:1302-1338)
__device__ inline uint2 hq_pack3(const Bc1Ep& e, uint32_t sels)
{
    uint32_t lc16 = ((uint32_t)e.lr << 11) | ((uint32_t)e.lg << 5) | (uint32_t)e.lb;
    uint32_t hc16 = ((uint32_t)e.hr << 11) | ((uint32_t)e.hg << 5) | (uint32_t)e.hb;
    uint32_t packed = sels;
    if (lc16 > hc16) {
        const uint32_t t = lc16; lc16 = hc16; hc16 = t;
        packed = 0;
        for (int i = 0; i < 16; i++) {
            const uint32_t s = hq_sel(sels, i);
            packed |= (s < 2 ? s ^ 1u : s) << (2 * i); // s_sel_trans_inv = {1, 0, 2, 3}
        }
    }
    return make_uint2(lc16 | (hc16 << 16), packed);
}

// ---- rgbcx::encode_bc1(pDst, pPixels, flags, total_orderings_to_try, total_orderings_to_try3) :2263-2604
__device__ inline uint2 bc1_hq_encode_block(const uint32_t px[16], uint32_t flags, uint32_t to4, uint32_t to3)
{
    const uint32_t rgb0 = px[0] & 0xFFFFFFu;
    bool solid = true;
    for (int i = 1; i < 16; i++) solid &= ((px[i] & 0xFFFFFFu) == rgb0);
    if (solid) return hq_solid(rgb0 & 0xFF, (rgb0 >> 8) & 0xFF, rgb0 >> 16, (flags & (RF_USE3 | RF_BLACK3)) != 0);

    HqPx p;
    HqStats S;
    bool any_black = false;
    S.gray = true;
    S.min_r = S.min_g = S.min_b = 255; S.max_r = S.max_g = S.max_b = 0; S.total_r = S.total_g = S.total_b = 0;
    for (int i = 0; i < 16; i++) {
        const int r = px[i] & 0xFF, g = (px[i] >> 8) & 0xFF, b = (px[i] >> 16) & 0xFF;
        p.r[i] = r; p.g[i] = g; p.b[i] = b;
        S.gray &= (r == g) && (r == b);
        any_black |= (r | g | b) < 4;
        S.max_r = max(S.max_r, r); S.max_g = max(S.max_g, g); S.max_b = max(S.max_b, b);
        S.min_r = min(S.min_r, r); S.min_g = min(S.min_g, g); S.min_b = min(S.min_b, b);
        S.total_r += r; S.total_g += g; S.total_b += b;
    }
    S.avg_r = (S.total_r + 8) >> 4; S.avg_g = (S.total_g + 8) >> 4; S.avg_b = (S.total_b + 8) >> 4;

    const bool needs_block_error = (flags & (RF_LIKELY | RF_USE3 | RF_FULL | RF_ROUNDS_MASK)) != 0 || (any_black && (flags & RF_BLACK3) != 0);
    // level 5 asked for transparent black on a block without near-black texels: the reference takes its other branch and
    // nothing after it applies (cur_err stays UINT32_MAX, no 3-colour trial, no search rounds)
    if (!needs_block_error) return bc1_encode_block(px, 5);

    HqResult R;
    R.three = false;
    R.sels = 0;
    R.e.lr = R.e.lg = R.e.lb = R.e.hr = R.e.hg = R.e.hb = 0;
    Bc1Ep orig = R.e;
    uint32_t cur_err = 0xFFFFFFFFu;

    const int total_rounds = (flags & RF_TRYALL) ? 2 : 1;
    for (int round = 0; round < total_rounds; round++) {
        uint32_t mflags = flags;
        if (round == 1) mflags = (flags & ~(uint32_t)(RF_2DLS | RF_BBOX)) | RF_BBOX;
        Bc1Ep re;
        hq_pick_initial(px, p, mflags, S, re);
        const Bc1Ep orig_re = re;
        uint32_t rsels, rerr = hq_find4(flags, p, re, rsels, 0xFFFFFFFFu);
        const int passes = (flags & RF_TWO_LS) ? 2 : 1;
        for (int pass = 0; pass < passes; pass++) {
            float xl[3], xh[3];
            Bc1Ep t;
            if (!hq_ls4(p, rsels, S.total_r, S.total_g, S.total_b, xl, xh)) hq_single4(S.avg_r, S.avg_g, S.avg_b, t);
            else hq_round_swapped(xl, xh, t);
            if (hq_same(re, t)) break;
            uint32_t tsels;
            const uint32_t terr = hq_find4(flags, p, t, tsels, rerr);
            if (terr < rerr) { re = t; rerr = terr; rsels = tsels; }
            else break;
        }
        if (rerr <= cur_err) {
            cur_err = rerr;
            R.e = re;
            orig = orig_re;
            R.sels = rsels;
        }
    }

    if (cur_err && (flags & RF_LIKELY)) {
        const int iters = (flags & RF_ITERATIVE) ? 2 : 1;
        for (int it = 0; it < iters; it++) {
            const uint32_t orig_err = cur_err;
            const uint32_t base = hq_order_index4(R.sels);
            HqSums sums;
            hq_sorted_sums(p, R.e, S.total_r, S.total_g, S.total_b, sums);
            const uint32_t q_total = min(max(to4, 1u), (uint32_t)HQ_MAX4);
            for (uint32_t q = 0; q < q_total; q++) {
                const uint32_t s = g_hq_best4[base * HQ_MAX4 + q];
                Bc1Ep t;
                if (s == 15 || s == 700 || s == 753 || s == 515) hq_single4(S.avg_r, S.avg_g, S.avg_b, t);
                else {
                    float xl[3], xh[3];
                    hq_ls4_sums(s, sums, S.total_r, S.total_g, S.total_b, xl, xh);
                    hq_round_swapped(xl, xh, t);
                }
                uint32_t tsels;
                const uint32_t terr = hq_find4(flags, p, t, tsels, cur_err);
                if (terr < cur_err) { cur_err = terr; R.e = t; R.sels = tsels; }
            }
            if (!cur_err || cur_err == orig_err) break;
        }
    }

    if ((flags & (RF_USE3 | RF_BLACK3)) && cur_err) {
        if (flags & RF_USE3) hq_try3(p, flags, cur_err, S, orig, to3, R);
        if (any_black && (flags & RF_BLACK3)) hq_try3_black(p, flags, cur_err, R);
    }
    if ((flags & RF_ROUNDS_MASK) && cur_err) hq_endpoint_search(p, any_black, flags, R, cur_err);

    return R.three ? hq_pack3(R.e, R.sels) : bc1_pack4(R.e, R.sels);
}

} // namespace hkbcn
