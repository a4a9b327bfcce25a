/* TEST INFRASTRUCTURE ONLY.
 * oracle_encode.c: serial plain-C restatement of the reference's deterministic encoders
 * (BC1 level 5, BC2, BC3 hq, BC4 hq / fast, BC5 hq).  Compile with -ffp-contract=off
 * (oracle/Makefile): the reference build never fuses multiply-add. */
#include "oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ BC4
 * rgbcx.h:391-423  bc4_block::get_block_values */
static void bc4_values(uint32_t l, uint32_t h, int v[8])
{
    v[0] = (int)l;
    v[1] = (int)h;
    if (l > h) {
        for (int i = 1; i < 7; i++) v[1 + i] = (int)((l * (7 - i) + h * i) / 7);
    } else {
        for (int i = 1; i < 5; i++) v[1 + i] = (int)((l * (5 - i) + h * i) / 5);
        v[6] = 0;
        v[7] = 255;
    }
}

/* rgbcx.cpp:2730-2884  encode_bc4_hq(dst, px, stride, search_rad = 3, modes = ALL) */
void oracle_encode_bc4_hq(const uint8_t* px, uint32_t stride, uint8_t* dst)
{
    int mn = 255, mx = 0;
    for (int i = 0; i < 16; i++) {
        int p = px[i * stride];
        if (p < mn) mn = p;
        if (p > mx) mx = p;
    }
    if (mn == mx) { /* :2744-2754, mode-6 flag set */
        dst[0] = dst[1] = (uint8_t)mn;
        memset(dst + 2, 0, 6);
        return;
    }
    uint32_t best = 0xFFFFFFFFu;
    for (int mode = 0; mode < 2; mode++)
        for (int lo = -3; lo <= 3; lo++)
            for (int hi = -3; hi <= 3; hi++) {
                int e0 = mx + hi, e1 = mn + lo;
                e0 = e0 < 0 ? 0 : e0 > 255 ? 255 : e0;
                e1 = e1 < 0 ? 0 : e1 > 255 ? 255 : e1;
                if (e0 == e1) continue;
                int is6 = e0 <= e1; /* is_alpha6_block() */
                if ((mode == 0 && is6) || (mode == 1 && !is6)) { int t = e0; e0 = e1; e1 = t; }
                int v[8];
                bc4_values((uint32_t)e0, (uint32_t)e1, v);
                uint32_t err = 0;
                uint8_t sel[16];
                int i;
                for (i = 0; i < 16; i++) { /* :2845-2866 */
                    uint32_t be = 0xFFFFFFFFu, bi = 0;
                    for (uint32_t j = 0; j < 8; j++) {
                        int d = v[j] - px[i * stride];
                        uint32_t e = (uint32_t)(d * d);
                        if (e < be) { be = e; bi = j; if (!e) break; }
                    }
                    err += be;
                    if (err >= best) break;
                    sel[i] = (uint8_t)bi;
                }
                if (err < best) {
                    best = err;
                    uint64_t s = 0;
                    for (i = 0; i < 16; i++) s |= (uint64_t)sel[i] << (i * 3);
                    dst[0] = (uint8_t)e0;
                    dst[1] = (uint8_t)e1;
                    for (i = 0; i < 6; i++) dst[2 + i] = (uint8_t)(s >> (8 * i));
                }
            }
}

/* rgbcx.cpp:2608-2728  encode_bc4 (non-hq) */
void oracle_encode_bc4_fast(const uint8_t* px, uint32_t stride, uint8_t* dst)
{
    int mn = 255, mx = 0;
    for (int i = 0; i < 16; i++) {
        int p = px[i * stride];
        if (p < mn) mn = p;
        if (p > mx) mx = p;
    }
    dst[0] = (uint8_t)mx;
    dst[1] = (uint8_t)mn;
    if (mx == mn) { memset(dst + 2, 0, 6); return; }
    const int delta = mx - mn, bias = 4 - mn * 14;
    static const uint32_t tran[8] = {1, 7, 6, 5, 4, 3, 2, 0};
    uint64_t s = 0;
    for (int i = 0; i < 16; i++) {
        int vv = px[i * stride] * 14 + bias, c = 0;
        for (int k = 13; k >= 1; k -= 2) c += vv >= delta * k;
        s |= (uint64_t)tran[c] << (i * 3);
    }
    for (int i = 0; i < 6; i++) dst[2 + i] = (uint8_t)(s >> (8 * i));
}

/* rgbcx.cpp:2921-2927 encode_bc5_hq(dst, px, 0, 1, 2) */
void oracle_encode_bc5(const uint8_t* rg, uint8_t* dst)
{
    oracle_encode_bc4_hq(rg, 2, dst);
    oracle_encode_bc4_hq(rg + 1, 2, dst + 8);
}

/* ------------------------------------------------------------------ BC1
 * a12 tables: rgbcx.cpp:485-527 (prepare_bc1_single_color_table, cBC1Ideal), built once. */
static uint8_t m5_hi[256], m5_lo[256], m6_hi[256], m6_lo[256];
static float mid5[32], mid6[64];
static int tables_ready;

static void match_table(uint8_t* thi, uint8_t* tlo, int size)
{
    for (int i = 0; i < 256; i++) {
        int lowest = 256;
        for (int lo = 0; lo < size; lo++)
            for (int hi = 0; hi < size; hi++) {
                int lo_e = size == 32 ? (lo << 3) | (lo >> 2) : (lo << 2) | (lo >> 4);
                int hi_e = size == 32 ? (hi << 3) | (hi >> 2) : (hi << 2) | (hi >> 4);
                int e = abs((hi_e * 2 + lo_e) / 3 - i) + (abs(hi_e - lo_e) * 3) / 100; /* interp_5_6_ideal + 3% */
                if (e < lowest || (e == lowest && lo == hi)) { thi[i] = (uint8_t)hi; tlo[i] = (uint8_t)lo; lowest = e; }
            }
    }
}
static void init_tables(void)
{
    if (tables_ready) return;
    match_table(m5_hi, m5_lo, 32);
    match_table(m6_hi, m6_lo, 64);
    /* rgbcx.cpp:693-694 g_midpoint5/6: 6-decimal literals of the level midpoints, last = 1e+37 */
    char buf[32];
    for (int i = 0; i < 31; i++) {
        int a = (i << 3) | (i >> 2), b = ((i + 1) << 3) | ((i + 1) >> 2);
        snprintf(buf, sizeof buf, "%.6f", (a + b) / 510.0);
        mid5[i] = strtof(buf, NULL);
    }
    for (int i = 0; i < 63; i++) {
        int a = (i << 2) | (i >> 4), b = ((i + 1) << 2) | ((i + 1) >> 4);
        snprintf(buf, sizeof buf, "%.6f", (a + b) / 510.0);
        mid6[i] = strtof(buf, NULL);
    }
    mid5[31] = mid6[63] = 1e+37f;
    tables_ready = 1;
}

static int to5(uint32_t v) { v = v * 31 + 128; return (uint8_t)((v + (v >> 8)) >> 8); }
static int to6(uint32_t v) { v = v * 63 + 128; return (uint8_t)((v + (v >> 8)) >> 8); }

static void put16(uint8_t* d, uint32_t v) { d[0] = (uint8_t)v; d[1] = (uint8_t)(v >> 8); }

/* rgbcx.cpp:629-691 encode_bc1_solid_block, allow_3color = false */
static void bc1_solid(uint8_t* dst, int r, int g, int b)
{
    uint32_t mask = 0xAA;
    int max16 = (m5_hi[r] << 11) | (m6_hi[g] << 5) | m5_hi[b];
    int min16 = (m5_lo[r] << 11) | (m6_lo[g] << 5) | m5_lo[b];
    if (min16 == max16) {
        mask = 0;
        if (min16 > 0) min16--;
        else { max16 = 1; min16 = 0; mask = 0x55; }
    }
    if (max16 < min16) { int t = max16; max16 = min16; min16 = t; mask ^= 0x55; }
    put16(dst, (uint32_t)max16);
    put16(dst + 2, (uint32_t)min16);
    memset(dst + 4, (int)mask, 4);
}

/* rgbcx.cpp:915-950 bc1_find_sels4_noerr, colours per :869-879 (cBC1Ideal) */
static void find_sels(const uint8_t* px, const int l[3], const int h[3], uint8_t sels[16])
{
    int c[4][3];
    c[0][0] = (l[0] << 3) | (l[0] >> 2); c[0][1] = (l[1] << 2) | (l[1] >> 4); c[0][2] = (l[2] << 3) | (l[2] >> 2);
    c[3][0] = (h[0] << 3) | (h[0] >> 2); c[3][1] = (h[1] << 2) | (h[1] >> 4); c[3][2] = (h[2] << 3) | (h[2] >> 2);
    for (int k = 0; k < 3; k++) {
        c[1][k] = (c[0][k] * 2 + c[3][k]) / 3;
        c[2][k] = (c[3][k] * 2 + c[0][k]) / 3;
    }
    int a[3] = {c[3][0] - c[0][0], c[3][1] - c[0][1], c[3][2] - c[0][2]};
    int dots[4];
    for (int i = 0; i < 4; i++) dots[i] = c[i][0] * a[0] + c[i][1] * a[1] + c[i][2] * a[2];
    int t0 = dots[0] + dots[1], t1 = dots[1] + dots[2], t2 = dots[2] + dots[3];
    a[0] *= 2; a[1] *= 2; a[2] *= 2;
    static const uint8_t s_sels[4] = {3, 2, 1, 0};
    for (int i = 0; i < 16; i++) {
        int d = px[i * 4] * a[0] + px[i * 4 + 1] * a[1] + px[i * 4 + 2] * a[2];
        sels[i] = s_sels[(d <= t0) + (d < t1) + (d < t2)];
    }
}

/* rgbcx.cpp:1194-1228 precise_round_565, one endpoint */
static void round565(const float x[3], int out[3])
{
    out[0] = (int)(x[0] * 31.0f);
    out[1] = (int)(x[1] * 63.0f);
    out[2] = (int)(x[2] * 31.0f);
    if ((uint32_t)out[0] > 31U) out[0] = (~out[0] >> 31) & 31;
    if ((uint32_t)out[2] > 31U) out[2] = (~out[2] >> 31) & 31;
    if ((uint32_t)out[1] > 63U) out[1] = (~out[1] >> 31) & 63;
    out[0] = (out[0] + (x[0] > mid5[out[0]])) & 31;
    out[1] = (out[1] + (x[1] > mid6[out[1]])) & 63;
    out[2] = (out[2] + (x[2] > mid5[out[2]])) & 31;
}

/* rgbcx.cpp:2263-2371 + :2600-2603: encode_bc1 with flags = TwoLeastSquaresPasses | UseFasterMSEEval
 * (level 5, no 3-colour), i.e. the !needs_block_error branch. */
void oracle_encode_bc1(const uint8_t* px, uint8_t* dst)
{
    init_tables();
    const int fr = px[0], fg = px[1], fb = px[2];
    int j;
    for (j = 15; j >= 1; --j)
        if (px[j * 4] != fr || px[j * 4 + 1] != fg || px[j * 4 + 2] != fb) break;
    if (j == 0) { bc1_solid(dst, fr, fg, fb); return; }

    int mn[3] = {fr, fg, fb}, mx[3] = {fr, fg, fb}, tot[3] = {fr, fg, fb};
    int gray = (fr == fg) && (fr == fb);
    for (int i = 1; i < 16; i++) {
        const uint8_t* p = px + i * 4;
        gray &= (p[0] == p[1]) && (p[0] == p[2]);
        for (int k = 0; k < 3; k++) {
            if (p[k] > mx[k]) mx[k] = p[k];
            if (p[k] < mn[k]) mn[k] = p[k];
            tot[k] += p[k];
        }
    }
    int avg[3] = {(tot[0] + 8) >> 4, (tot[1] + 8) >> 4, (tot[2] + 8) >> 4};
    int l[3], h[3];

    /* encode_bc1_pick_initial, rgbcx.cpp:1813-2175 */
    if (gray) {
        if (mx[0] - mn[0] < 2) {
            l[0] = l[2] = h[0] = h[2] = to5((uint32_t)fr);
            l[1] = h[1] = to6((uint32_t)fr);
        } else {
            l[0] = l[2] = to5((uint32_t)mn[0]); l[1] = to6((uint32_t)mn[0]);
            h[0] = h[2] = to5((uint32_t)mx[0]); h[1] = to6((uint32_t)mx[0]);
        }
    } else {
        int icov[6] = {0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 16; i++) {
            int r = px[i * 4] - avg[0], g = px[i * 4 + 1] - avg[1], b = px[i * 4 + 2] - avg[2];
            icov[0] += r * r; icov[1] += r * g; icov[2] += r * b; icov[3] += g * g; icov[4] += g * b; icov[5] += b * b;
        }
        int sr = 306, sg = 601, sb = 117;
        float xr = (float)(mx[0] - mn[0]), xg = (float)(mx[1] - mn[1]), xb = (float)(mx[2] - mn[2]);
        if (icov[2] < 0) xr = -xr;
        if (icov[4] < 0) xg = -xg;
        float cov[6];
        for (int i = 0; i < 6; i++) cov[i] = (float)icov[i] * (1.0f / 255.0f);
        for (int it = 0; it < 4; it++) {
            float r = xr * cov[0] + xg * cov[1] + xb * cov[2];
            float g = xr * cov[1] + xg * cov[3] + xb * cov[4];
            float b = xr * cov[2] + xg * cov[4] + xb * cov[5];
            xr = r; xg = g; xb = b;
        }
        float k = fabsf(xr);
        if (fabsf(xg) > k) k = fabsf(xg);
        if (fabsf(xb) > k) k = fabsf(xb);
        if (k >= 2) {
            float m = 2048.0f / k;
            sr = (int)(xr * m); sg = (int)(xg * m); sb = (int)(xb * m);
        }
        sr = (int)((uint32_t)sr << 4); sg = (int)((uint32_t)sg << 4); sb = (int)((uint32_t)sb << 4);
        int lowd = 0x7FFFFFFF, highd = (int)0x80000000;
        for (int i = 0; i < 16; i++) {
            int d = ((px[i * 4] * sr + px[i * 4 + 1] * sg + px[i * 4 + 2] * sb) & ~0xF) + i;
            if (d < lowd) lowd = d;
            if (d > highd) highd = d;
        }
        const uint8_t *pl = px + (lowd & 15) * 4, *ph = px + (highd & 15) * 4;
        l[0] = to5(pl[0]); l[1] = to6(pl[1]); l[2] = to5(pl[2]);
        h[0] = to5(ph[0]); h[1] = to6(ph[1]); h[2] = to5(ph[2]);
    }

    uint8_t sels[16];
    find_sels(px, l, h, sels);

    for (int pass = 0; pass < 2; pass++) { /* :2333-2370 */
        int tl[3], th[3];
        /* compute_least_squares_endpoints4_rgb, :727-773 */
        static const uint32_t wv[4] = {0x000009, 0x010204, 0x040201, 0x090000};
        uint32_t uq[3] = {0, 0, 0}, wacc = 0;
        for (int i = 0; i < 16; i++) {
            wacc += wv[sels[i]];
            for (int k = 0; k < 3; k++) uq[k] += sels[i] * (uint32_t)px[i * 4 + k];
        }
        float z00 = (float)((wacc >> 16) & 0xFF), z10 = (float)((wacc >> 8) & 0xFF), z11 = (float)(wacc & 0xFF), z01 = z10;
        float det = z00 * z11 - z01 * z10;
        if (fabs(det) < 1e-8f) {
            tl[0] = m5_hi[avg[0]]; tl[1] = m6_hi[avg[1]]; tl[2] = m5_hi[avg[2]];
            th[0] = m5_lo[avg[0]]; th[1] = m6_lo[avg[1]]; th[2] = m5_lo[avg[2]];
        } else {
            det = (3.0f / 255.0f) / det;
            float iz00 = z11 * det, iz01 = -z01 * det, iz10 = -z10 * det, iz11 = z00 * det;
            float xl[3], xh[3];
            for (int k = 0; k < 3; k++) {
                int q10 = tot[k] * 3 - (int)uq[k];
                xl[k] = iz00 * (float)uq[k] + iz01 * q10;
                xh[k] = iz10 * (float)uq[k] + iz11 * q10;
            }
            round565(xl, th); /* the call site hands (trial_h*, trial_l*) to (xl, xh): :2355 */
            round565(xh, tl);
        }
        if (!memcmp(l, tl, sizeof l) && !memcmp(h, th, sizeof h)) break;
        find_sels(px, tl, th, sels);
        memcpy(l, tl, sizeof l);
        memcpy(h, th, sizeof h);
    }

    /* bc1_encode4, :1245-1300 */
    uint32_t lc = (uint32_t)((l[0] << 11) | (l[1] << 5) | l[2]), hc = (uint32_t)((h[0] << 11) | (h[1] << 5) | h[2]);
    if (lc == hc) {
        uint8_t mask = 0;
        if (hc > 0) hc--;
        else { hc = 0; lc = 1; mask = 0x55; }
        put16(dst, lc); put16(dst + 2, hc);
        memset(dst + 4, mask, 4);
    } else {
        uint8_t inv = 0;
        if (lc < hc) { uint32_t t = lc; lc = hc; hc = t; inv = 0x55; }
        put16(dst, lc); put16(dst + 2, hc);
        static const uint8_t trans[4] = {0, 2, 3, 1};
        uint32_t packed = 0;
        for (int i = 0; i < 16; i++) packed |= (uint32_t)trans[sels[i]] << (i * 2);
        for (int i = 0; i < 4; i++) dst[4 + i] = (uint8_t)(packed >> (8 * i)) ^ inv;
    }
}

/* ImageEncoders.cpp:145-159 Encode_BC2 */
void oracle_encode_bc2(const uint8_t* px, uint8_t* dst)
{
    for (int i = 0; i < 8; i++) {
        int q1 = (int)(px[i * 8 + 3] * (15.0f / 255.0f) + 0.5f), q2 = (int)(px[i * 8 + 7] * (15.0f / 255.0f) + 0.5f);
        if (q1 > 15) q1 = 15;
        if (q2 > 15) q2 = 15;
        dst[i] = (uint8_t)(q1 | (q2 << 4));
    }
    oracle_encode_bc1(px, dst + 8);
}

/* rgbcx.cpp:2905-2911 encode_bc3_hq(level 5) */
void oracle_encode_bc3(const uint8_t* px, uint8_t* dst)
{
    oracle_encode_bc4_hq(px + 3, 4, dst);
    oracle_encode_bc1(px, dst + 8);
}
