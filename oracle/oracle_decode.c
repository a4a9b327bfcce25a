/* TEST INFRASTRUCTURE ONLY.
 * oracle_decode.c: serial plain-C restatement of the decoders Hork wires behind Decode_BC*
 * (ImageEncoders.cpp:77-136): bcdec for BC1-BC6H, bc7decomp for BC7. */
#include "oracle.h"
#include <string.h>

static uint16_t rd16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
static uint64_t rd64(const uint8_t* p)
{
    uint64_t v = 0;
    for (int i = 0; i < 8; i++) v |= (uint64_t)p[i] << (8 * i);
    return v;
}

/* bcdec.h:93-151 bcdec__color_block */
static void color_block(const uint8_t* blk, uint8_t* dst, size_t pitch, int opaque_only)
{
    uint32_t c0 = rd16(blk), c1 = rd16(blk + 2), ref[4][4];
    uint32_t e[2][3] = {{(((c0 >> 11) & 31) * 527 + 23) >> 6, (((c0 >> 5) & 63) * 259 + 33) >> 6, ((c0 & 31) * 527 + 23) >> 6},
                        {(((c1 >> 11) & 31) * 527 + 23) >> 6, (((c1 >> 5) & 63) * 259 + 33) >> 6, ((c1 & 31) * 527 + 23) >> 6}};
    for (int k = 0; k < 3; k++) {
        ref[0][k] = e[0][k];
        ref[1][k] = e[1][k];
        if (c0 > c1 || opaque_only) {
            ref[2][k] = (2 * e[0][k] + e[1][k] + 1) / 3;
            ref[3][k] = (e[0][k] + 2 * e[1][k] + 1) / 3;
        } else {
            ref[2][k] = (e[0][k] + e[1][k] + 1) >> 1;
            ref[3][k] = 0;
        }
    }
    ref[0][3] = ref[1][3] = ref[2][3] = 255;
    ref[3][3] = (c0 > c1 || opaque_only) ? 255 : 0;
    uint32_t idx = (uint32_t)blk[4] | ((uint32_t)blk[5] << 8) | ((uint32_t)blk[6] << 16) | ((uint32_t)blk[7] << 24);
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++, idx >>= 2)
            for (int k = 0; k < 4; k++) dst[y * pitch + x * 4 + k] = (uint8_t)ref[idx & 3][k];
}

/* bcdec.h:170-210 bcdec__smooth_alpha_block */
static void smooth_alpha(const uint8_t* blk, uint8_t* dst, size_t pitch, int pixel_size)
{
    uint32_t a[8];
    a[0] = blk[0];
    a[1] = blk[1];
    if (a[0] > a[1]) {
        for (uint32_t i = 1; i < 7; i++) a[1 + i] = ((7 - i) * a[0] + i * a[1] + 1) / 7;
    } else {
        for (uint32_t i = 1; i < 5; i++) a[1 + i] = ((5 - i) * a[0] + i * a[1] + 1) / 5;
        a[6] = 0;
        a[7] = 255;
    }
    uint64_t idx = rd64(blk) >> 16;
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++, idx >>= 3) dst[y * pitch + x * pixel_size] = (uint8_t)a[idx & 7];
}

void oracle_decode_bc1(const uint8_t* blk, uint8_t* dst, size_t pitch) { color_block(blk, dst, pitch, 0); }
void oracle_decode_bc2(const uint8_t* blk, uint8_t* dst, size_t pitch)
{
    color_block(blk + 8, dst, pitch, 1);
    for (int y = 0; y < 4; y++) { /* bcdec.h:153-168 sharp alpha */
        uint32_t row = rd16(blk + 2 * y);
        for (int x = 0; x < 4; x++) dst[y * pitch + x * 4 + 3] = (uint8_t)(((row >> (4 * x)) & 15) * 17);
    }
}
void oracle_decode_bc3(const uint8_t* blk, uint8_t* dst, size_t pitch)
{
    color_block(blk + 8, dst, pitch, 1);
    smooth_alpha(blk, dst + 3, pitch, 4);
}
void oracle_decode_bc4(const uint8_t* blk, uint8_t* dst, size_t pitch) { smooth_alpha(blk, dst, pitch, 1); }
void oracle_decode_bc5(const uint8_t* blk, uint8_t* dst, size_t pitch)
{
    smooth_alpha(blk, dst, pitch, 2);
    smooth_alpha(blk + 8, dst + 1, pitch, 2);
}

/* ------------------------------------------------------------------ shared format tables
 * D3D11 BC7/BC6H format constants: two-subset shapes (BC6H uses the first 32) as 16-bit masks,
 * bit i = subset of pixel i; three-subset shapes 2 bits per pixel; fix-up (anchor) pixels. */
static const uint16_t shape2[64] = {
    0xCCCC, 0x8888, 0xEEEE, 0xECC8, 0xC880, 0xFEEC, 0xFEC8, 0xEC80, 0xC800, 0xFFEC, 0xFE80, 0xE800, 0xFFE8, 0xFF00, 0xFFF0, 0xF000,
    0xF710, 0x008E, 0x7100, 0x08CE, 0x008C, 0x7310, 0x3100, 0x8CCE, 0x088C, 0x3110, 0x6666, 0x366C, 0x17E8, 0x0FF0, 0x718E, 0x399C,
    0xAAAA, 0xF0F0, 0x5A5A, 0x33CC, 0x3C3C, 0x55AA, 0x9696, 0xA55A, 0x73CE, 0x13C8, 0x324C, 0x3BDC, 0x6996, 0xC33C, 0x9966, 0x0660,
    0x0272, 0x04E4, 0x4E40, 0x2720, 0xC936, 0x936C, 0x39C6, 0x639C, 0x9336, 0x9CC6, 0x817E, 0xE718, 0xCCF0, 0x0FCC, 0x7744, 0xEE22};
static const uint32_t shape3[64] = {
    0xAA685050, 0x6A5A5040, 0x5A5A4200, 0x5450A0A8, 0xA5A50000, 0xA0A05050, 0x5555A0A0, 0x5A5A5050, 0xAA550000, 0xAA555500, 0xAAAA5500,
    0x90909090, 0x94949494, 0xA4A4A4A4, 0xA9A59450, 0x2A0A4250, 0xA5945040, 0x0A425054, 0xA5A5A500, 0x55A0A0A0, 0xA8A85454, 0x6A6A4040,
    0xA4A45000, 0x1A1A0500, 0x0050A4A4, 0xAAA59090, 0x14696914, 0x69691400, 0xA08585A0, 0xAA821414, 0x50A4A450, 0x6A5A0200, 0xA9A58000,
    0x5090A0A8, 0xA8A09050, 0x24242424, 0x00AA5500, 0x24924924, 0x24499224, 0x50A50A50, 0x500AA550, 0xAAAA4444, 0x66660000, 0xA5A0A5A0,
    0x50A050A0, 0x69286928, 0x44AAAA44, 0x66666600, 0xAA444444, 0x54A854A8, 0x95809580, 0x96969600, 0xA85454A8, 0x80959580, 0xAA141414,
    0x96960000, 0xAAAA1414, 0xA05050A0, 0xA0A5A5A0, 0x96000000, 0x40804080, 0xA9A8A9A8, 0xAAAAAA44, 0x2A4A5254};
static const uint8_t anchor2[64] = {15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 2, 8, 2, 2, 8, 8, 15, 2, 8, 2, 2, 8, 8, 2, 2,
                                    15, 15, 6, 8, 2, 8, 15, 15, 2, 8, 2, 2, 2, 15, 15, 6, 6, 2, 6, 8, 15, 15, 2, 2, 15, 15, 15, 15, 15, 2, 2, 15};
static const uint8_t anchor3a[64] = {3, 3, 15, 15, 8, 3, 15, 15, 8, 8, 6, 6, 6, 5, 3, 3, 3, 3, 8, 15, 3, 3, 6, 10, 5, 8, 8, 6, 8, 5, 15, 15,
                                     8, 15, 3, 5, 6, 10, 8, 15, 15, 3, 15, 5, 15, 15, 15, 15, 3, 15, 5, 5, 5, 8, 5, 10, 5, 10, 8, 13, 15, 12, 3, 3};
static const uint8_t anchor3b[64] = {15, 8, 8, 3, 15, 15, 3, 8, 15, 15, 15, 15, 15, 15, 15, 8, 15, 8, 15, 3, 15, 8, 15, 8, 3, 15, 6, 10, 15, 15, 10, 8,
                                     15, 3, 15, 10, 10, 8, 9, 10, 6, 15, 8, 15, 3, 6, 6, 8, 15, 3, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 3, 15, 15, 8};
static const uint8_t w2[4] = {0, 21, 43, 64}, w3[8] = {0, 9, 18, 27, 37, 46, 55, 64},
                     w4[16] = {0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64};

typedef struct { uint64_t lo, hi; } bits_t;
static uint32_t take(bits_t* b, int n) /* LSB-first bit reader (bcdec.h:217-229) */
{
    if (n == 0) return 0;
    uint32_t v = (uint32_t)(b->lo & ((1ull << n) - 1));
    b->lo = (b->lo >> n) | (b->hi << (64 - n));
    b->hi >>= n;
    return v;
}

/* ------------------------------------------------------------------ BC7
 * bc7decomp.cpp:168-634 (unpack_bc7 and the four per-mode-family routines) */
void oracle_decode_bc7(const uint8_t* blk, uint8_t* dst, size_t pitch)
{
    static const uint8_t NS[8] = {3, 2, 3, 2, 1, 1, 1, 2}, PB[8] = {4, 6, 6, 6, 0, 0, 0, 6}, CB[8] = {4, 6, 5, 7, 5, 7, 7, 5},
                         AB[8] = {0, 0, 0, 0, 6, 8, 7, 5}, IB[8] = {3, 3, 2, 2, 2, 2, 4, 2}, IB2[8] = {0, 0, 0, 0, 3, 2, 0, 0},
                         EPB[8] = {1, 0, 0, 1, 0, 0, 1, 1}, SPB[8] = {0, 1, 0, 0, 0, 0, 0, 0};
    uint8_t out[16][4];
    int mode = 0;
    while (mode < 8 && !(blk[0] & (1 << mode))) mode++;
    if (mode == 8) { /* reserved: transparent black (bc7decomp.cpp:628-630) */
        for (int y = 0; y < 4; y++) memset(dst + y * pitch, 0, 16);
        return;
    }
    bits_t b = {rd64(blk), rd64(blk + 8)};
    take(&b, mode + 1);
    uint32_t part = take(&b, PB[mode]);
    uint32_t rot = (mode == 4 || mode == 5) ? take(&b, 2) : 0;
    uint32_t isel = (mode == 4) ? take(&b, 1) : 0;
    int ne = NS[mode] * 2;
    uint32_t ep[6][4];
    for (int c = 0; c < 4; c++)
        for (int e = 0; e < ne; e++) {
            int nb = c < 3 ? CB[mode] : AB[mode];
            ep[e][c] = nb ? take(&b, nb) : 255;
        }
    uint32_t pb[6] = {0, 0, 0, 0, 0, 0};
    if (EPB[mode]) for (int e = 0; e < ne; e++) pb[e] = take(&b, 1);
    if (SPB[mode]) { pb[0] = pb[1] = take(&b, 1); pb[2] = pb[3] = take(&b, 1); }
    for (int e = 0; e < ne; e++)
        for (int c = 0; c < 4; c++) {
            int nb = c < 3 ? CB[mode] : AB[mode];
            if (!nb) continue;
            uint32_t v = ep[e][c];
            int tb = nb;
            if (EPB[mode] || SPB[mode]) { v = (v << 1) | pb[e]; tb++; }
            v <<= (8 - tb);
            v |= v >> tb;
            ep[e][c] = v;
        }
    int subset[16], a1 = -1, a2 = -1;
    for (int i = 0; i < 16; i++) subset[i] = NS[mode] == 1 ? 0 : NS[mode] == 2 ? (shape2[part] >> i) & 1 : (shape3[part] >> (2 * i)) & 3;
    if (NS[mode] == 2) a1 = anchor2[part];
    if (NS[mode] == 3) { a1 = anchor3a[part]; a2 = anchor3b[part]; }
    uint32_t i1[16], i2[16];
    for (int i = 0; i < 16; i++) i1[i] = take(&b, IB[mode] - (i == 0 || i == a1 || i == a2));
    if (IB2[mode]) for (int i = 0; i < 16; i++) i2[i] = take(&b, IB2[mode] - (i == 0));
    for (int i = 0; i < 16; i++) {
        const uint32_t *l = ep[subset[i] * 2], *h = ep[subset[i] * 2 + 1];
        int cb = IB[mode], ab = IB[mode];
        uint32_t ci = i1[i], ai = i1[i];
        if (IB2[mode]) {
            if (isel) { cb = IB2[mode]; ci = i2[i]; }
            else { ab = IB2[mode]; ai = i2[i]; }
        }
        uint32_t cw = cb == 2 ? w2[ci] : cb == 3 ? w3[ci] : w4[ci];
        uint32_t aw = ab == 2 ? w2[ai] : ab == 3 ? w3[ai] : w4[ai];
        for (int c = 0; c < 3; c++) out[i][c] = (uint8_t)((l[c] * (64 - cw) + h[c] * cw + 32) >> 6);
        out[i][3] = (uint8_t)((l[3] * (64 - aw) + h[3] * aw + 32) >> 6);
        if (rot) { uint8_t t = out[i][3]; out[i][3] = out[i][rot - 1]; out[i][rot - 1] = t; }
    }
    for (int y = 0; y < 4; y++) memcpy(dst + y * pitch, out[y * 4], 16);
}

/* ------------------------------------------------------------------ BC6H
 * bcdec.h:375-883 bcdec_bc6h_half, :885-902 bcdec_bc6h_float.
 * Header layouts per mode written as (field, lsb, count, reversed) runs in stream order;
 * fields: 0..3 = rw,rx,ry,rz; 4..7 = g*; 8..11 = b*; 12 = partition d. */
typedef struct { uint8_t field, lsb, cnt, rev; } run_t;
#define R_(f, l, c) {f, l, c, 0}
#define RV(f, l, c) {f, l, c, 1}
enum { RW, RX, RY, RZ, GW, GX, GY, GZ, BW, BX, BY, BZ, DD };
static const run_t lay0[] = {R_(GY,4,1),R_(BY,4,1),R_(BZ,4,1),R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,5),R_(GZ,4,1),R_(GY,0,4),R_(GX,0,5),R_(BZ,0,1),R_(GZ,0,4),R_(BX,0,5),R_(BZ,1,1),R_(BY,0,4),R_(RY,0,5),R_(BZ,2,1),R_(RZ,0,5),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay1[] = {R_(GY,5,1),R_(GZ,4,1),R_(GZ,5,1),R_(RW,0,7),R_(BZ,0,1),R_(BZ,1,1),R_(BY,4,1),R_(GW,0,7),R_(BY,5,1),R_(BZ,2,1),R_(GY,4,1),R_(BW,0,7),R_(BZ,3,1),R_(BZ,5,1),R_(BZ,4,1),R_(RX,0,6),R_(GY,0,4),R_(GX,0,6),R_(GZ,0,4),R_(BX,0,6),R_(BY,0,4),R_(RY,0,6),R_(RZ,0,6),R_(DD,0,5),{0,0,0,0}};
static const run_t lay2[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,5),R_(RW,10,1),R_(GY,0,4),R_(GX,0,4),R_(GW,10,1),R_(BZ,0,1),R_(GZ,0,4),R_(BX,0,4),R_(BW,10,1),R_(BZ,1,1),R_(BY,0,4),R_(RY,0,5),R_(BZ,2,1),R_(RZ,0,5),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay3[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,4),R_(RW,10,1),R_(GZ,4,1),R_(GY,0,4),R_(GX,0,5),R_(GW,10,1),R_(GZ,0,4),R_(BX,0,4),R_(BW,10,1),R_(BZ,1,1),R_(BY,0,4),R_(RY,0,4),R_(BZ,0,1),R_(BZ,2,1),R_(RZ,0,4),R_(GY,4,1),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay4[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,4),R_(RW,10,1),R_(BY,4,1),R_(GY,0,4),R_(GX,0,4),R_(GW,10,1),R_(BZ,0,1),R_(GZ,0,4),R_(BX,0,5),R_(BW,10,1),R_(BY,0,4),R_(RY,0,4),R_(BZ,1,1),R_(BZ,2,1),R_(RZ,0,4),R_(BZ,4,1),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay5[] = {R_(RW,0,9),R_(BY,4,1),R_(GW,0,9),R_(GY,4,1),R_(BW,0,9),R_(BZ,4,1),R_(RX,0,5),R_(GZ,4,1),R_(GY,0,4),R_(GX,0,5),R_(BZ,0,1),R_(GZ,0,4),R_(BX,0,5),R_(BZ,1,1),R_(BY,0,4),R_(RY,0,5),R_(BZ,2,1),R_(RZ,0,5),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay6[] = {R_(RW,0,8),R_(GZ,4,1),R_(BY,4,1),R_(GW,0,8),R_(BZ,2,1),R_(GY,4,1),R_(BW,0,8),R_(BZ,3,1),R_(BZ,4,1),R_(RX,0,6),R_(GY,0,4),R_(GX,0,5),R_(BZ,0,1),R_(GZ,0,4),R_(BX,0,5),R_(BZ,1,1),R_(BY,0,4),R_(RY,0,6),R_(RZ,0,6),R_(DD,0,5),{0,0,0,0}};
static const run_t lay7[] = {R_(RW,0,8),R_(BZ,0,1),R_(BY,4,1),R_(GW,0,8),R_(GY,5,1),R_(GY,4,1),R_(BW,0,8),R_(GZ,5,1),R_(BZ,4,1),R_(RX,0,5),R_(GZ,4,1),R_(GY,0,4),R_(GX,0,6),R_(GZ,0,4),R_(BX,0,5),R_(BZ,1,1),R_(BY,0,4),R_(RY,0,5),R_(BZ,2,1),R_(RZ,0,5),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay8[] = {R_(RW,0,8),R_(BZ,1,1),R_(BY,4,1),R_(GW,0,8),R_(BY,5,1),R_(GY,4,1),R_(BW,0,8),R_(BZ,5,1),R_(BZ,4,1),R_(RX,0,5),R_(GZ,4,1),R_(GY,0,4),R_(GX,0,5),R_(BZ,0,1),R_(GZ,0,4),R_(BX,0,6),R_(BY,0,4),R_(RY,0,5),R_(BZ,2,1),R_(RZ,0,5),R_(BZ,3,1),R_(DD,0,5),{0,0,0,0}};
static const run_t lay9[] = {R_(RW,0,6),R_(GZ,4,1),R_(BZ,0,1),R_(BZ,1,1),R_(BY,4,1),R_(GW,0,6),R_(GY,5,1),R_(BY,5,1),R_(BZ,2,1),R_(GY,4,1),R_(BW,0,6),R_(GZ,5,1),R_(BZ,3,1),R_(BZ,5,1),R_(BZ,4,1),R_(RX,0,6),R_(GY,0,4),R_(GX,0,6),R_(GZ,0,4),R_(BX,0,6),R_(BY,0,4),R_(RY,0,6),R_(RZ,0,6),R_(DD,0,5),{0,0,0,0}};
static const run_t lay10[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,10),R_(GX,0,10),R_(BX,0,10),{0,0,0,0}};
static const run_t lay11[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,9),R_(RW,10,1),R_(GX,0,9),R_(GW,10,1),R_(BX,0,9),R_(BW,10,1),{0,0,0,0}};
static const run_t lay12[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,8),RV(RW,10,2),R_(GX,0,8),RV(GW,10,2),R_(BX,0,8),RV(BW,10,2),{0,0,0,0}};
static const run_t lay13[] = {R_(RW,0,10),R_(GW,0,10),R_(BW,0,10),R_(RX,0,4),RV(RW,10,6),R_(GX,0,4),RV(GW,10,6),R_(BX,0,4),RV(BW,10,6),{0,0,0,0}};
static const run_t* const layouts[14] = {lay0, lay1, lay2, lay3, lay4, lay5, lay6, lay7, lay8, lay9, lay10, lay11, lay12, lay13};
/* actual_bits_count, bcdec.h:376-381 */
static const uint8_t wbits[14] = {10, 7, 11, 11, 11, 9, 8, 8, 8, 6, 10, 11, 12, 16};
static const uint8_t dbits[3][14] = {{5, 6, 5, 4, 4, 5, 6, 5, 5, 6, 10, 9, 8, 4}, {5, 6, 4, 5, 4, 5, 5, 6, 5, 6, 10, 9, 8, 4}, {5, 6, 4, 4, 5, 5, 5, 5, 6, 6, 10, 9, 8, 4}};

static int sext(int v, int bits) { return (int)((uint32_t)v << (32 - bits)) >> (32 - bits); }

static int unquantize(int val, int bits, int sgn) /* bcdec.h:289-329 */
{
    if (!sgn) {
        if (bits >= 15) return val;
        if (!val) return 0;
        if (val == (1 << bits) - 1) return 0xFFFF;
        return ((val << 16) + 0x8000) >> bits;
    }
    if (bits >= 16) return val;
    int s = 0, unq;
    if (val < 0) { s = 1; val = -val; }
    if (val == 0) unq = 0;
    else if (val >= (1 << (bits - 1)) - 1) unq = 0x7FFF;
    else unq = ((val << 15) + 0x4000) >> (bits - 1);
    return s ? -unq : unq;
}
static uint16_t finish(int val, int sgn) /* bcdec.h:331-346 */
{
    if (!sgn) return (uint16_t)((val * 31) >> 6);
    val = val < 0 ? -(((-val) * 31) >> 5) : (val * 31) >> 5;
    int s = 0;
    if (val < 0) { s = 0x8000; val = -val; }
    return (uint16_t)(s | val);
}
static float half_to_float(uint16_t h) /* bcdec.h:348-373 */
{
    union { uint32_t u; float f; } o, magic = {113u << 23};
    const uint32_t shifted_exp = 0x7c00u << 13;
    o.u = (uint32_t)(h & 0x7fff) << 13;
    uint32_t e = shifted_exp & o.u;
    o.u += (127 - 15) << 23;
    if (e == shifted_exp) o.u += (128 - 16) << 23;
    else if (e == 0) { o.u += 1 << 23; o.f -= magic.f; }
    o.u |= (uint32_t)(h & 0x8000) << 16;
    return o.f;
}

void oracle_decode_bc6h(const uint8_t* blk, uint8_t* dst, size_t pitch, int sgn, int f32out)
{
    uint16_t px[16][3];
    bits_t b = {rd64(blk), rd64(blk + 8)};
    uint32_t code = take(&b, 2);
    if (code > 1) code |= take(&b, 3) << 2;
    static const int8_t code2mode[32] = {0, 1, 2, 10, -1, -1, 3, 11, -1, -1, 4, 12, -1, -1, 5, 13, -1, -1, 6, -1, -1, -1, 7, -1, -1, -1, 8, -1, -1, -1, 9, -1};
    int mode = code2mode[code];
    if (mode < 0) { /* reserved modes: zeros (bcdec.h:819-833) */
        memset(px, 0, sizeof px);
    } else {
        int f[13] = {0};
        for (const run_t* r = layouts[mode]; r->cnt; r++) {
            uint32_t v = take(&b, r->cnt);
            if (r->rev) { uint32_t t = 0; for (int i = 0; i < r->cnt; i++) t |= ((v >> i) & 1) << (r->cnt - 1 - i); v = t; }
            f[r->field] |= (int)(v << r->lsb);
        }
        const int W = wbits[mode], two = mode < 10, ne = two ? 4 : 2, part = two ? f[DD] : 0;
        const int transformed = mode != 9 && mode != 10;
        if (sgn) for (int c = 0; c < 3; c++) f[c * 4] = sext(f[c * 4], W);
        if (transformed || sgn)
            for (int i = 1; i < ne; i++) for (int c = 0; c < 3; c++) f[c * 4 + i] = sext(f[c * 4 + i], dbits[c][mode]);
        if (transformed)
            for (int i = 1; i < ne; i++) for (int c = 0; c < 3; c++) {
                int v = (f[c * 4 + i] + f[c * 4]) & ((1 << W) - 1);
                f[c * 4 + i] = sgn ? sext(v, W) : v;
            }
        for (int i = 0; i < 16; i++) {
            int fix = (i == 0) || (two && i == anchor2[part]);
            int ib = (two ? 3 : 4) - fix;
            uint32_t idx = take(&b, ib);
            int w = two ? w3[idx] : w4[idx];
            int s = two ? (shape2[part] >> i) & 1 : 0;
            for (int c = 0; c < 3; c++) {
                int e0 = unquantize(f[c * 4 + s * 2], W, sgn), e1 = unquantize(f[c * 4 + s * 2 + 1], W, sgn);
                px[i][c] = finish((e0 * (64 - w) + e1 * w + 32) >> 6, sgn);
            }
        }
    }
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++)
            for (int c = 0; c < 3; c++) {
                if (f32out) { float v = half_to_float(px[y * 4 + x][c]); memcpy(dst + y * pitch + (x * 3 + c) * 4, &v, 4); }
                else memcpy(dst + y * pitch + (x * 3 + c) * 2, &px[y * 4 + x][c], 2);
            }
}

/* ------------------------------------------------------------------ surfaces
 * the gather / scatter loops of ImageEncoders.cpp:215-557 (encode) and the per-block
 * Decode_BC*(src, dst, RowStride) applied to every block (decode). */
int oracle_compress(int fmt, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h)
{
    static const int bpp_t[8] = {4, 4, 4, 1, 2, 16, 16, 4}, bb_t[8] = {8, 16, 16, 8, 16, 16, 16, 16};
    if (fmt < 0 || fmt > 4) return -1;
    const int bpp = bpp_t[fmt], bb = bb_t[fmt];
    uint8_t block[64];
    for (uint32_t by = 0; by < h / 4; by++)
        for (uint32_t bx = 0; bx < w / 4; bx++, dst += bb) {
            for (int r = 0; r < 4; r++) memcpy(block + r * 4 * bpp, src + ((size_t)(by * 4 + r) * w + bx * 4) * bpp, 4 * bpp);
            switch (fmt) {
            case 0: oracle_encode_bc1(block, dst); break;
            case 1: oracle_encode_bc2(block, dst); break;
            case 2: oracle_encode_bc3(block, dst); break;
            case 3: oracle_encode_bc4_hq(block, 1, dst); break;
            case 4: oracle_encode_bc5(block, dst); break;
            }
        }
    return 0;
}

int oracle_decompress(int fmt, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, size_t row_stride, int f32out)
{
    static const int bb_t[8] = {8, 16, 16, 8, 16, 16, 16, 16};
    if (fmt < 0 || fmt > 7) return -1;
    const int opp = fmt == 3 ? 1 : fmt == 4 ? 2 : (fmt == 5 || fmt == 6) ? (f32out ? 12 : 6) : 4;
    for (uint32_t by = 0; by < h / 4; by++)
        for (uint32_t bx = 0; bx < w / 4; bx++, src += bb_t[fmt]) {
            uint8_t* d = dst + (size_t)by * 4 * row_stride + (size_t)bx * 4 * opp;
            switch (fmt) {
            case 0: oracle_decode_bc1(src, d, row_stride); break;
            case 1: oracle_decode_bc2(src, d, row_stride); break;
            case 2: oracle_decode_bc3(src, d, row_stride); break;
            case 3: oracle_decode_bc4(src, d, row_stride); break;
            case 4: oracle_decode_bc5(src, d, row_stride); break;
            case 5: oracle_decode_bc6h(src, d, row_stride, 0, f32out); break;
            case 6: oracle_decode_bc6h(src, d, row_stride, 1, f32out); break;
            case 7: oracle_decode_bc7(src, d, row_stride); break;
            }
        }
    return 0;
}
