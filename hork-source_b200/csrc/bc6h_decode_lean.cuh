// bc6h_decode_lean.cuh -- instruction-lean BC6H block decoder (sm_100a), all 14 modes, unsigned and signed.
//
// Bit-exact with bcdec_bc6h_half / bcdec_bc6h_float (ThirdParty/bcdec/bcdec.h:375-902), the decoders behind
// Hk::TextureBlockCompression::Decode_BC6h_f16 / _f32 (Hork/Image/ImageEncoders.cpp:102-110).
//
//   * header: one generated straight-line parser per mode (bc6h_parse_gen.cuh), every bit run a constant-position extract;
//     sign extension, delta transform and unquantisation with the mode's widths as compile-time constants;
//   * texels (unsigned): with A = a*256 + 128 and D = 4(b - a), x = D*w + A = 4(a(64-w) + bw + 32), v = x >> 8 is the
//     interpolated value and (v*124) >> 8 == (v*31) >> 6 is bcdec's finish_unquantize, so a texel channel costs
//     IMAD, SHF, IMAD and the final shift is absorbed by the PRMT that packs two halfs into an output word;
//   * index streams are re-expanded to fixed-width fields exactly as for BC7 modes 1 and 6 (same bit positions: 82 / 65).
// Plain integer code: tools/decode_host_check.cpp compiles it for the CPU against the reference.
#pragma once
#include <math.h>
#include <string.h>
#include "bc7_decode_lean.cuh"

namespace hkbcn {
namespace lean6 {

HKD uint32_t gb(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, int pos, int n) { return lean7::getbits(w0, w1, w2, w3, pos, n); }
// the n (2 or 6) low bits of v in reverse order (bcdec__bitstream_read_bits_r, bcdec.h:235-246)
HKD uint32_t rev_bits(uint32_t v, int n)
{
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 6; i++)
        if (i < n) r |= ((v >> i) & 1u) << (n - 1 - i);
    return r;
}

} // namespace lean6
} // namespace hkbcn
#include "bc6h_parse_gen.cuh"
namespace hkbcn {
namespace lean6 {

// actual_bits_count (bcdec.h:376-381): endpoint width and the three delta widths per mode
HKD constexpr int mode_W(int m) { return m == 0 ? 10 : m == 1 ? 7 : m <= 4 ? 11 : m == 5 ? 9 : m <= 8 ? 8 : m == 9 ? 6 : m == 10 ? 10 : m == 11 ? 11 : m == 12 ? 12 : 16; }
HKD constexpr int mode_D(int m, int c)
{
    return m == 0 ? 5 : m == 1 ? 6 : m == 2 ? (c == 0 ? 5 : 4) : m == 3 ? (c == 1 ? 5 : 4) : m == 4 ? (c == 2 ? 5 : 4) : m == 5 ? 5
         : m == 6 ? (c == 0 ? 6 : 5) : m == 7 ? (c == 1 ? 6 : 5) : m == 8 ? (c == 2 ? 6 : 5) : m == 9 ? 6 : m == 10 ? 10 : m == 11 ? 9 : m == 12 ? 8 : 4;
}

HKD int sext(int v, int bits) { return (int)((uint32_t)v << (32 - bits)) >> (32 - bits); }

// bcdec__unquantize (bcdec.h:289-329)
template <int W, bool SIGNED>
HKD int unquantize(int v)
{
    if (!SIGNED) {
        if (W >= 15) return v;
        int u = ((v << 16) + 0x8000) >> W;
        u = (v == (1 << W) - 1) ? 0xFFFF : u;
        return v == 0 ? 0 : u;
    } else {
        if (W >= 16) return v;
        const int a = v < 0 ? -v : v;
        int u = ((a << 15) + 0x4000) >> (W - 1);
        u = (a >= (1 << (W - 1)) - 1) ? 0x7FFF : u;
        u = a == 0 ? 0 : u;
        return v < 0 ? -u : u;
    }
}

// header -> 12 unquantised endpoint values e[c * 4 + i] (i = 0, 1: region 0's A, B; 2, 3: region 1's) and the partition
template <int MODE, bool SIGNED>
HKD void endpoints(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, int e[12], uint32_t& part)
{
    int f[13];
    if (MODE == 0) parse_mode0(w0, w1, w2, w3, f);
    else if (MODE == 1) parse_mode1(w0, w1, w2, w3, f);
    else if (MODE == 2) parse_mode2(w0, w1, w2, w3, f);
    else if (MODE == 3) parse_mode3(w0, w1, w2, w3, f);
    else if (MODE == 4) parse_mode4(w0, w1, w2, w3, f);
    else if (MODE == 5) parse_mode5(w0, w1, w2, w3, f);
    else if (MODE == 6) parse_mode6(w0, w1, w2, w3, f);
    else if (MODE == 7) parse_mode7(w0, w1, w2, w3, f);
    else if (MODE == 8) parse_mode8(w0, w1, w2, w3, f);
    else if (MODE == 9) parse_mode9(w0, w1, w2, w3, f);
    else if (MODE == 10) parse_mode10(w0, w1, w2, w3, f);
    else if (MODE == 11) parse_mode11(w0, w1, w2, w3, f);
    else if (MODE == 12) parse_mode12(w0, w1, w2, w3, f);
    else parse_mode13(w0, w1, w2, w3, f);
    constexpr int W = mode_W(MODE);
    constexpr int NE = MODE < 10 ? 4 : 2;
    constexpr bool TR = MODE != 9 && MODE != 10;
    part = (uint32_t)f[12];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        int base = f[c * 4];
        if (SIGNED) base = sext(base, W);
        e[c * 4] = unquantize<W, SIGNED>(base);
#pragma unroll
        for (int i = 1; i < 4; i++) {
            if (i >= NE) { e[c * 4 + i] = 0; continue; }
            int v = f[c * 4 + i];
            if (TR || SIGNED) v = sext(v, mode_D(MODE, c));
            if (TR) {
                v = (v + base) & ((1 << W) - 1);
                if (SIGNED) v = sext(v, W);
            }
            e[c * 4 + i] = unquantize<W, SIGNED>(v);
        }
    }
}

// ---- texels on the FMA pipe ---------------------------------------------------------------------------------------
// The integer version of the texel loop (IMAD, SHF, IMAD per channel, six LOP3 selects per two-region texel, SHF / PRMT
// weight lookups) ran the ALU pipe -- half rate on sm_100: LOP3, SHF, PRMT, SEL, ISETP -- at 73 % with the FMA pipe at 36 %
// (profiles/r1_bc6h_decode_lean_ncu.txt).  Everything below is exact in fp32 (all values are integers or half-integers
// below 2^24), so the interpolation moves to FFMA / FADD and the weights and region masks come from shared-memory tables:
//   unsigned: x = D*w + A with A = 256a + 128, D = 4(b - a) is 4(a(64-w) + bw + 32) < 2^24; v = x >> 8.
//             y = (x - 126) + 1.5*2^31 rounds to a multiple of 256 (the ulp there) and can never tie (x is a multiple of 4),
//             so the float's bit pattern is bits(1.5*2^31) + v; (bits * 124 - bits(C) * 124) = v * 124 has
//             bcdec's (v * 31) >> 6 in bytes 1..2, where the PRMT that packs two halfs picks it up.
//   signed:   x = D*w + A with A = 64a + 32, D = b - a; v = x >> 6 = round((x - 31.5) / 64), ulp 64 at 1.5*2^29.
//             r = RZ(|z| * 31/2048 + k) with z = 64 v (exact), k = 1.5*2^23 (+ 32768 when v <= -2, i.e. when bcdec sets the
//             sign bit: the magnitude (|v| * 31) >> 5 is zero for |v| <= 1): the float's low 16 bits are the half.
//             The "v <= -2" flag is sat(-z/64 - 1), one FFMA.SAT -- no compare, no select, nothing on the ALU pipe.
// Two channels of a texel share one packed instruction (fma / add .f32x2, SASS FFMA2 / FADD2: one issue slot, two results,
// each rounded as the scalar instruction rounds it -- and everything here is exact anyway); weights and region masks are
// stored as duplicated pairs so that one 8-byte shared-memory load yields the packed operand.
#ifdef HKBCN_HOST_EMU
static inline float hk_fma(float a, float b, float c) { return fmaf(a, b, c); }
static inline float hk_add(float a, float b) { return a + b; }
static inline uint32_t hk_fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float hk_bitsf(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline float hk_fma_sat(float a, float b, float c) { const float r = fmaf(a, b, c); return r < 0.0f ? 0.0f : r > 1.0f ? 1.0f : r; }
static inline float hk_fma_rz_abs(float a, float b, float c)
{   // |a| * b + c rounded toward zero; the exact sum fits a double
    const double t = fabs((double)a) * (double)b + (double)c;
    float f = (float)t;
    if (fabs((double)f) > fabs(t)) f = nextafterf(f, 0.0f);
    return f;
}
struct fp2 { float x, y; };
static inline fp2 hk_pack2(float a, float b) { return fp2{a, b}; }
static inline fp2 hk_fma2(fp2 a, fp2 b, fp2 c) { return fp2{fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)}; }
static inline fp2 hk_add2(fp2 a, fp2 b) { return fp2{a.x + b.x, a.y + b.y}; }
static inline float hk_lo(fp2 a) { return a.x; }
static inline float hk_hi(fp2 a) { return a.y; }
static inline fp2 hk_ld2(const void* p) { fp2 r; memcpy(&r, p, 8); return r; }
#else
HKD float hk_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
HKD float hk_add(float a, float b) { return __fadd_rn(a, b); }
HKD uint32_t hk_fbits(float f) { return __float_as_uint(f); }
HKD float hk_bitsf(uint32_t u) { return __uint_as_float(u); }
HKD float hk_fma_sat(float a, float b, float c) { float r; asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
HKD float hk_fma_rz_abs(float a, float b, float c) { return __fmaf_rz(fabsf(a), b, c); }
typedef unsigned long long fp2;
HKD fp2 hk_pack2(float a, float b) { fp2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
HKD fp2 hk_fma2(fp2 a, fp2 b, fp2 c) { fp2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
HKD fp2 hk_add2(fp2 a, fp2 b) { fp2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
HKD float hk_lo(fp2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return x; }
HKD float hk_hi(fp2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return y; }
HKD fp2 hk_ld2(const void* p) { return *(const fp2*)p; }
#endif

// aWeight3 / aWeight4 as floats and the region of every texel of the 32 two-region shapes as 0.0f / 1.0f, every value twice
// (one 8-byte load = the packed operand of two channels).  m is texel-major so that a warp's lookups (one texel, 32 shapes)
// fall into different banks.
struct FTab {
    float w3[8][2], w4[16][2];
    float m[16][32][2];
};
inline void fill_ftab(FTab* F, const uint16_t* shape2)
{
    static const int w3[8] = {0, 9, 18, 27, 37, 46, 55, 64}, w4[16] = {0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64};
    for (int i = 0; i < 8; i++) F->w3[i][0] = F->w3[i][1] = (float)w3[i];
    for (int i = 0; i < 16; i++) F->w4[i][0] = F->w4[i][1] = (float)w4[i];
    for (int s = 0; s < 32; s++)
        for (int i = 0; i < 16; i++) F->m[i][s][0] = F->m[i][s][1] = ((shape2[s] >> i) & 1) ? 1.0f : 0.0f;
}

// x = D * w + A (see above) -> the finished half: unsigned in bytes 1..2 of the result, signed in bytes 0..1
template <bool SIGNED>
HKD uint32_t finish_y(float y)
{
    if (!SIGNED) return hk_fbits(y) * 124u - 0x4F400000u * 124u;
    const float z = hk_add(y, -805306368.0f);                   // 64 v
    const float n = hk_fma_sat(z, -0.015625f, -1.0f);           // v <= -2 ? 1 : 0
    const float k = hk_fma(n, 32768.0f, 12582912.0f);           // 1.5 * 2^23 (+ the sign bit of the half)
    return hk_fbits(hk_fma_rz_abs(z, 0.01513671875f, k));       // |v| * 31 / 32, floored
}
template <bool SIGNED> HKD constexpr float finish_magic() { return SIGNED ? 805306368.0f : 3221225472.0f; } // 1.5 * 2^29 / 1.5 * 2^31

// Texel rows of one block.  sink.row(r, w) receives the six output words of texel row r (RGB16F x 4, bcdec's layout).
template <bool SIGNED, bool TWO, class Sink>
HKD void texels(const int e[12], uint32_t part, uint32_t w2, uint32_t w3, const lean7::Tables& T, const FTab& F, Sink& sink)
{
    float A[2][3], D[2][3];
#pragma unroll
    for (int s = 0; s < (TWO ? 2 : 1); s++)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const int a = e[c * 4 + 2 * s], b = e[c * 4 + 2 * s + 1];
            A[s][c] = SIGNED ? hk_add((float)(a * 64), 0.5f) : (float)(a * 256 + 2); // A - 31.5 / A - 126: see above
            D[s][c] = SIGNED ? (float)(b - a) : (float)(4 * (b - a));
        }
    uint32_t lo, hi;
    const float* mrow = &F.m[0][0][0];
    if (TWO) {
        // 46 index bits from bit 82; implicit zeros after texel 0's two bits and the anchor's (as BC7 mode 1)
        unsigned long long S = ((unsigned long long)(w3 >> 18) << 32) | hk_funnel_r(w2, w3, 18);
        S += S & ~3ull;
        S += S & (((unsigned long long)T.ins3[part][1] << 32) | T.ins3[part][0]);
        lo = (uint32_t)S;
        hi = (uint32_t)(S >> 32);
        mrow = &F.m[0][part][0];
#pragma unroll
        for (int c = 0; c < 3; c++) { // region 1 as an offset from region 0
            A[1][c] = hk_add(A[1][c], -A[0][c]);
            D[1][c] = hk_add(D[1][c], -D[0][c]);
        }
    } else {
        lo = w2 & ~1u; // texel 0: 3 bits at 65..67 -> a 4-bit field with its top bit clear, after >> 1 below
        hi = w3;
    }
    // channels R, G as one packed pair, B scalar
    const fp2 A0 = hk_pack2(A[0][0], A[0][1]), D0 = hk_pack2(D[0][0], D[0][1]);
    const fp2 A1 = TWO ? hk_pack2(A[1][0], A[1][1]) : A0, D1 = TWO ? hk_pack2(D[1][0], D[1][1]) : D0;
    const fp2 K = hk_pack2(finish_magic<SIGNED>(), finish_magic<SIGNED>());
#pragma unroll
    for (int r = 0; r < 4; r++) {
        uint32_t y[12];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = 4 * r + k;
            fp2 wd, a01, d01;
            float a2, d2;
            if (TWO) {
                // byte offset of the weight pair: 8 * ((S >> 3i) & 7)
                const int p = 3 * i;
                const uint32_t off = p == 0 ? (lo << 3) : p <= 29 ? (lo >> (p - 3)) : p >= 35 ? (hi >> (p - 35)) : hk_funnel_r(lo, hi, p - 3);
                wd = hk_ld2((const char*)F.w3 + (off & 0x38u));
                const fp2 md = hk_ld2(mrow + 64 * i);
                a01 = hk_fma2(md, A1, A0); d01 = hk_fma2(md, D1, D0);
                a2 = hk_fma(hk_lo(md), A[1][2], A[0][2]); d2 = hk_fma(hk_lo(md), D[1][2], D[0][2]);
            } else {
                const uint32_t word = i < 8 ? lo : hi;
                const int q = 4 * (i & 7);
                const uint32_t off = i == 0 ? ((word << 2) & 0x38u) : q == 0 ? ((word << 3) & 0x78u) : ((word >> (q - 3)) & 0x78u);
                wd = hk_ld2((const char*)F.w4 + off);
                a01 = A0; d01 = D0;
                a2 = A[0][2]; d2 = D[0][2];
            }
            const fp2 y01 = hk_add2(hk_fma2(d01, wd, a01), K);
            const float y2 = hk_add(hk_fma(d2, hk_lo(wd), a2), finish_magic<SIGNED>());
            y[3 * k + 0] = finish_y<SIGNED>(hk_lo(y01));
            y[3 * k + 1] = finish_y<SIGNED>(hk_hi(y01));
            y[3 * k + 2] = finish_y<SIGNED>(y2);
        }
        uint32_t o[6];
#pragma unroll
        for (int j = 0; j < 6; j++) o[j] = hk_prmt(y[2 * j], y[2 * j + 1], SIGNED ? 0x5410 : 0x6521); // halfs in bytes 0..1 / 1..2
        sink.row(r, o);
    }
}

// mode index 0..13 from the block's low five bits, 14 = reserved (bcdec writes zeros)
HKD uint32_t mode_of(uint32_t w0)
{
    const uint32_t code = w0 & 31u;
    if ((code & 2u) == 0) return code & 1u;
    if ((code & 3u) == 2u) return 2u + (code >> 2);
    return code < 16u ? 10u + (code >> 2) : 14u;
}

// The fourteen modes differ in the header only: the switch leaves the twelve unquantised endpoints, and the texel code (two
// thirds of the work) runs once for all two-region modes and once for all one-region modes, so a warp that straddles a
// mode boundary of the sorted tile diverges over the parse, not over the texels, and the kernel carries two texel bodies
// instead of fourteen.
template <bool SIGNED, class Sink>
HKD void decode_bc6h(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, const lean7::Tables& T, const FTab& F, Sink& sink)
{
    int e[12];
    uint32_t part = 0;
    const uint32_t mode = mode_of(w0);
    switch (mode) {
    case 0: endpoints<0, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 1: endpoints<1, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 2: endpoints<2, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 3: endpoints<3, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 4: endpoints<4, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 5: endpoints<5, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 6: endpoints<6, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 7: endpoints<7, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 8: endpoints<8, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 9: endpoints<9, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 10: endpoints<10, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 11: endpoints<11, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 12: endpoints<12, SIGNED>(w0, w1, w2, w3, e, part); break;
    case 13: endpoints<13, SIGNED>(w0, w1, w2, w3, e, part); break;
    default:
#pragma unroll
        for (int i = 0; i < 12; i++) e[i] = 0;
    }
    if (mode < 10) texels<SIGNED, true>(e, part, w2, w3, T, F, sink);
    else if (mode < 14) texels<SIGNED, false>(e, part, w2, w3, T, F, sink);
    else {
        const uint32_t z[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int r = 0; r < 4; r++) sink.row(r, z);
    }
}

} // namespace lean6
} // namespace hkbcn
