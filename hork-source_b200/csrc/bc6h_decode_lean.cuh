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

// One interpolated + finished channel value, returned with the half in bytes 1..2 (bits 8..23).
// unsigned: A = a*256 + 128, D = 4(b - a).  signed: A = a*64 + 32, D = b - a (bcdec.h:331-346 finish_unquantize).
template <bool SIGNED>
HKD uint32_t channel(int A, int D, int w)
{
    if (!SIGNED) {
        const uint32_t v = (uint32_t)(D * w + A) >> 8;
        return v * 124u;
    } else {
        const int v = (D * w + A) >> 6;
        const int a = v < 0 ? -v : v;
        const uint32_t mag = (uint32_t)(a * 31) >> 5;
        return (mag | ((v < 0 && mag != 0) ? 0x8000u : 0u)) << 8;
    }
}

// Texel rows of one block.  sink.row(r, w) receives the six output words of texel row r (RGB16F x 4, bcdec's layout).
template <bool SIGNED, bool TWO, class Sink>
HKD void texels(const int e[12], uint32_t part, uint32_t w2, uint32_t w3, const lean7::Tables& T, Sink& sink)
{
    int A[2][3], D[2][3];
#pragma unroll
    for (int s = 0; s < (TWO ? 2 : 1); s++)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const int a = e[c * 4 + 2 * s], b = e[c * 4 + 2 * s + 1];
            A[s][c] = SIGNED ? a * 64 + 32 : a * 256 + 128;
            D[s][c] = SIGNED ? b - a : 4 * (b - a);
        }
    uint32_t lo, hi, shape = 0;
    if (TWO) {
        // 46 index bits from bit 82; implicit zeros after texel 0's two bits and the anchor's (as BC7 mode 1)
        unsigned long long S = ((unsigned long long)(w3 >> 18) << 32) | hk_funnel_r(w2, w3, 18);
        S += S & ~3ull;
        S += S & (((unsigned long long)T.ins3[part][1] << 32) | T.ins3[part][0]);
        lo = (uint32_t)S;
        hi = (uint32_t)(S >> 32);
        // subset bits back from the nibble table: bit 4k+3 of nib[h] = subset of texel 8h+k
        shape = part; // resolved per texel below through T.nib
    } else {
        lo = w2 & ~1u; // texel 0: 3 bits at 65..67 -> a 4-bit field with its top bit clear, after >> 1 below
        hi = w3;
    }
    (void)shape;
    const uint32_t W3lo = 0x1B120900u, W3hi = 0x40372E25u; // aWeight3 = {0, 9, 18, 27, 37, 46, 55, 64}
    const uint32_t n0 = TWO ? T.nib[part][0] : 0u, n1 = TWO ? T.nib[part][1] : 0u;
#pragma unroll
    for (int r = 0; r < 4; r++) {
        uint32_t y[12];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = 4 * r + k;
            int w;
            int a0, a1, a2, d0, d1, d2;
            if (TWO) {
                const int p = 3 * i;
                const uint32_t idx = (p + 3 <= 32 ? (lo >> p) : p >= 32 ? (hi >> (p - 32)) : hk_funnel_r(lo, hi, p)) & 7u;
                w = (int)hk_prmt(W3lo, W3hi, idx);
                const uint32_t nb = i < 8 ? n0 : n1;
                const int m = -(int)((nb >> (4 * (i & 7) + 3)) & 1u); // all ones for subset 1
                a0 = A[0][0] ^ ((A[0][0] ^ A[1][0]) & m); d0 = D[0][0] ^ ((D[0][0] ^ D[1][0]) & m);
                a1 = A[0][1] ^ ((A[0][1] ^ A[1][1]) & m); d1 = D[0][1] ^ ((D[0][1] ^ D[1][1]) & m);
                a2 = A[0][2] ^ ((A[0][2] ^ A[1][2]) & m); d2 = D[0][2] ^ ((D[0][2] ^ D[1][2]) & m);
            } else {
                const uint32_t word = i < 8 ? lo : hi;
                const uint32_t idx = i == 0 ? ((word >> 1) & 7u) : ((word >> (4 * (i & 7))) & 15u);
                w = (int)((idx * 68u + 8u) >> 4); // aWeight4
                a0 = A[0][0]; a1 = A[0][1]; a2 = A[0][2];
                d0 = D[0][0]; d1 = D[0][1]; d2 = D[0][2];
            }
            y[3 * k + 0] = channel<SIGNED>(a0, d0, w);
            y[3 * k + 1] = channel<SIGNED>(a1, d1, w);
            y[3 * k + 2] = channel<SIGNED>(a2, d2, w);
        }
        uint32_t o[6];
#pragma unroll
        for (int j = 0; j < 6; j++) o[j] = hk_prmt(y[2 * j], y[2 * j + 1], 0x6521); // halfs sit in bytes 1..2
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
HKD void decode_bc6h(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, const lean7::Tables& T, Sink& sink)
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
    if (mode < 10) texels<SIGNED, true>(e, part, w2, w3, T, sink);
    else if (mode < 14) texels<SIGNED, false>(e, part, w2, w3, T, sink);
    else {
        const uint32_t z[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int r = 0; r < 4; r++) sink.row(r, z);
    }
}

} // namespace lean6
} // namespace hkbcn
