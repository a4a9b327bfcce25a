// bcn_decode_lean.cuh -- instruction-lean BC1..BC5 block decoders (sm_100a), register-only, branch-free per block.
//
// Bit-exact with bcdec__color_block / __sharp_alpha_block / __smooth_alpha_block (ThirdParty/bcdec/bcdec.h:93-210), the
// decoders behind Hk::TextureBlockCompression::Decode_BC1..5 (Hork/Image/ImageEncoders.cpp:77-100).
//
// Why this exists: the decoders are HBM-bound only if a block costs about 100 instructions; the straightforward
// per-pixel select (16 x 4-way select, scalar /3 and /7) costs 370 (ncu: issue slots 74 % busy at 45 % of DRAM peak).
// Here the arithmetic runs on two 16-bit lanes per register, palettes are byte tables looked up with PRMT (four texels
// per instruction), and divisions are exact multiply-shifts:
//   565 -> 888      (c*527+23)>>6, (c*259+33)>>6 on lanes {endpoint0, endpoint1}
//   /3, n <= 766    (y + (y>>8) + 43) >> 8 with y = n*85                 (fits a 16-bit lane; checked for every n)
//   /7, n <= 1786   (n*9363) >> 16;   /5, n <= 1276   (n*13108) >> 16     (checked for every n)
// Every function is plain integer code on uint32_t, so tools/decode_host_check.cpp compiles it for the CPU and compares
// it with the oracle over exhaustive / random blocks.
#pragma once
#include <stdint.h>

#ifdef HKBCN_HOST_EMU
#define HKD static inline
// PRMT (default mode) on the host: byte k of the result = byte (sel nibble k & 7) of {a, b}; nibble bit 3 replicates the sign
static inline uint32_t hk_prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    const uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int k = 0; k < 4; k++) {
        const uint32_t n = (sel >> (4 * k)) & 0xF;
        uint32_t byte = (uint32_t)(v >> (8 * (n & 7))) & 0xFF;
        if (n & 8) byte = (byte & 0x80) ? 0xFF : 0x00;
        r |= byte << (8 * k);
    }
    return r;
}
static inline uint32_t hk_funnel_r(uint32_t lo, uint32_t hi, uint32_t s) { return (uint32_t)((((uint64_t)hi << 32) | lo) >> (s & 31)); }
#else
#include <cuda_runtime.h>
#define HKD __device__ __forceinline__
HKD uint32_t hk_prmt(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
HKD uint32_t hk_funnel_r(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_r(lo, hi, s); }
#endif

namespace hkbcn {
namespace lean {

// Byte tables of one block: entry e of channel c is byte e of tab[c] (colour: 4 entries; alpha: 8 entries in lo/hi).
struct ColorTab {
    uint32_t r, g, b, a; // a: 0xFFFFFFFF, or 0x00FFFFFF for BC1's 3-colour mode (entry 3 = transparent black)
};

// bcdec__color_block (bcdec.h:93-151).  w = c0 | c1 << 16 (the block's first word).
HKD ColorTab color_tables(uint32_t w, bool onlyOpaque)
{
    // lanes: low half = endpoint 0, high half = endpoint 1
    const uint32_t R = (w >> 11) & 0x001F001Fu, G = (w >> 5) & 0x003F003Fu, B = w & 0x001F001Fu;
    const uint32_t R8 = ((R * 527u + 0x00170017u) >> 6) & 0x00FF00FFu;
    const uint32_t G8 = ((G * 259u + 0x00210021u) >> 6) & 0x00FF00FFu;
    const uint32_t B8 = ((B * 527u + 0x00170017u) >> 6) & 0x00FF00FFu;
    const uint32_t Rs = hk_prmt(R8, 0, 0x1032), Gs = hk_prmt(G8, 0, 0x1032), Bs = hk_prmt(B8, 0, 0x1032);
    ColorTab t;
    if ((w & 0xFFFFu) > (w >> 16) || onlyOpaque) {
        // entries 2, 3 = (2a+b+1)/3, (a+2b+1)/3: lane0 = 2*e0+e1, lane1 = 2*e1+e0; y = (n+1)*85; result in bytes 1 and 3
        const uint32_t Yr = (2u * R8 + Rs) * 85u + 0x00550055u, Yg = (2u * G8 + Gs) * 85u + 0x00550055u, Yb = (2u * B8 + Bs) * 85u + 0x00550055u;
        const uint32_t Zr = Yr + hk_prmt(Yr, 0, 0x4341) + 0x002B002Bu;
        const uint32_t Zg = Yg + hk_prmt(Yg, 0, 0x4341) + 0x002B002Bu;
        const uint32_t Zb = Yb + hk_prmt(Yb, 0, 0x4341) + 0x002B002Bu;
        t.r = hk_prmt(R8, Zr, 0x7520); // [e0, e1, e2, e3]
        t.g = hk_prmt(G8, Zg, 0x7520);
        t.b = hk_prmt(B8, Zb, 0x7520);
        t.a = 0xFFFFFFFFu;
    } else {
        // 3-colour mode: entry 2 = (a+b+1)>>1, entry 3 = 0 with alpha 0
        const uint32_t Hr = ((R8 + Rs + 0x00010001u) >> 1) & 0xFFu, Hg = ((G8 + Gs + 0x00010001u) >> 1) & 0xFFu, Hb = ((B8 + Bs + 0x00010001u) >> 1) & 0xFFu;
        t.r = hk_prmt(R8, Hr, 0x7420); // [e0, e1, half, 0]  (byte 7 of {R8, Hr} is 0)
        t.g = hk_prmt(G8, Hg, 0x7420);
        t.b = hk_prmt(B8, Hb, 0x7420);
        t.a = 0x00FFFFFFu;
    }
    return t;
}

// 16 two-bit selectors -> PRMT selectors: s[0] low half = row 0, high half = row 1; s[1] = rows 2, 3 (nibble k = texel k)
HKD void spread2(uint32_t idx, uint32_t s[2])
{
#pragma unroll
    for (int h = 0; h < 2; h++) {
        uint32_t x = (idx >> (16 * h)) & 0xFFFFu;
        x = (x | (x << 8)) & 0x00FF00FFu;
        x = (x | (x << 4)) & 0x0F0F0F0Fu;
        x = (x | (x << 2)) & 0x33333333u;
        s[h] = x;
    }
}

// 24 bits = 8 three-bit selectors -> 8 nibbles (low half = first row, high half = second row)
HKD uint32_t spread3(uint32_t x)
{
    uint32_t t = x + 15u * (x & 0x00FFF000u);    // rows to 16-bit halves:      (x & 0xFFF) | (x & 0xFFF000) << 4
    t = t + 3u * (t & 0x0FC00FC0u);              // pairs to bytes:             (t & 0x3F)  | (t & 0xFC0) << 2
    return t + (t & 0x38383838u);                // selectors to nibbles:       (t & 0x07)  | (t & 0x38) << 1
}

// bcdec__smooth_alpha_block (bcdec.h:170-210): the 8-entry palette as bytes of lo (entries 0-3) and hi (4-7)
HKD void alpha_table(uint32_t w0, uint32_t& lo, uint32_t& hi)
{
    const uint32_t a0 = w0 & 0xFFu, a1 = (w0 >> 8) & 0xFFu;
    const bool eight = a0 > a1;
    const uint32_t d = eight ? 7u : 5u, M = eight ? 9363u : 13108u;
    const uint32_t step = a1 - a0; // mod 2^32; numerators stay in range
    uint32_t n = d * a0 + 1u + step;
    const uint32_t v2 = (n * M) >> 16; n += step;
    const uint32_t v3 = (n * M) >> 16; n += step;
    const uint32_t v4 = (n * M) >> 16; n += step;
    const uint32_t v5 = (n * M) >> 16; n += step;
    uint32_t v6 = (n * M) >> 16; n += step;
    uint32_t v7 = (n * M) >> 16;
    if (!eight) { v6 = 0; v7 = 255; }
    lo = (w0 & 0xFFFFu) | (v2 << 16) | (v3 << 24);
    hi = v4 | (v5 << 8) | (v6 << 16) | (v7 << 24);
}

// the 16 alpha values of a BC4-style block as four row words (byte k = texel k of the row)
HKD void alpha_rows(uint32_t w0, uint32_t w1, uint32_t row[4])
{
    uint32_t lo, hi;
    alpha_table(w0, lo, hi);
    const uint32_t s01 = spread3(hk_funnel_r(w0, w1, 16) & 0x00FFFFFFu); // block bits 16..39
    const uint32_t s23 = spread3(w1 >> 8);                                // block bits 40..63
    row[0] = hk_prmt(lo, hi, s01);
    row[1] = hk_prmt(lo, hi, s01 >> 16);
    row[2] = hk_prmt(lo, hi, s23);
    row[3] = hk_prmt(lo, hi, s23 >> 16);
}

// bcdec__sharp_alpha_block (bcdec.h:153-168): 4-bit alpha * 17; row r = 16 bits
HKD uint32_t sharp_alpha_row(uint32_t bits16)
{
    uint32_t x = bits16 & 0xFFFFu;
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    return x * 17u;
}

// four channel rows (byte k = texel k) -> four RGBA8 texels
HKD void transpose_row(uint32_t r, uint32_t g, uint32_t b, uint32_t a, uint32_t px[4])
{
    const uint32_t rg0 = hk_prmt(r, g, 0x5140), rg1 = hk_prmt(r, g, 0x7362);
    const uint32_t ba0 = hk_prmt(b, a, 0x5140), ba1 = hk_prmt(b, a, 0x7362);
    px[0] = hk_prmt(rg0, ba0, 0x5410);
    px[1] = hk_prmt(rg0, ba0, 0x7632);
    px[2] = hk_prmt(rg1, ba1, 0x5410);
    px[3] = hk_prmt(rg1, ba1, 0x7632);
}

// one texel row of a colour block: sel = the row's four selector nibbles (low 16 bits used), arow = the row's alpha bytes
HKD void color_row(const ColorTab& t, uint32_t sel, uint32_t arow, uint32_t px[4])
{
    transpose_row(hk_prmt(t.r, 0, sel), hk_prmt(t.g, 0, sel), hk_prmt(t.b, 0, sel), arow, px);
}

// ---- whole blocks: px[4 * row + k] = 0xAABBGGRR ------------------------------------------------
HKD void decode_bc1(uint32_t w0, uint32_t w1, uint32_t px[16])
{
    const ColorTab t = color_tables(w0, false);
    uint32_t s[2];
    spread2(w1, s);
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const uint32_t sel = s[r >> 1] >> (16 * (r & 1));
        color_row(t, sel, hk_prmt(t.a, 0, sel), px + 4 * r);
    }
}
HKD void decode_bc2(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t px[16])
{
    const ColorTab t = color_tables(w2, true);
    uint32_t s[2];
    spread2(w3, s);
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const uint32_t sel = s[r >> 1] >> (16 * (r & 1));
        color_row(t, sel, sharp_alpha_row((r < 2 ? w0 : w1) >> (16 * (r & 1))), px + 4 * r);
    }
}
HKD void decode_bc3(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t px[16])
{
    const ColorTab t = color_tables(w2, true);
    uint32_t s[2], a[4];
    spread2(w3, s);
    alpha_rows(w0, w1, a);
#pragma unroll
    for (int r = 0; r < 4; r++) color_row(t, s[r >> 1] >> (16 * (r & 1)), a[r], px + 4 * r);
}
// BC5 row: interleave the two channel rows -> 4 x RG8 = two words
HKD void bc5_row(uint32_t r, uint32_t g, uint32_t& lo, uint32_t& hi)
{
    lo = hk_prmt(r, g, 0x5140);
    hi = hk_prmt(r, g, 0x7362);
}

} // namespace lean
} // namespace hkbcn
