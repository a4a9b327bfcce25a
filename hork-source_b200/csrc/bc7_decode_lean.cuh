// bc7_decode_lean.cuh -- instruction-lean BC7 block decoders for modes 1, 3, 5, 6, 7 (sm_100a).
//
// Bit-exact with bc7decomp::unpack_bc7 (ThirdParty/bc7enc_rdo/bc7decomp.cpp:594-634), the decoder behind
// Hk::TextureBlockCompression::Decode_BC7 (Hork/Image/ImageEncoders.cpp:112-136).  Modes 0, 2 and 4 (never emitted by
// the reference's encoder presets, bc7enc.cpp:2428-2437) stay on the general decoder in bcn_decode.cu.
//
// The interpolation (l*(64-w) + h*w + 32) >> 6 runs on two 16-bit lanes per register: with A = l*256 + 128 and
// D = h - l (an ordinary 32-bit difference of the packed lanes), V = D*(4w) + A has its lanes equal to
// 4*(l*(64-w) + h*w + 32) <= 65408, so the results are bytes 1 and 3 of V and a single PRMT assembles the texel.
//   * two-subset modes (1, 3, 7): lanes = {subset 0, subset 1} of one channel; the 2 x 2^ib palette entries go to a
//     per-thread column of shared memory and each texel is one LDS addressed by (subset << 3 | index), a nibble stream
//     built once per block (index bits re-expanded to a fixed width by inserting the implicit anchor zeros).
//   * one-subset modes (5, 6): lanes = {r, g} and {b, a}; two (three for mode 5) IMADs and a PRMT per texel.
// Everything is plain integer code, compiled for the CPU by tools/decode_host_check.cpp and compared with the reference.
#pragma once
#include "bcn_decode_lean.cuh"
#ifdef HKBCN_HOST_EMU
#include "bcn_tables.cuh"
#endif

namespace hkbcn {
namespace lean7 {

// per-shape constants of the two-subset partitions, copied to shared memory by the kernel
struct Tables {
    uint32_t nib[64][2];  // bit 4k+3 of nib[s][h] = subset of texel 8h+k
    uint32_t ins3[64][2]; // 64-bit mask of the bits at/above the anchor's implicit zero in the 3-bit index stream (after texel 0's)
    uint32_t ins2[64];    // same for the 2-bit stream (32-bit)
};

// host-side fill (also used by the host check)
inline void fill_tables(Tables* T, const uint16_t* shape2, const uint8_t* anchor2)
{
    for (int s = 0; s < 64; s++) {
        T->nib[s][0] = T->nib[s][1] = 0;
        for (int i = 0; i < 16; i++)
            if ((shape2[s] >> i) & 1) T->nib[s][i >> 3] |= 8u << (4 * (i & 7));
        const int p3 = 3 * anchor2[s] + 2, p2 = 2 * anchor2[s] + 1;
        const uint64_t m3 = ~((1ull << p3) - 1ull);
        T->ins3[s][0] = (uint32_t)m3;
        T->ins3[s][1] = (uint32_t)(m3 >> 32);
        T->ins2[s] = ~((1u << p2) - 1u);
    }
}

// bits [POS, POS+N) of the 128-bit block (N <= 32), POS and N compile-time after inlining
HKD uint32_t getbits(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, int pos, int n)
{
    const int wi = pos >> 5, sh = pos & 31;
    const uint32_t a = wi == 0 ? w0 : wi == 1 ? w1 : wi == 2 ? w2 : w3;
    const uint32_t b = wi == 0 ? w1 : wi == 1 ? w2 : wi == 2 ? w3 : 0u;
    const uint32_t v = sh ? (sh + n > 32 ? hk_funnel_r(a, b, sh) : (a >> sh)) : a;
    return n >= 32 ? v : (v & ((1u << n) - 1u));
}

// four packed fields of NB bits (NB = 5, 6, 7) -> one per byte
template <int NB>
HKD uint32_t spread_to_bytes(uint32_t x)
{
    if (NB == 6) {
        x = x + 15u * (x & 0x00FFF000u);
        return x + 3u * (x & 0x0FC00FC0u);
    } else if (NB == 7) {
        x = x + 3u * (x & 0x0FFFC000u);
        return x + (x & 0x3F803F80u);
    } else { // 5
        x = x + 63u * (x & 0x000FFC00u);
        return x + 7u * (x & 0x03E003E0u);
    }
}

// Per-thread palette scratch of the two-subset modes: 16 entries, entry e of thread t at byte e * 1024 + 4 * t of a
// 16 KB shared array (entry-major: conflict-free whatever the lanes' entries are).  On the host: a plain array.
#ifdef HKBCN_HOST_EMU
struct Pal {
    uint32_t* a;
    void st(int e, uint32_t v) const { a[e] = v; }
    uint32_t ld(uint32_t c, int i) const { return a[(c >> (4 * i)) & 15u]; } // entry = nibble i of c
};
#else
struct Pal {
    char*    base; // the shared array (uniform)
    uint32_t t4;   // 4 * threadIdx.x (< 1024)
    __device__ __forceinline__ void st(int e, uint32_t v) const { *(uint32_t*)(base + (e * 1024 + t4)) = v; }
    __device__ __forceinline__ uint32_t ld(uint32_t c, int i) const
    {
        const uint32_t sh = (4 * i >= 10) ? (c >> (4 * i - 10)) : (c << (10 - 4 * i));
        return *(const uint32_t*)(base + ((sh & 0x3C00u) | t4)); // SHF + LOP3 + LDS [R + UR]
    }
};
#endif

// Palette of a two-subset mode: Xb = bytes [s0.lo, s0.hi, s1.lo, s1.hi] of one channel (8-bit endpoints).
// Writes entry (s << 3 | k) for k < 2^IB.
template <int IB, bool HAS_ALPHA>
HKD void two_subset_palette(uint32_t Rb, uint32_t Gb, uint32_t Bb, uint32_t Ab, const Pal& pal)
{
    const uint32_t K = 0x80u;
    const uint32_t Ar = hk_prmt(Rb, K, 0x2404), Ag = hk_prmt(Gb, K, 0x2404), Ab_ = hk_prmt(Bb, K, 0x2404), Aa = hk_prmt(Ab, K, 0x2404);
    const uint32_t Dr = hk_prmt(Rb, 0, 0x4341) - (Rb & 0x00FF00FFu), Dg = hk_prmt(Gb, 0, 0x4341) - (Gb & 0x00FF00FFu);
    const uint32_t Db = hk_prmt(Bb, 0, 0x4341) - (Bb & 0x00FF00FFu), Da = hk_prmt(Ab, 0, 0x4341) - (Ab & 0x00FF00FFu);
#pragma unroll
    for (int k = 0; k < (1 << IB); k++) {
        const uint32_t w4 = 4u * (IB == 2 ? (k * 64 + 1) / 3 : (k * 64 + 3) / 7); // {0,21,43,64}, {0,9,18,27,37,46,55,64}
        const uint32_t Vr = Dr * w4 + Ar, Vg = Dg * w4 + Ag, Vb = Db * w4 + Ab_;
        const uint32_t t = hk_prmt(Vr, Vg, 0x7351); // [r.s0, g.s0, r.s1, g.s1]
        uint32_t u;
        if (HAS_ALPHA) u = hk_prmt(Vb, Da * w4 + Aa, 0x7351);
        else u = hk_prmt(Vb, 0xFFFFFFFFu, 0x7371);   // [b.s0, FF, b.s1, FF]
        pal.st(k, hk_prmt(t, u, 0x5410));
        pal.st(8 + k, hk_prmt(t, u, 0x7632));
    }
}

// texels of a two-subset block from the combined nibble streams (subset << 3 | index)
HKD void two_subset_texels(uint32_t c0, uint32_t c1, const Pal& pal, uint32_t px[16])
{
#pragma unroll
    for (int i = 0; i < 8; i++) {
        px[i] = pal.ld(c0, i);
        px[8 + i] = pal.ld(c1, i);
    }
}

// ---- mode 1: 2 subsets, 6-bit RGB + shared p-bit per subset, 3-bit indices (bc7decomp.cpp:168-260 family)
HKD void decode_mode1(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, const Tables& T, const Pal& pal, uint32_t px[16])
{
    const uint32_t part = (w0 >> 2) & 63u;
    const uint32_t pb = ((w2 >> 16) & 1u) * 0x0101u + ((w2 >> 17) & 1u) * 0x01010000u;
    uint32_t Rb = spread_to_bytes<6>(w0 >> 8) * 2u + pb;
    uint32_t Gb = spread_to_bytes<6>(w1 & 0x00FFFFFFu) * 2u + pb;
    uint32_t Bb = spread_to_bytes<6>(hk_funnel_r(w1, w2, 24) & 0x00FFFFFFu) * 2u + pb;
    Rb = Rb * 2u + ((Rb >> 6) & 0x01010101u); // 7 -> 8 bits: (v << 1) | (v >> 6)
    Gb = Gb * 2u + ((Gb >> 6) & 0x01010101u);
    Bb = Bb * 2u + ((Bb >> 6) & 0x01010101u);
    two_subset_palette<3, false>(Rb, Gb, Bb, 0xFFFFFFFFu, pal);
    // 46 index bits from bit 82; implicit zeros after texel 0's two bits and the anchor's
    unsigned long long S = ((unsigned long long)(w3 >> 18) << 32) | hk_funnel_r(w2, w3, 18);
    S += S & ~3ull;
    S += S & (((unsigned long long)T.ins3[part][1] << 32) | T.ins3[part][0]);
    const uint32_t lo = (uint32_t)S, hi = (uint32_t)(S >> 32);
    const uint32_t c0 = lean::spread3(lo & 0x00FFFFFFu) | T.nib[part][0];
    const uint32_t c1 = lean::spread3(hk_funnel_r(lo, hi, 24) & 0x00FFFFFFu) | T.nib[part][1];
    two_subset_texels(c0, c1, pal, px);
}

// 16 two-bit indices (32 bits) -> two nibble streams
HKD void spread2_stream(uint32_t S, uint32_t& c0, uint32_t& c1)
{
    uint32_t s[2];
    lean::spread2(S, s);
    c0 = s[0];
    c1 = s[1];
}

// bit i of the low nibble -> bit 0 of byte i
HKD uint32_t pbits_to_bytes(uint32_t nib) { return ((nib & 15u) * 0x00204081u) & 0x01010101u; }

// ---- mode 3: 2 subsets, 7-bit RGB + p-bit per endpoint, 2-bit indices
HKD void decode_mode3(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, const Tables& T, const Pal& pal, uint32_t px[16])
{
    const uint32_t part = (w0 >> 4) & 63u;
    const uint32_t pb = pbits_to_bytes(hk_funnel_r(w2, w3, 30)); // bits 94..97
    const uint32_t Rb = spread_to_bytes<7>(hk_funnel_r(w0, w1, 10) & 0x0FFFFFFFu) * 2u + pb;
    const uint32_t Gb = spread_to_bytes<7>(hk_funnel_r(w1, w2, 6) & 0x0FFFFFFFu) * 2u + pb;
    const uint32_t Bb = spread_to_bytes<7>((w2 >> 2) & 0x0FFFFFFFu) * 2u + pb;
    two_subset_palette<2, false>(Rb, Gb, Bb, 0xFFFFFFFFu, pal);
    uint32_t S = w3 >> 2; // 30 index bits from bit 98
    S += S & ~1u;
    S += S & T.ins2[part];
    uint32_t c0, c1;
    spread2_stream(S, c0, c1);
    two_subset_texels(c0 | T.nib[part][0], c1 | T.nib[part][1], pal, px);
}

// ---- mode 7: 2 subsets, 5-bit RGBA + p-bit per endpoint, 2-bit indices
HKD void decode_mode7(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, const Tables& T, const Pal& pal, uint32_t px[16])
{
    const uint32_t part = (w0 >> 8) & 63u;
    const uint32_t pb = pbits_to_bytes(hk_funnel_r(w2, w3, 30));
    uint32_t Rb = spread_to_bytes<5>(hk_funnel_r(w0, w1, 14) & 0x000FFFFFu) * 2u + pb;
    uint32_t Gb = spread_to_bytes<5>((w1 >> 2) & 0x000FFFFFu) * 2u + pb;
    uint32_t Bb = spread_to_bytes<5>(hk_funnel_r(w1, w2, 22) & 0x000FFFFFu) * 2u + pb;
    uint32_t Ab = spread_to_bytes<5>((w2 >> 10) & 0x000FFFFFu) * 2u + pb;
    Rb = Rb * 4u + ((Rb >> 4) & 0x03030303u); // 6 -> 8 bits: (v << 2) | (v >> 4)
    Gb = Gb * 4u + ((Gb >> 4) & 0x03030303u);
    Bb = Bb * 4u + ((Bb >> 4) & 0x03030303u);
    Ab = Ab * 4u + ((Ab >> 4) & 0x03030303u);
    two_subset_palette<2, true>(Rb, Gb, Bb, Ab, pal);
    uint32_t S = w3 >> 2;
    S += S & ~1u;
    S += S & T.ins2[part];
    uint32_t c0, c1;
    spread2_stream(S, c0, c1);
    two_subset_texels(c0 | T.nib[part][0], c1 | T.nib[part][1], pal, px);
}

// packed lanes {x, y} (bytes 0 and 2) -> A = lanes * 256 + 128
HKD uint32_t lanes_A(uint32_t L) { return (L << 8) | 0x00800080u; }

// ---- mode 6: 1 subset, 7-bit RGBA + p-bit per endpoint, 4-bit indices
HKD void decode_mode6(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t px[16])
{
    // fields: r0 7, r1 14, g0 21, g1 28, b0 35, b1 42, a0 49, a1 56, p0 63, p1 64
    const uint32_t p0 = w1 >> 31, p1 = w2 & 1u;
    const uint32_t Lrg = ((((w0 >> 7) & 0x7Fu) | (((w0 >> 21) & 0x7Fu) << 16)) << 1) | (p0 * 0x00010001u);
    const uint32_t Hrg = ((((w0 >> 14) & 0x7Fu) | ((hk_funnel_r(w0, w1, 28) & 0x7Fu) << 16)) << 1) | (p1 * 0x00010001u);
    const uint32_t Lba = ((((w1 >> 3) & 0x7Fu) | (((w1 >> 17) & 0x7Fu) << 16)) << 1) | (p0 * 0x00010001u);
    const uint32_t Hba = ((((w1 >> 10) & 0x7Fu) | (((w1 >> 24) & 0x7Fu) << 16)) << 1) | (p1 * 0x00010001u);
    const uint32_t Arg = lanes_A(Lrg), Aba = lanes_A(Lba);
    const uint32_t Drg = (Hrg - Lrg) << 2, Dba = (Hba - Lba) << 2;
    // 63 index bits from bit 65: texel 0 has 3 bits (65..67), texel i >= 1 sits at bit 64 + 4i
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t word = i < 8 ? w2 : w3;
        const uint32_t idx = i == 0 ? ((w2 >> 1) & 7u) : ((word >> (4 * (i & 7))) & 15u);
        const uint32_t w = (idx * 68u + 8u) >> 4; // g_bc7_weights4
        px[i] = hk_prmt(Drg * w + Arg, Dba * w + Aba, 0x7531);
    }
}

// four-entry byte table [e0, e1, e2, e3] of one channel for the 2-bit weights {0, 21, 43, 64}: lanes = two entries,
// V = l * 4(64 - w) + h * 4w + 128, results in bytes 1 and 3
HKD uint32_t table2(uint32_t l, uint32_t h)
{
    const uint32_t V01 = l * (256u | (172u << 16)) + h * (84u << 16) + 0x00800080u;
    const uint32_t V23 = l * 84u + h * (172u | (256u << 16)) + 0x00800080u;
    return hk_prmt(V01, V23, 0x7531);
}

// ---- mode 5: 1 subset, 7-bit RGB + 8-bit A, 2-bit colour and 2-bit alpha indices, rotation
// Channel byte tables looked up four texels at a time (as BC3 in bcn_decode_lean.cuh); the rotation swaps which
// table / selector stream feeds which output channel, once per block.
HKD void decode_mode5(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t px[16])
{
    // fields: rot 6, r0 8, r1 15, g0 22, g1 29, b0 36, b1 43, a0 50, a1 58, colour indices 66 (31 bits), alpha indices 97 (31 bits)
    const uint32_t rot = (w0 >> 6) & 3u;
    uint32_t r0 = (w0 >> 8) & 0x7Fu, r1 = (w0 >> 15) & 0x7Fu, g0 = (w0 >> 22) & 0x7Fu, g1 = hk_funnel_r(w0, w1, 29) & 0x7Fu;
    uint32_t b0 = (w1 >> 4) & 0x7Fu, b1 = (w1 >> 11) & 0x7Fu;
    r0 = r0 * 2u + (r0 >> 6); r1 = r1 * 2u + (r1 >> 6);
    g0 = g0 * 2u + (g0 >> 6); g1 = g1 * 2u + (g1 >> 6);
    b0 = b0 * 2u + (b0 >> 6); b1 = b1 * 2u + (b1 >> 6);
    const uint32_t a0 = (w1 >> 18) & 0xFFu, a1 = hk_funnel_r(w1, w2, 26) & 0xFFu;
    const uint32_t tr = table2(r0, r1), tg = table2(g0, g1), tb = table2(b0, b1), ta = table2(a0, a1);
    uint32_t cS = hk_funnel_r(w2, w3, 2) & 0x7FFFFFFFu, aS = w3 >> 1;
    cS += cS & ~1u; // texel 0's implicit zero
    aS += aS & ~1u;
    uint32_t cs[2], as[2];
    lean::spread2(cS, cs);
    lean::spread2(aS, as);
    // output channel c = (table, selectors): the rotated channel trades places with alpha
    const uint32_t tR = rot == 1 ? ta : tr, tG = rot == 2 ? ta : tg, tB = rot == 3 ? ta : tb;
    const uint32_t tA = rot == 0 ? ta : rot == 1 ? tr : rot == 2 ? tg : tb;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t sR = rot == 1 ? as[h] : cs[h], sG = rot == 2 ? as[h] : cs[h], sB = rot == 3 ? as[h] : cs[h], sA = rot == 0 ? as[h] : cs[h];
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const int sh = 16 * q;
            lean::transpose_row(hk_prmt(tR, 0, sR >> sh), hk_prmt(tG, 0, sG >> sh), hk_prmt(tB, 0, sB >> sh), hk_prmt(tA, 0, sA >> sh), px + 4 * (2 * h + q));
        }
    }
}

// Decode one block with the lean paths; returns false for modes 0, 2, 4 (caller uses the general decoder).
HKD bool decode_bc7(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, const Tables& T, const Pal& pal, uint32_t px[16])
{
    const uint32_t low = w0 & 0xFFu;
    if (low & 1u) return false;
    if (low & 2u) { decode_mode1(w0, w1, w2, w3, T, pal, px); return true; }
    if (low & 4u) return false;
    if (low & 8u) { decode_mode3(w0, w1, w2, w3, T, pal, px); return true; }
    if (low & 16u) return false;
    if (low & 32u) { decode_mode5(w0, w1, w2, w3, px); return true; }
    if (low & 64u) { decode_mode6(w0, w1, w2, w3, px); return true; }
    if (low & 128u) { decode_mode7(w0, w1, w2, w3, T, pal, px); return true; }
#pragma unroll
    for (int i = 0; i < 16; i++) px[i] = 0; // reserved mode: transparent black (bc7decomp.cpp:628-630)
    return true;
}

#ifdef HKBCN_HOST_EMU
// host check entry: lean path or "not covered" (returns the reference's answer untouched by the caller)
static inline bool decode_bc7_host(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t px[16])
{
    static Tables T;
    static bool init = false;
    if (!init) { fill_tables(&T, kShape2, kAnchor2); init = true; }
    uint32_t scratch[16];
    const Pal pal{scratch};
    return decode_bc7(w0, w1, w2, w3, T, pal, px);
}
#endif

} // namespace lean7
} // namespace hkbcn
