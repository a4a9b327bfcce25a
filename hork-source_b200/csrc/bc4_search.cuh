// bc4_search.cuh -- exhaustive-radius BC4 endpoint search for one 16-value block.
//
// Behaviour contract: byte-identical to rgbcx::encode_bc4_hq(dst, px, stride,
// search_rad = 3, BC4_USE_ALL_MODES) (ThirdParty/bc7enc_rdo/rgbcx.cpp:2730-2884),
// the call behind Encode_BC4/Encode_BC5/Encode_BC3 with bMaxQuality
// (Hork/Image/ImageEncoders.cpp:161-185).
//
// The reference walks 2 modes x 7 lo_delta x 7 hi_delta = 98 trials serially and keeps
// the FIRST trial with strictly lower total error; per pixel it keeps the FIRST of the 8
// palette values with the lowest squared error.  Its two early-outs (break on a zero
// distance, abandon a trial once its running error reaches the best) do not change that
// result, so here every trial is evaluated in full and trials are compared on the key
// (error << 7 | trial_index), whose minimum is exactly "first strictly lower".
//
// B200 mapping: two lanes per block -- lane parity = mode (8-value / 6-value), 49 trials
// each -- pixels held as 8 x u16x2 registers so that one VABSDIFF4 + one VIMNMX.U16x2
// handles a palette entry for two pixels, and IDP.4A squares-and-sums the minima.
#pragma once
#include <stdint.h>

namespace hkbcn {

// 8 palette values of a BC4 block with endpoints (e0, e1), as get_block_values
// (rgbcx.h:391-423): e0 > e1 -> 8-value mode (/7 floor), else 6-value mode (/5 floor, 0, 255).
__device__ __forceinline__ void bc4_palette(uint32_t e0, uint32_t e1, uint32_t v[8])
{
    v[0] = e0;
    v[1] = e1;
    if (e0 > e1) {
        v[2] = (e0 * 6 + e1) / 7;
        v[3] = (e0 * 5 + e1 * 2) / 7;
        v[4] = (e0 * 4 + e1 * 3) / 7;
        v[5] = (e0 * 3 + e1 * 4) / 7;
        v[6] = (e0 * 2 + e1 * 5) / 7;
        v[7] = (e0 + e1 * 6) / 7;
    } else {
        v[2] = (e0 * 4 + e1) / 5;
        v[3] = (e0 * 3 + e1 * 2) / 5;
        v[4] = (e0 * 2 + e1 * 3) / 5;
        v[5] = (e0 + e1 * 4) / 5;
        v[6] = 0;
        v[7] = 255;
    }
}

// Endpoints of trial (mode, lo_delta, hi_delta) (rgbcx.cpp:2806-2820). Returns false when
// the clamped endpoints coincide (trial skipped).
__device__ __forceinline__ bool bc4_trial_endpoints(int minv, int maxv, int mode, int lo, int hi, uint32_t& e0, uint32_t& e1)
{
    int a = min(max(maxv + hi, 0), 255);
    int b = min(max(minv + lo, 0), 255);
    if (a == b) return false;
    // mode 0 wants e0 > e1 (8 values); mode 1 wants e0 <= e1 (6 values)
    bool swap = (mode == 0) ? (a <= b) : (a > b);
    e0 = swap ? b : a;
    e1 = swap ? a : b;
    return true;
}

// Lane-local search over one mode. px2[i] = pixel(2i) | pixel(2i+1) << 16.
// Returns min over the 49 trials of (err << 7 | trial_index); 0xFFFFFFFF if none valid.
__device__ __forceinline__ uint32_t bc4hq_search_mode(const uint32_t px2[8], int minv, int maxv, int mode)
{
    uint32_t best = 0xFFFFFFFFu;
#pragma unroll 1
    for (int lo = -3; lo <= 3; lo++) {
#pragma unroll 1
        for (int hi = -3; hi <= 3; hi++) {
            uint32_t e0, e1;
            if (!bc4_trial_endpoints(minv, maxv, mode, lo, hi, e0, e1)) continue;
            uint32_t v[8];
            bc4_palette(e0, e1, v);
            uint32_t err = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] *= 0x00010001u;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                uint32_t m = __vabsdiffu4(px2[i], v[0]);
#pragma unroll
                for (int j = 1; j < 8; j++) m = __vminu2(m, __vabsdiffu4(px2[i], v[j]));
                err = __dp4a(m, m, err);
            }
            uint32_t key = (err << 7) | (uint32_t)(mode * 49 + (lo + 3) * 7 + (hi + 3));
            best = min(best, key);
        }
    }
    return best;
}

// Emit the 8-byte block for the winning trial index (selectors: first minimum per pixel,
// rgbcx.cpp:2847-2862; packing :2871-2881).
__device__ __forceinline__ uint2 bc4hq_emit(const uint32_t px2[8], int minv, int maxv, uint32_t trial)
{
    int mode = trial >= 49 ? 1 : 0;
    int t = trial - mode * 49;
    int lo = t / 7 - 3, hi = t % 7 - 3;
    uint32_t e0, e1;
    bc4_trial_endpoints(minv, maxv, mode, lo, hi, e0, e1);
    uint32_t v[8];
    bc4_palette(e0, e1, v);
    uint64_t sels = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        int p = (px2[i >> 1] >> ((i & 1) * 16)) & 0xFF;
        int d0 = p - (int)v[0];
        uint32_t be = d0 * d0, bi = 0;
#pragma unroll
        for (int j = 1; j < 8; j++) {
            int d = p - (int)v[j];
            uint32_t e = d * d;
            if (e < be) { be = e; bi = j; }
        }
        sels |= (uint64_t)bi << (i * 3);
    }
    uint64_t blk = (uint64_t)e0 | ((uint64_t)e1 << 8) | (sels << 16);
    return make_uint2((uint32_t)blk, (uint32_t)(blk >> 32));
}

// Solid block (rgbcx.cpp:2744-2754, mode-6 flag set): {v, v, selectors 0}.
__device__ __forceinline__ uint2 bc4_solid(uint32_t v) { return make_uint2(v | (v << 8), 0u); }

// Non-hq encoder rgbcx::encode_bc4 (rgbcx.cpp:2608-2728): endpoints = (max, min), per-pixel
// selector by the 7 scaled thresholds; always 8-value mode.
__device__ __forceinline__ uint2 bc4_fast(const uint32_t px2[8], int minv, int maxv)
{
    if (minv == maxv) return make_uint2((uint32_t)maxv | ((uint32_t)minv << 8), 0u);
    const int delta = maxv - minv;
    const int bias = 4 - minv * 14;
    uint64_t sels = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        int p = (px2[i >> 1] >> ((i & 1) * 16)) & 0xFF;
        int vv = p * 14 + bias;
        int c = (vv >= delta * 13) + (vv >= delta * 11) + (vv >= delta * 9) + (vv >= delta * 7) + (vv >= delta * 5) + (vv >= delta * 3) +
                (vv >= delta);
        // s_tran = {1,7,6,5,4,3,2,0}
        uint32_t s = (c == 0) ? 1u : (c == 7) ? 0u : (uint32_t)(8 - c);
        sels |= (uint64_t)s << (i * 3);
    }
    uint64_t blk = (uint64_t)maxv | ((uint64_t)minv << 8) | (sels << 16);
    return make_uint2((uint32_t)blk, (uint32_t)(blk >> 32));
}

} // namespace hkbcn
