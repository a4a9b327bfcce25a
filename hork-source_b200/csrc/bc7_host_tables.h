// bc7_host_tables.h -- host-side generation of the BC7 encoder's lookup tables (a12): the loops of
// bc7enc_compress_block_init (ThirdParty/bc7enc_rdo/bc7enc.cpp:124-285) and the least-squares weight tables (:52-56).
#pragma once
#include "bc7_thread.cuh"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace hkbcn {
namespace bc7 {

// returns false if the perceptual weight scaling of bc7enc_compress_block does not give the kernel's constants
inline bool fill_tables(Tables* H)
{
    memset(H, 0, sizeof(*H));
    // perceptual weight scaling of bc7enc_compress_block (bc7enc.cpp:2409-2417) must give the constants the kernel uses
    const float pr_weight = (.5f / (1.0f - .2126f)) * (.5f / (1.0f - .2126f));
    const float pb_weight = (.5f / (1.0f - .0722f)) * (.5f / (1.0f - .0722f));
    if ((int)(128 * 4.0f) != bc7::WY || (int)(64 * 4.0f * pr_weight) != bc7::WCR || (int)(16 * 4.0f * pb_weight) != bc7::WCB || 32 * 4 != bc7::WA)
        return false;
    for (uint32_t p = 0; p < 2; p++) {
        for (uint32_t i = 0; i < 32; i++) {
            uint32_t vl = ((i << 1) | p) << 2;
            vl |= (vl >> 6);
            uint32_t in = i + 1 < 31 ? i + 1 : 31;
            uint32_t vh = ((in << 1) | p) << 2;
            vh |= (vh >> 6);
            H->mid7[i][p] = (i == 31) ? 1.0f : (vl / 255.0f + vh / 255.0f) / 2.0f;
        }
        for (uint32_t i = 0; i < 64; i++) {
            uint32_t vl = ((i << 1) | p) << 1;
            vl |= (vl >> 7);
            uint32_t in = i + 1 < 63 ? i + 1 : 63;
            uint32_t vh = ((in << 1) | p) << 1;
            vh |= (vh >> 7);
            H->mid1[i][p] = (i == 63) ? 1.0f : (vl / 255.0f + vh / 255.0f) / 2.0f;
        }
    }
    for (uint32_t i = 0; i < 128; i++) {
        uint32_t vl = (i << 1);
        vl |= (vl >> 7);
        uint32_t in = i + 1 < 127 ? i + 1 : 127;
        uint32_t vh = in << 1;
        vh |= (vh >> 7);
        H->mid5[i] = (i == 127) ? 1.0f : (vl / 255.0f + vh / 255.0f) / 2.0f;
    }
    // mode 1: 6-bit endpoints + shared p-bit, selector 2 of 8 (weight 18)
    for (int c = 0; c < 256; c++)
        for (uint32_t lp = 0; lp < 2; lp++) {
            uint32_t best_err = 0xFFFF, bl = 0, bh = 0;
            for (uint32_t l = 0; l < 64; l++) {
                uint32_t low = ((l << 1) | lp) << 1;
                low |= (low >> 7);
                for (uint32_t h = 0; h < 64; h++) {
                    uint32_t high = ((h << 1) | lp) << 1;
                    high |= (high >> 7);
                    const int k = (int)((low * (64 - 18) + high * 18 + 32) >> 6);
                    const uint32_t err = (uint32_t)((k - c) * (k - c));
                    if (err < best_err) { best_err = err; bl = l; bh = h; }
                }
            }
            H->opt1[c][lp] = best_err | (bl << 16) | (bh << 24);
        }
    // mode 7: 5-bit endpoints + per-endpoint p-bits, selector 1 of 4 (weight 21)
    for (int c = 0; c < 256; c++)
        for (uint32_t hp = 0; hp < 2; hp++)
            for (uint32_t lp = 0; lp < 2; lp++) {
                uint32_t best_err = 0xFFFF, bl = 0, bh = 0;
                for (uint32_t l = 0; l < 32; l++) {
                    uint32_t low = ((l << 1) | lp) << 2;
                    low |= (low >> 6);
                    for (uint32_t h = 0; h < 32; h++) {
                        uint32_t high = ((h << 1) | hp) << 2;
                        high |= (high >> 6);
                        const int k = (int)((low * (64 - 21) + high * 21 + 32) >> 6);
                        const uint32_t err = (uint32_t)((k - c) * (k - c));
                        if (err < best_err) { best_err = err; bl = l; bh = h; }
                    }
                }
                H->opt7[c][hp][lp] = best_err | (bl << 16) | (bh << 24);
            }
    // g_bc7_weights{2,3,4}x (bc7enc.cpp:52-56): w*w, (1-w)*w, (1-w)*(1-w), w written to 6 decimals
    auto gen_wx = [](float (*dst)[4], int n) {
        for (int i = 0; i < n; i++) {
            const int k = n == 4 ? (i * 64 + 1) / 3 : n == 8 ? (i * 64 + 3) / 7 : (i * 64 + 7) / 15;
            const double w = k / 64.0;
            const double v[4] = {w * w, (1.0 - w) * w, (1.0 - w) * (1.0 - w), w};
            for (int c = 0; c < 4; c++) {
                char buf[32];
                snprintf(buf, sizeof(buf), "%.6f", v[c]);
                dst[i][c] = strtof(buf, nullptr);
            }
        }
    };
    gen_wx(H->wx4, 16);
    gen_wx(H->wx3, 8);
    gen_wx(H->wx2, 4);
    return true;
}

} // namespace bc7
} // namespace hkbcn
