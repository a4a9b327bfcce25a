// mip_resize.cu -- device-side mip cascade (sm_100a): level i from level i-1, the resampling the importers run on the CPU
// between two Compress* calls.
//
// Replaces the stbir_resize step of the import cascade, Hork/Image/Image.cpp:1755-1803 (uint8: BC1-5, BC7) and
// :1852-1899 (float: BC6H): ThirdParty/stb/stb_image_resize.h v0.97 stbir_resize (:2541) in its downsampling form --
// contributors and coefficients per input pixel (stbir__calculate_sample_range_downsample :1024,
// stbir__calculate_coefficients_downsample :1092), per-output normalisation (:1126), edge modes (:942), the five kernels
// (:762-850), uint8 decode with the sRGB table / division by 255 (:1252-1300), alpha weighting with STBIR_ALPHA_EPSILON
// (:1375-1400), un-weighting and encode (:1696-1790).  ImageMipmapConfig defaults: wrap + Mitchell (Image.h:361-371).
//
// Tap tables are built on the host with the reference's float arithmetic step by step; the kernel accumulates in the same
// order (input columns ascending inside a row, rows ascending) without FMA contraction, so the float sums are the
// reference's.  Differences left: the sRGB decode table is computed (pow) instead of typed in, and the sRGB encode is the
// exact transfer function instead of the reference's 104-entry piecewise-linear table -- contract +-1 LSB, >= 50 dB.
//
// One thread per output texel; a level is 8 x 8 taps at most per texel and a few hundred microseconds at 4096^2, against
// 25 ms of BC7 encode for the same chain.  Algorithmic HBM bytes per output texel: 4 source texels read + 1 written.
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>
#include "hkbcn_kernels.h"
#include "build/stbir_srgb.inc"

namespace hkbcn {

namespace {

float filter_kernel(int filter, float x, float scale)
{
    x = fabsf(x);
    switch (filter) {
    case 1: { // trapezoid ("box")
        const float halfscale = scale / 2, t = 0.5f + halfscale;
        if (x >= t) return 0;
        const float r = 0.5f - halfscale;
        return x <= r ? 1.0f : (t - x) / scale;
    }
    case 2: return x <= 1.0f ? 1 - x : 0;
    case 3: // cubic b-spline
        if (x < 1.0f) return (4 + x * x * (3 * x - 6)) / 6;
        if (x < 2.0f) return (8 + x * (-12 + x * (6 - x))) / 6;
        return 0;
    case 4: // catmull-rom
        if (x < 1.0f) return 1 - x * x * (2.5f - 1.5f * x);
        if (x < 2.0f) return 2 - x * (4 + x * (0.5f * x - 2.5f));
        return 0;
    default: // 5: mitchell
        if (x < 1.0f) return (16 + x * x * (21 * x - 36)) / 18;
        if (x < 2.0f) return (32 + x * (-60 + x * (36 - 7 * x))) / 18;
        return 0;
    }
}
float filter_support(int filter, float scale) { return filter == 1 ? 0.5f + scale / 2 : filter == 2 ? 1.0f : 2.0f; }

// stbir__edge_wrap_slow (:942); -1 = the texel reads as zero (STBIR_EDGE_ZERO)
int edge_index(int edge, int n, int max)
{
    if (n >= 0 && n < max) return n;
    switch (edge) {
    case 1: return n < 0 ? 0 : max - 1;                                                  // clamp
    case 2: if (n < 0) return n > -max ? -n : max - 1; return n >= 2 * max ? 0 : 2 * max - n - 1; // reflect
    case 3: { if (n >= 0) return n % max; int m = (-n) % max; return m ? max - m : 0; }   // wrap
    default: return -1;                                                                   // zero
    }
}

} // namespace

// Taps of one axis: output i reads idx[i*T + k] with weight w[i*T + k], k < T (padded with weight 0), ascending input order.
// The contributors of an output are consecutive inputs first[i] .. first[i] + cnt[i] - 1 (before edge mapping; map[] maps them).
// Upsampling (out_size > in_size: the -resample pre-pass, CreateImage's resample to a power of two): stbir looks at every
// OUTPUT pixel instead (stbir__calculate_sample_range_upsample :1009, stbir__calculate_coefficients_upsample :1038): the
// inputs inside the filter's support around the output's centre, kernel(centre - input centre, 1 / scale), leading and
// trailing zero weights dropped, the rest normalised to sum 1.
static void build_axis_taps_up(int in_size, int out_size, int filter, int edge, AxisTaps& t)
{
    const float scale = (float)out_size / (float)in_size, shift = 0.0f;
    const float out_radius = filter_support(filter, 1 / scale) * scale;
    std::vector<std::vector<std::pair<int, float>>> per_out((size_t)out_size);
    int lo_n = 0, hi_n = in_size - 1;
    for (int n = 0; n < out_size; n++) {
        const float out_center = (float)n + 0.5f;
        const float in_lo = (out_center - out_radius + shift) / scale, in_hi = (out_center + out_radius + shift) / scale;
        const float in_center_of_out = (out_center + shift) / scale;
        int first = (int)floor(in_lo + 0.5);
        const int last = (int)floor(in_hi - 0.5);
        std::vector<float> c;
        float total = 0;
        for (int i = 0; i <= last - first; i++) {
            const float in_center = (float)(i + first) + 0.5f;
            const float k = filter_kernel(filter, in_center_of_out - in_center, 1 / scale);
            if (c.empty() && !k) { // a leading zero weight moves the first contributor up
                first++;
                i--;
                continue;
            }
            c.push_back(k);
            total += k;
        }
        const float fs = 1 / total;
        for (float& k : c) k *= fs;
        while (!c.empty() && !c.back()) c.pop_back();
        for (size_t i = 0; i < c.size(); i++) per_out[(size_t)n].push_back({first + (int)i, c[i]});
        if (!c.empty()) {
            lo_n = std::min(lo_n, first);
            hi_n = std::max(hi_n, first + (int)c.size() - 1);
        }
    }
    int T = 0;
    for (auto& v : per_out) T = std::max(T, (int)v.size());
    const int margin = std::max(-lo_n, hi_n - (in_size - 1));
    t.T = T;
    t.margin = margin;
    t.uniform = false;
    t.idx.assign((size_t)out_size * T, 0);
    t.w.assign((size_t)out_size * T, 0.0f);
    t.first.assign((size_t)out_size, 0);
    t.cnt.assign((size_t)out_size, 0);
    t.map.assign((size_t)(in_size + 2 * margin), 0);
    for (int n = -margin; n < in_size + margin; n++) {
        const int e = edge_index(edge, n, in_size);
        t.map[(size_t)(n + margin)] = e < 0 ? 0 : e;
    }
    for (int i = 0; i < out_size; i++) {
        t.first[(size_t)i] = per_out[(size_t)i].empty() ? 0 : per_out[(size_t)i][0].first;
        t.cnt[(size_t)i] = (int)per_out[(size_t)i].size();
        for (size_t k = 0; k < per_out[(size_t)i].size(); k++) {
            const int e = edge_index(edge, per_out[(size_t)i][k].first, in_size);
            t.idx[(size_t)i * T + k] = e < 0 ? 0 : e;
            t.w[(size_t)i * T + k] = e < 0 ? 0.0f : per_out[(size_t)i][k].second;
        }
    }
}

void build_axis_taps(int in_size, int out_size, int filter, int edge, AxisTaps& t)
{
    if (out_size > in_size) { // stbir__use_upsampling: scale > 1
        build_axis_taps_up(in_size, out_size, filter, edge, t);
        return;
    }
    std::vector<int>& idx = t.idx;
    std::vector<float>& w = t.w;
    int& T = t.T;
    const float scale = (float)out_size / (float)in_size, shift = 0.0f;
    const float support = filter_support(filter, scale);
    const float in_radius = support / scale;
    const int margin = (int)ceil(support * 2 / scale) / 2;
    std::vector<std::vector<std::pair<int, float>>> per_out((size_t)out_size);
    for (int n = -margin; n < in_size + margin; n++) {
        const float center = (float)n + 0.5f;
        const float lo = (center - in_radius) * scale - shift, hi = (center + in_radius) * scale - shift;
        const float out_center_of_in = center * scale - shift;
        const int first = (int)floor(lo + 0.5), last0 = (int)floor(hi - 0.5);
        std::vector<float> c;
        for (int i = first; i <= last0; i++) c.push_back(filter_kernel(filter, ((float)i + 0.5f) - out_center_of_in, scale) * scale);
        int last = last0;
        for (int i = last0 - first; i >= 0; i--) { // trailing zero weights are dropped before normalisation (:1114-1121)
            if (c[(size_t)i] != 0.0f) break;
            last = first + i - 1;
        }
        for (int i = first; i <= last; i++)
            if (i >= 0 && i < out_size) per_out[(size_t)i].push_back({n, c[(size_t)(i - first)]});
    }
    T = 0;
    for (auto& v : per_out) {
        float total = 0;
        for (auto& p : v) total += p.second;
        const float s = 1 / total;
        for (auto& p : v) p.second *= s;
        if ((int)v.size() > T) T = (int)v.size();
    }
    idx.assign((size_t)out_size * T, 0);
    w.assign((size_t)out_size * T, 0.0f);
    t.margin = margin;
    t.first.assign((size_t)out_size, 0);
    t.cnt.assign((size_t)out_size, 0);
    t.map.assign((size_t)(in_size + 2 * margin), 0);
    for (int n = -margin; n < in_size + margin; n++) {
        const int e = edge_index(edge, n, in_size);
        t.map[(size_t)(n + margin)] = e < 0 ? 0 : e;
    }
    for (int i = 0; i < out_size; i++) {
        t.first[(size_t)i] = per_out[(size_t)i].empty() ? 0 : per_out[(size_t)i][0].first;
        t.cnt[(size_t)i] = (int)per_out[(size_t)i].size();
    }
    t.uniform = T <= 8 && in_size == 2 * out_size;
    for (int i = 0; i < out_size && t.uniform; i++) {
        if (t.cnt[(size_t)i] != T || t.first[(size_t)i] != 2 * i + t.first[0]) t.uniform = false;
        for (int k = 0; k < T && t.uniform; k++)
            if (per_out[(size_t)i][(size_t)k].second != per_out[0][(size_t)k].second) t.uniform = false;
    }
    for (int i = 0; i < out_size; i++)
        for (size_t k = 0; k < per_out[(size_t)i].size(); k++) {
            const int e = edge_index(edge, per_out[(size_t)i][k].first, in_size);
            idx[(size_t)i * T + k] = e < 0 ? 0 : e;
            w[(size_t)i * T + k] = e < 0 ? 0.0f : per_out[(size_t)i][k].second; // a zero texel adds nothing
            if (e < 0) t.uniform = false; // STBIR_EDGE_ZERO: the border outputs' weights differ
        }
}

// ---------------------------------------------------------------- kernels (ResizeParams: hkbcn_kernels.h)
__device__ __forceinline__ uint8_t encode_linear8(float f)
{
    f = f < 0 ? 0 : f > 1 ? 1 : f;
    return (uint8_t)(int)((double)(f * 255.0f) + 0.5);
}
#if HK_HAVE_STBIR_SRGB
// stbir__linear_to_srgb_uchar (stb_image_resize.h:675-698): a piecewise-linear fit of the transfer function over 104 segments
// (13 binades x 8), bias and scale of each in build/stbir_srgb.inc (tools/gen_stbir_tables.py reads them from the engine tree)
static __device__ const uint32_t g_srgb8_tab4[104] = {HK_FP32_TO_SRGB8_TAB4};
__device__ __forceinline__ uint8_t encode_srgb8(float in)
{
    constexpr uint32_t MINU = (127u - 13u) << 23; // 2^-13 maps to 0, 1 - eps to 255; written so that NaN maps to 0
    if (!(in > __uint_as_float(MINU))) in = __uint_as_float(MINU);
    if (in > __uint_as_float(0x3f7fffffu)) in = __uint_as_float(0x3f7fffffu);
    const uint32_t u = __float_as_uint(in);
    const uint32_t tab = g_srgb8_tab4[(u - MINU) >> 20];
    const uint32_t bias = (tab >> 16) << 9, scale = tab & 0xffffu, t = (u >> 12) & 0xffu;
    return (uint8_t)((bias + scale * t) >> 16);
}
bool mip_srgb_is_stbir() { return true; }
#else
// built without the engine tree: the exact transfer function, within 1 LSB of stbir's piecewise-linear fit
__device__ __forceinline__ uint8_t encode_srgb8(float f)
{
    f = !(f > 0) ? 0 : f > 1 ? 1 : f;
    const float v = f <= 0.0031308f ? f * 12.92f : 1.055f * powf(f, 1.0f / 2.4f) - 0.055f;
    return (uint8_t)(int)(v * 255.0f + 0.5f);
}
bool mip_srgb_is_stbir() { return false; }
#endif

template <int CH>
__global__ void __launch_bounds__(256) mip_resize_u8_kernel(const ResizeParams P)
{
    __shared__ float lut[256];
    __shared__ float alut[256]; // alpha decode i / 255.0f: the same IEEE division, once per CTA instead of once per tap (it was
                                // half of the kernel's instructions)
    lut[threadIdx.x] = P.lut[threadIdx.x];
    alut[threadIdx.x] = (float)threadIdx.x / 255.0f;
    __syncthreads();
    const int ox = blockIdx.x * 32 + (threadIdx.x & 31), oy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (ox >= P.out_w || oy >= P.out_h) return;
    float acc[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) acc[c] = 0;
    for (int ky = 0; ky < P.Ty; ky++) {
        const float wy = P.wy[oy * P.Ty + ky];
        const uint8_t* row = P.src + (size_t)P.iy[oy * P.Ty + ky] * P.in_w * CH;
        float h[CH];
#pragma unroll
        for (int c = 0; c < CH; c++) h[c] = 0;
        for (int kx = 0; kx < P.Tx; kx++) {
            const float wx = P.wx[ox * P.Tx + kx];
            const uint8_t* t = row + (size_t)P.ix[ox * P.Tx + kx] * CH;
            float v[CH];
            int q[CH];
            if constexpr (CH == 4) {
                const uchar4 q4 = *(const uchar4*)t;
                q[0] = q4.x; q[1] = q4.y; q[2] = q4.z; q[3] = q4.w;
            } else {
#pragma unroll
                for (int c = 0; c < CH; c++) q[c] = t[c];
            }
#pragma unroll
            for (int c = 0; c < CH; c++) v[c] = ((c == P.alpha) ? alut : lut)[q[c]];
            if (P.weight) {
                // alpha weighting (:1375-1400): the epsilon keeps the colour of fully transparent texels
                float a = 0;
#pragma unroll
                for (int c = 0; c < CH; c++) if (c == P.alpha) a = v[c];
                a += 1.0f / (1 << 20) / (1 << 20) / (1 << 20) / (1 << 20);
#pragma unroll
                for (int c = 0; c < CH; c++) v[c] = (c == P.alpha) ? a : v[c] * a;
            }
#pragma unroll
            for (int c = 0; c < CH; c++) h[c] += v[c] * wx;
        }
#pragma unroll
        for (int c = 0; c < CH; c++) acc[c] += h[c] * wy;
    }
    if (P.weight) {
        float a = 0;
#pragma unroll
        for (int c = 0; c < CH; c++) if (c == P.alpha) a = acc[c];
        const float r = a != 0 ? 1.0f / a : 0.0f;
#pragma unroll
        for (int c = 0; c < CH; c++) if (c != P.alpha) acc[c] *= r;
    }
    uint8_t o[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) o[c] = (P.srgb && c != P.alpha) ? encode_srgb8(acc[c]) : encode_linear8(acc[c]);
    uint8_t* d = P.dst + ((size_t)oy * P.out_w + ox) * CH;
    if constexpr (CH == 4) *(uchar4*)d = make_uchar4(o[0], o[1], o[2], o[3]);
    else {
#pragma unroll
        for (int c = 0; c < CH; c++) d[c] = o[c];
    }
}

// RGBA32F, no alpha weighting, linear (Image.cpp:1868-1875)
__global__ void __launch_bounds__(256) mip_resize_f32_kernel(const ResizeParams P)
{
    const int ox = blockIdx.x * 32 + (threadIdx.x & 31), oy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (ox >= P.out_w || oy >= P.out_h) return;
    float4 acc = make_float4(0, 0, 0, 0);
    for (int ky = 0; ky < P.Ty; ky++) {
        const float wy = P.wy[oy * P.Ty + ky];
        const float4* row = (const float4*)P.src + (size_t)P.iy[oy * P.Ty + ky] * P.in_w;
        float4 h = make_float4(0, 0, 0, 0);
        for (int kx = 0; kx < P.Tx; kx++) {
            const float wx = P.wx[ox * P.Tx + kx];
            const float4 v = row[P.ix[ox * P.Tx + kx]];
            h.x += v.x * wx; h.y += v.y * wx; h.z += v.z * wx; h.w += v.w * wx;
        }
        acc.x += h.x * wy; acc.y += h.y * wy; acc.z += h.z * wy; acc.w += h.w * wy;
    }
    ((float4*)P.dst)[(size_t)oy * P.out_w + ox] = acc;
}

// RGB32F (Float3 normals, 4-byte aligned): the averaged-normal cascade of the roughness bake (Image.cpp:2530-2542)
__global__ void __launch_bounds__(256) mip_resize_f32x3_kernel(const ResizeParams P)
{
    const int ox = blockIdx.x * 32 + (threadIdx.x & 31), oy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (ox >= P.out_w || oy >= P.out_h) return;
    float ax = 0, ay = 0, az = 0;
    for (int ky = 0; ky < P.Ty; ky++) {
        const float wy = P.wy[oy * P.Ty + ky];
        const float* row = (const float*)P.src + (size_t)P.iy[oy * P.Ty + ky] * P.in_w * 3;
        float hx = 0, hy = 0, hz = 0;
        for (int kx = 0; kx < P.Tx; kx++) {
            const float wx = P.wx[ox * P.Tx + kx];
            const float* v = row + (size_t)P.ix[ox * P.Tx + kx] * 3;
            hx += v[0] * wx; hy += v[1] * wx; hz += v[2] * wx;
        }
        ax += hx * wy; ay += hy * wy; az += hz * wy;
    }
    float* d = (float*)P.dst + ((size_t)oy * P.out_w + ox) * 3;
    d[0] = ax; d[1] = ay; d[2] = az;
}

// ---- the general resampler (ResampleRawImage, Image.cpp:2188-2243: any size to any size, 1..4 channels, uint8 or float, the
// alpha channel weighted or not): one thread per output texel, the same sums in the same order as above.  Used for the
// -resample / resample-to-a-power-of-two pre-pass, which runs once per image; the cascade keeps its own kernels.
template <int CH, bool F32>
__global__ void __launch_bounds__(256) resample_gather_kernel(const ResizeParams P)
{
    __shared__ float lut[F32 ? 1 : 256], alut[F32 ? 1 : 256];
    if (!F32) {
        lut[threadIdx.x] = P.lut[threadIdx.x];
        alut[threadIdx.x] = (float)threadIdx.x / 255.0f;
        __syncthreads();
    }
    const int ox = blockIdx.x * 32 + (threadIdx.x & 31), oy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (ox >= P.out_w || oy >= P.out_h) return;
    float acc[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) acc[c] = 0;
    for (int ky = 0; ky < P.Ty; ky++) {
        const float wy = P.wy[oy * P.Ty + ky];
        const size_t row = (size_t)P.iy[oy * P.Ty + ky] * P.in_w;
        float h[CH];
#pragma unroll
        for (int c = 0; c < CH; c++) h[c] = 0;
        for (int kx = 0; kx < P.Tx; kx++) {
            const float wx = P.wx[ox * P.Tx + kx];
            const size_t t = (row + (size_t)P.ix[ox * P.Tx + kx]) * CH;
            float v[CH];
#pragma unroll
            for (int c = 0; c < CH; c++) {
                if (F32) v[c] = ((const float*)P.src)[t + c];
                else v[c] = ((c == P.alpha) ? alut : lut)[P.src[t + c]];
            }
            if (P.weight) { // alpha weighting (:1375-1400)
                float a = 0;
#pragma unroll
                for (int c = 0; c < CH; c++) if (c == P.alpha) a = v[c];
                a += 1.0f / (1 << 20) / (1 << 20) / (1 << 20) / (1 << 20);
#pragma unroll
                for (int c = 0; c < CH; c++) v[c] = (c == P.alpha) ? a : v[c] * a;
            }
#pragma unroll
            for (int c = 0; c < CH; c++) h[c] += v[c] * wx;
        }
#pragma unroll
        for (int c = 0; c < CH; c++) acc[c] += h[c] * wy;
    }
    if (P.weight) {
        float a = 0;
#pragma unroll
        for (int c = 0; c < CH; c++) if (c == P.alpha) a = acc[c];
        const float r = a != 0 ? 1.0f / a : 0.0f;
#pragma unroll
        for (int c = 0; c < CH; c++) if (c != P.alpha) acc[c] *= r;
    }
    const size_t o = ((size_t)oy * P.out_w + ox) * CH;
#pragma unroll
    for (int c = 0; c < CH; c++) {
        if (F32) ((float*)P.dst)[o + c] = acc[c];
        else P.dst[o + c] = (P.srgb && c != P.alpha) ? encode_srgb8(acc[c]) : encode_linear8(acc[c]);
    }
}

cudaError_t launch_resample_general(const ResizeParams& P, int channels, bool f32, cudaStream_t stream)
{
    const dim3 grid((unsigned)(P.out_w + 31) / 32, (unsigned)(P.out_h + 7) / 8);
#define HK_RS(CH)                                                                           \
    case CH:                                                                                \
        if (f32) resample_gather_kernel<CH, true><<<grid, 256, 0, stream>>>(P);             \
        else resample_gather_kernel<CH, false><<<grid, 256, 0, stream>>>(P);                \
        break;
    switch (channels) {
        HK_RS(1) HK_RS(2) HK_RS(3) HK_RS(4)
    default: return cudaErrorInvalidValue;
    }
#undef HK_RS
    return cudaGetLastError();
}

// ---- the separable form: stbir's own structure (horizontal pass per input row, then the vertical pass), per output tile ----
// A CTA owns a 32 x 8 output tile.  It stages the (at most 72 x 24) input texels the tile's taps reach -- decoded once: table
// lookup, alpha weighting -- in shared memory, filters every staged row horizontally to the tile's 32 columns, then filters
// vertically.  Same products, same sums, same order as the gather kernels above (inputs ascending inside a row, rows
// ascending, no contraction), so the results are bit-identical to them and to stbir; what changes is the count: a gather
// thread evaluates Ty x Tx = 64 Mitchell taps and 64 decodes per output, here it is about 22 horizontal + 8 vertical taps
// and 6 decodes.
constexpr int MT_W = 32, MT_H = 8, MT_SX = 72, MT_SY = 24;

template <int CH, bool F32>
__global__ void __launch_bounds__(256) mip_resize_tile_kernel(const ResizeParams P)
{
    __shared__ __align__(16) float dec[MT_SY * MT_SX * CH];
    __shared__ __align__(16) float hbuf[MT_SY * MT_W * CH];
    __shared__ float s_wx[MT_W * 8], s_wy[MT_H * 8];
    __shared__ int s_fx[MT_W], s_cx[MT_W], s_fy[MT_H], s_cy[MT_H];
    __shared__ float lut[F32 ? 1 : 256], alut[F32 ? 1 : 256];
    const int tid = threadIdx.x;
    const int ox0 = blockIdx.x * MT_W, oy0 = blockIdx.y * MT_H;
    const int tw = min(MT_W, P.out_w - ox0), th = min(MT_H, P.out_h - oy0);
    if (!F32) {
        lut[tid] = P.lut[tid];
        alut[tid] = (float)tid / 255.0f; // alpha decode: the reference's IEEE division, once per CTA
    }
    if (tid < tw) {
        s_fx[tid] = P.firstx[ox0 + tid];
        s_cx[tid] = P.cntx[ox0 + tid];
    }
    if (tid >= 32 && tid < 32 + th) {
        s_fy[tid - 32] = P.firsty[oy0 + tid - 32];
        s_cy[tid - 32] = P.cnty[oy0 + tid - 32];
    }
    for (int i = tid; i < tw * P.Tx; i += 256) s_wx[i] = P.wx[(size_t)ox0 * P.Tx + i];
    for (int i = tid; i < th * P.Ty; i += 256) s_wy[i] = P.wy[(size_t)oy0 * P.Ty + i];
    __syncthreads();
    const int nx_lo = s_fx[0], sx = s_fx[tw - 1] + s_cx[tw - 1] - nx_lo;
    const int ny_lo = s_fy[0], sy = s_fy[th - 1] + s_cy[th - 1] - ny_lo;
    // ---- stage + decode
    for (int i = tid; i < sy * sx; i += 256) {
        const int r = i / sx, c = i - r * sx;
        const int iy = P.mapy[ny_lo + r + P.marginy], ix = P.mapx[nx_lo + c + P.marginx];
        float v[CH];
        if (F32) {
            const float* t = (const float*)P.src + ((size_t)iy * P.in_w + ix) * CH;
            if (CH == 4) {
                const float4 q = *(const float4*)t;
                v[0] = q.x; v[1] = q.y; v[2] = q.z; v[CH - 1] = q.w;
            } else {
#pragma unroll
                for (int k = 0; k < CH; k++) v[k] = t[k];
            }
        } else {
            const uint8_t* t = P.src + ((size_t)iy * P.in_w + ix) * CH;
            int q[CH];
            if (CH == 4) {
                const uchar4 q4 = *(const uchar4*)t;
                q[0] = q4.x; q[1] = q4.y; q[2] = q4.z; q[CH - 1] = q4.w;
            } else {
#pragma unroll
                for (int k = 0; k < CH; k++) q[k] = t[k];
            }
#pragma unroll
            for (int k = 0; k < CH; k++) v[k] = ((k == P.alpha) ? alut : lut)[q[k]];
            if (P.weight) {
                // alpha weighting (:1375-1400): the epsilon keeps the colour of fully transparent texels
                float a = 0;
#pragma unroll
                for (int k = 0; k < CH; k++) if (k == P.alpha) a = v[k];
                a += 1.0f / (1 << 20) / (1 << 20) / (1 << 20) / (1 << 20);
#pragma unroll
                for (int k = 0; k < CH; k++) v[k] = (k == P.alpha) ? a : v[k] * a;
            }
        }
        float* d = dec + ((size_t)r * MT_SX + c) * CH;
#pragma unroll
        for (int k = 0; k < CH; k++) d[k] = v[k];
    }
    __syncthreads();
    // ---- horizontal pass: every staged row to the tile's columns (inputs ascending)
    for (int i = tid; i < sy * MT_W; i += 256) {
        const int r = i >> 5, x = i & 31;
        if (x >= tw) continue;
        const float* row = dec + ((size_t)r * MT_SX + (s_fx[x] - nx_lo)) * CH;
        const float* wx = s_wx + x * P.Tx;
        float h[CH];
#pragma unroll
        for (int k = 0; k < CH; k++) h[k] = 0;
        const int n = s_cx[x];
        for (int kx = 0; kx < n; kx++) {
            const float w = wx[kx];
#pragma unroll
            for (int k = 0; k < CH; k++) h[k] += row[kx * CH + k] * w;
        }
        float* d = hbuf + ((size_t)r * MT_W + x) * CH;
#pragma unroll
        for (int k = 0; k < CH; k++) d[k] = h[k];
    }
    __syncthreads();
    // ---- vertical pass (rows ascending), un-weighting, encode
    const int x = tid & 31, y = tid >> 5;
    if (x >= tw || y >= th) return;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) acc[k] = 0;
    {
        const float* col = hbuf + ((size_t)(s_fy[y] - ny_lo) * MT_W + x) * CH;
        const float* wy = s_wy + y * P.Ty;
        const int n = s_cy[y];
        for (int ky = 0; ky < n; ky++) {
            const float w = wy[ky];
#pragma unroll
            for (int k = 0; k < CH; k++) acc[k] += col[(size_t)ky * MT_W * CH + k] * w;
        }
    }
    const int ox = ox0 + x, oy = oy0 + y;
    if (F32) {
        float* d = (float*)P.dst + ((size_t)oy * P.out_w + ox) * CH;
        if (CH == 4) *(float4*)d = make_float4(acc[0], acc[1], acc[2], acc[CH - 1]);
        else {
#pragma unroll
            for (int k = 0; k < CH; k++) d[k] = acc[k];
        }
        return;
    }
    if (P.weight) {
        float a = 0;
#pragma unroll
        for (int k = 0; k < CH; k++) if (k == P.alpha) a = acc[k];
        const float r = a != 0 ? 1.0f / a : 0.0f;
#pragma unroll
        for (int k = 0; k < CH; k++) if (k != P.alpha) acc[k] *= r;
    }
    uint8_t o[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) o[k] = (P.srgb && k != P.alpha) ? encode_srgb8(acc[k]) : encode_linear8(acc[k]);
    uint8_t* d = P.dst + ((size_t)oy * P.out_w + ox) * CH;
    if (CH == 4) *(uchar4*)d = make_uchar4(o[0], o[1], o[2], o[CH - 1]);
    else {
#pragma unroll
        for (int k = 0; k < CH; k++) d[k] = o[k];
    }
}

// ---- the same for the common case: both axes exact 2 : 1 with output-independent taps (every level of a power-of-two chain
// under wrap / clamp / reflect).  The weights come as kernel parameters, the staged columns are dealt over eight shared-memory
// lanes (column c at [c & 7][c >> 3]) so that a thread computing four adjacent outputs of a row reads its 6 + Tx inputs without
// bank conflicts -- 14 loads for 4 outputs instead of 8 each -- and stores during the decode are conflict-free as well.
// TT: taps per output on both axes as a compile-time constant (8: Mitchell / Catmull-Rom / B-spline at 2 : 1, the importers'
// default), so that the tap loops carry no run-time bounds; 0 = taken from the parameters.
template <int CH, bool F32, int TT>
__global__ void __launch_bounds__(256, 3) mip_resize_tile2_kernel(const ResizeParams P)
{
    const int TX = TT ? TT : P.Tx, TY = TT ? TT : P.Ty;
    constexpr int OCT = MT_SX / 8; // 9 slots per lane of columns
    __shared__ __align__(16) float dec[MT_SY * MT_SX * CH];
    __shared__ __align__(16) float hbuf[MT_SY * MT_W * CH];
    __shared__ int s_mx[MT_SX], s_my[MT_SY];
    __shared__ float lut[F32 ? 1 : 256], alut[F32 ? 1 : 256];
    const int tid = threadIdx.x;
    const int ox0 = blockIdx.x * MT_W, oy0 = blockIdx.y * MT_H;
    const int tw = min(MT_W, P.out_w - ox0), th = min(MT_H, P.out_h - oy0);
    const int nx_lo = 2 * ox0 + P.x0, sx = 2 * (tw - 1) + TX;
    const int ny_lo = 2 * oy0 + P.y0, sy = 2 * (th - 1) + TY;
    if (!F32) {
        lut[tid] = P.lut[tid];
        alut[tid] = (float)tid / 255.0f; // alpha decode: the reference's IEEE division, once per CTA
    }
    if (tid < sx) s_mx[tid] = P.mapx[nx_lo + tid + P.marginx];
    if (tid >= 128 && tid < 128 + sy) s_my[tid - 128] = P.mapy[ny_lo + tid - 128 + P.marginy];
    __syncthreads();
    // ---- stage + decode
    const float inv_sx = 1.0f / (float)sx;
    for (int i = tid; i < sy * sx; i += 256) {
        const int r = (int)(((float)i + 0.5f) * inv_sx), c = i - r * sx; // i < 1728: the float quotient is exact enough to floor
        float v[CH];
        if (F32) {
            const float* t = (const float*)P.src + ((size_t)s_my[r] * P.in_w + s_mx[c]) * CH;
            if (CH == 4) {
                const float4 q = *(const float4*)t;
                v[0] = q.x; v[1] = q.y; v[2] = q.z; v[CH - 1] = q.w;
            } else {
#pragma unroll
                for (int k = 0; k < CH; k++) v[k] = t[k];
            }
        } else {
            const uint8_t* t = P.src + ((size_t)s_my[r] * P.in_w + s_mx[c]) * CH;
            int q[CH];
            if (CH == 4) {
                const uchar4 q4 = *(const uchar4*)t;
                q[0] = q4.x; q[1] = q4.y; q[2] = q4.z; q[CH - 1] = q4.w;
            } else {
#pragma unroll
                for (int k = 0; k < CH; k++) q[k] = t[k];
            }
#pragma unroll
            for (int k = 0; k < CH; k++) v[k] = ((k == P.alpha) ? alut : lut)[q[k]];
            if (P.weight) {
                // alpha weighting (:1375-1400): the epsilon keeps the colour of fully transparent texels
                float a = 0;
#pragma unroll
                for (int k = 0; k < CH; k++) if (k == P.alpha) a = v[k];
                a += 1.0f / (1 << 20) / (1 << 20) / (1 << 20) / (1 << 20);
#pragma unroll
                for (int k = 0; k < CH; k++) v[k] = (k == P.alpha) ? a : v[k] * a;
            }
        }
        float* d = dec + ((size_t)r * MT_SX + (c & 7) * OCT + (c >> 3)) * CH;
#pragma unroll
        for (int k = 0; k < CH; k++) d[k] = v[k];
    }
    __syncthreads();
    // ---- horizontal pass: thread = (staged row, group of four adjacent outputs); inputs ascending per output
    if (tid < sy * 8) {
        const int r = tid >> 3, g = tid & 7;
        const float* row = dec + (size_t)r * MT_SX * CH;
        float h[4][CH];
#pragma unroll
        for (int o = 0; o < 4; o++)
#pragma unroll
            for (int k = 0; k < CH; k++) h[o][k] = 0;
        // input t of the group (column 8 g + t of the staged row) feeds output o as its tap t - 2 o
#pragma unroll
        for (int t = 0; t < 14; t++) {
            if (t < 6 + TX) {
                float v[CH];
                const float* s = row + ((size_t)(t & 7) * OCT + g + (t >> 3)) * CH;
                if (CH == 4) {
                    const float4 q = *(const float4*)s;
                    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[CH - 1] = q.w;
                } else {
#pragma unroll
                    for (int k = 0; k < CH; k++) v[k] = s[k];
                }
#pragma unroll
                for (int o = 0; o < 4; o++) {
                    const int kx = t - 2 * o;
                    if (kx >= 0 && kx < 8 && kx < TX) {
#pragma unroll
                        for (int k = 0; k < CH; k++) h[o][k] += v[k] * P.ux[kx];
                    }
                }
            }
        }
#pragma unroll
        for (int o = 0; o < 4; o++) {
            float* d = hbuf + ((size_t)r * MT_W + 4 * g + o) * CH;
#pragma unroll
            for (int k = 0; k < CH; k++) d[k] = h[o][k];
        }
    }
    __syncthreads();
    // ---- vertical pass (rows ascending), un-weighting, encode
    const int x = tid & 31, y = tid >> 5;
    if (x >= tw || y >= th) return;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) acc[k] = 0;
    {
        const float* col = hbuf + ((size_t)(2 * y) * MT_W + x) * CH;
#pragma unroll
        for (int ky = 0; ky < 8; ky++) {
            if (ky < TY) {
#pragma unroll
                for (int k = 0; k < CH; k++) acc[k] += col[(size_t)ky * MT_W * CH + k] * P.uy[ky];
            }
        }
    }
    const int ox = ox0 + x, oy = oy0 + y;
    if (F32) {
        float* d = (float*)P.dst + ((size_t)oy * P.out_w + ox) * CH;
        if (CH == 4) *(float4*)d = make_float4(acc[0], acc[1], acc[2], acc[CH - 1]);
        else {
#pragma unroll
            for (int k = 0; k < CH; k++) d[k] = acc[k];
        }
        return;
    }
    if (P.weight) {
        float a = 0;
#pragma unroll
        for (int k = 0; k < CH; k++) if (k == P.alpha) a = acc[k];
        const float r = a != 0 ? 1.0f / a : 0.0f;
#pragma unroll
        for (int k = 0; k < CH; k++) if (k != P.alpha) acc[k] *= r;
    }
    uint8_t o[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) o[k] = (P.srgb && k != P.alpha) ? encode_srgb8(acc[k]) : encode_linear8(acc[k]);
    uint8_t* d = P.dst + ((size_t)oy * P.out_w + ox) * CH;
    if (CH == 4) *(uchar4*)d = make_uchar4(o[0], o[1], o[2], o[CH - 1]);
    else {
#pragma unroll
        for (int k = 0; k < CH; k++) d[k] = o[k];
    }
}

// ---- RGBA8, exact 2 : 1, eight taps per axis (the importers' default: Mitchell; also Catmull-Rom and the B-spline): the
// level-1 kernel of every colour chain.  Same arithmetic in the same order as above; what changes is who computes what:
//   stage       the 70 x 32 input texels of a 32 x 13 output tile as RAW words (4 bytes each, not 16)
//   horizontal  warp = group of 4 adjacent output columns, lane = staged row: a thread reads its row's 14 inputs with four
//               16-byte loads (row pitch 76 words: the eight rows of a quarter-warp fall into disjoint banks), decodes and
//               alpha-weights them in registers and produces 4 horizontal results -- 4 loads of raw words where the general
//               tile kernel spends 14 loads of decoded texels, and no 16-byte decoded copy of every texel in shared memory
//   vertical    thread = (column, pair of output rows): the pair shares 6 of its 10 inputs
// The horizontal buffer is column-major ([x][row], pitch 33 float4) so that both the row-parallel writes and the
// column-parallel reads are conflict-free.
constexpr int M3_W = 32, M3_H = 13, M3_SX = 2 * M3_W + 6, M3_SY = 2 * M3_H + 6, M3_P = 76, M3_HP = 33;
static_assert(M3_SY == 32, "one staged row per lane");

// packed f32x2 arithmetic (two channels of a texel per instruction, each half rounded as the scalar instruction rounds).
// ptxas contracts mul.f32x2 + add.f32x2 into one FFMA2 whatever -fmad says, so a product that is then added is taken as
// fma(a, b, +0) with the zero pair coming from a kernel parameter: round(a * b), as the multiply gives it (a -0 product
// would become +0, and added to an accumulator that started at +0 either gives the same sum).
typedef unsigned long long fp2;
__device__ __forceinline__ fp2 m_pack2(float a, float b) { fp2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void m_unpack2(fp2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ fp2 m_fma2(fp2 a, fp2 b, fp2 c) { fp2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ fp2 m_add2(fp2 a, fp2 b) { fp2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ fp2 m_mul2(fp2 a, fp2 b) { fp2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
#ifndef M3_MINB
#define M3_MINB 6  // resident CTAs per SM asked of ptxas (40 registers)
#endif
// HAS_A: the fourth channel is alpha (decoded linearly; colour weighted by it when P.weight); else it is a colour channel
template <bool HAS_A>
__global__ void __launch_bounds__(256, M3_MINB) mip_resize_rgba8_t8_kernel(const ResizeParams P)
{
    __shared__ __align__(16) uint32_t raw[M3_SY * M3_P];
    __shared__ __align__(16) float4 hb[M3_W * M3_HP + 2]; // + 2: the last pair of rows loads two entries it does not use
    __shared__ float lut[256], alut[256];
    __shared__ int s_mx[M3_SX + 2], s_my[M3_SY];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ox0 = blockIdx.x * M3_W, oy0 = blockIdx.y * M3_H;
    const int tw = min(M3_W, P.out_w - ox0), th = min(M3_H, P.out_h - oy0);
    const int nx_lo = 2 * ox0 + P.x0, sx = 2 * (tw - 1) + 8;
    const int ny_lo = 2 * oy0 + P.y0, sy = 2 * (th - 1) + 8;
    {
        const float v = P.lut[tid];
        lut[tid] = v;
        alut[tid] = P.srgb ? (float)tid / 255.0f : v; // the linear table is v / 255 itself
    }
    const int ax = nx_lo - 1;
    const bool interior = tw == M3_W && (ax & 3) == 0 && ax >= 0 && ax + 72 <= P.in_w && (P.in_w & 3) == 0 && ((uintptr_t)P.src & 15) == 0;
    const int* __restrict__ my = P.mapy + ny_lo + P.marginy;
    if (!interior) {
        if (tid < sx) s_mx[tid] = P.mapx[nx_lo + tid + P.marginx];
        if (tid >= 128 && tid < 128 + sy) s_my[tid - 128] = my[tid - 128];
        __syncthreads();
    }
    // ---- stage: word c of staged row r goes to raw[r * M3_P + c + 1] (the + 1 makes the 16-byte groups of an interior tile,
    // whose first texel 2 ox0 - 3 sits one word past a 16-byte boundary of the image row, land on 16-byte boundaries here)
    if (interior) {
        // interior in x: 18 aligned 16-byte loads per row, all of a thread's loads in flight before the first store
        uint4 v[3];
        int at[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const int i = tid + 256 * k, r = (i * 3641) >> 16, g = i - 18 * r; // i / 18 for i < 768
            at[k] = r < sy ? r * M3_P + 4 * g : -1;
            if (r < sy) v[k] = __ldg((const uint4*)(P.src + ((size_t)__ldg(my + r) * P.in_w + ax + 4 * g) * 4));
        }
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (at[k] >= 0) *(uint4*)(raw + at[k]) = v[k];
    } else {
        const uint32_t magic = ((1u << 20) + (uint32_t)sx - 1) / (uint32_t)sx; // i / sx = i * magic >> 20 for i < 32 sx, 8 <= sx <= 70
        for (int i = tid; i < sy * sx; i += 256) {
            const int r = (int)(((uint32_t)i * magic) >> 20), c = i - r * sx;
            raw[r * M3_P + c + 1] = __ldg((const uint32_t*)(P.src + ((size_t)s_my[r] * P.in_w + s_mx[c]) * 4));
        }
    }
    __syncthreads();
    // ---- horizontal pass: warp = four adjacent output columns (4 w .. 4 w + 3), lane = staged row
    if (lane < sy && 4 * warp < tw) {
        const uint32_t* row = raw + lane * M3_P + 8 * warp;
        uint32_t q[16];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint4 v = *(const uint4*)(row + 4 * j);
            q[4 * j + 0] = v.x; q[4 * j + 1] = v.y; q[4 * j + 2] = v.z; q[4 * j + 3] = v.w;
        }
        const fp2 Z = *(const fp2*)P.zero2;
        fp2 h01[4], h23[4];
#pragma unroll
        for (int o = 0; o < 4; o++) h01[o] = h23[o] = Z;
        // input t of the group feeds output o as its tap t - 2 o; inputs ascending per output
#pragma unroll
        for (int t = 0; t < 14; t++) {
            const uint32_t w = q[t + 1];
            const float v0 = lut[w & 255u], v1 = lut[(w >> 8) & 255u], v2 = lut[(w >> 16) & 255u];
            float v3 = (HAS_A ? alut : lut)[w >> 24];
            fp2 v01 = m_pack2(v0, v1), v23;
            if (HAS_A && P.weight) {
                // alpha weighting (:1375-1400): the epsilon keeps the colour of fully transparent texels
                v3 += 1.0f / (1 << 20) / (1 << 20) / (1 << 20) / (1 << 20);
                v01 = m_mul2(v01, m_pack2(v3, v3));
                v23 = m_pack2(v2 * v3, v3);
            } else v23 = m_pack2(v2, v3);
#pragma unroll
            for (int o = 0; o < 4; o++) {
                const int kx = t - 2 * o;
                if (kx >= 0 && kx < 8) {
                    const fp2 W = *(const fp2*)(P.ux2 + 2 * kx);
                    h01[o] = m_add2(h01[o], m_fma2(v01, W, Z));
                    h23[o] = m_add2(h23[o], m_fma2(v23, W, Z));
                }
            }
        }
#pragma unroll
        for (int o = 0; o < 4; o++) *(ulonglong2*)(hb + (4 * warp + o) * M3_HP + lane) = make_ulonglong2(h01[o], h23[o]);
    }
    __syncthreads();
    // ---- vertical pass (rows ascending), un-weighting, encode: thread = (column, output rows 2 p and 2 p + 1)
    const int x = lane, p = warp;
    if (x >= tw || 2 * p >= th) return;
    const ulonglong2* col = (const ulonglong2*)(hb + x * M3_HP + 4 * p);
    ulonglong2 in[10];
#pragma unroll
    for (int j = 0; j < 10; j++) in[j] = col[j];
    const fp2 Z = *(const fp2*)P.zero2;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const int y = 2 * p + half;
        if (y >= th) break;
        fp2 a01 = Z, a23 = Z;
#pragma unroll
        for (int ky = 0; ky < 8; ky++) {
            const fp2 W = *(const fp2*)(P.uy2 + 2 * ky);
            a01 = m_add2(a01, m_fma2(in[2 * half + ky].x, W, Z));
            a23 = m_add2(a23, m_fma2(in[2 * half + ky].y, W, Z));
        }
        float acc[4];
        m_unpack2(a01, acc[0], acc[1]);
        m_unpack2(a23, acc[2], acc[3]);
        if (HAS_A && P.weight) {
            const float a = acc[3];
            const float r = a != 0 ? 1.0f / a : 0.0f;
            acc[0] *= r; acc[1] *= r; acc[2] *= r;
        }
        uint8_t o[4];
#pragma unroll
        for (int k = 0; k < 3; k++) o[k] = P.srgb ? encode_srgb8(acc[k]) : encode_linear8(acc[k]);
        o[3] = (P.srgb && !HAS_A) ? encode_srgb8(acc[3]) : encode_linear8(acc[3]);
        *(uchar4*)(P.dst + ((size_t)(oy0 + y) * P.out_w + ox0 + x) * 4) = make_uchar4(o[0], o[1], o[2], o[3]);
    }
}

// ---- R8 / RG8, exact 2 : 1, four taps per axis (the triangle filter: what the importers mip roughness and packed normal maps
// with -- BC4 / BC5 chains), no alpha, linear.  A 256-thread CTA owns 64 x 16 outputs and stages their 130 x 34 inputs as raw
// bytes (4.6 KB for R8); a thread computes four adjacent outputs of one row from its 10 x 4 window: the horizontal pass of each
// of the four rows in registers, then the vertical one -- the same sums in the same order as the kernels above, no
// intermediate buffer, no second barrier.  Texel c of a staged row (c = 0 is input 2 ox0 + x0) sits at byte c * CH + M4_SH,
// which puts the 4-byte words of an interior tile on 4-byte boundaries and a thread's window on an 8- (R8) or 16-byte one.
constexpr int M4_W = 64, M4_H = 16, M4_SX = 2 * M4_W + 2, M4_SY = 2 * M4_H + 2;
template <int CH>
__global__ void __launch_bounds__(256) mip_resize_u8_t4_kernel(const ResizeParams P)
{
    constexpr int SH = CH == 1 ? 3 : 2;               // bytes in front of texel 0: (2 ox0 - 1) * CH - SH is a multiple of 4
    constexpr int PITCH = CH == 1 ? 144 : 272;        // bytes per staged row (>= M4_SX * CH + SH, a multiple of 16)
    constexpr int WORDS = (M4_SX * CH + SH + 3) / 4;  // 34 / 66 words per row
    __shared__ __align__(16) uint8_t raw[M4_SY * PITCH];
    __shared__ float lut[256];
    __shared__ int s_mx[M4_SX], s_my[M4_SY];
    const int tid = threadIdx.x;
    const int ox0 = blockIdx.x * M4_W, oy0 = blockIdx.y * M4_H;
    const int nx_lo = 2 * ox0 + P.x0, ny_lo = 2 * oy0 + P.y0;
    lut[tid] = P.lut[tid];
    const long long ab = (long long)nx_lo * CH - SH; // byte offset within an input row of the first staged word
    const bool interior = ab >= 0 && (ab & 3) == 0 && ab + 4 * WORDS <= (long long)P.in_w * CH && ((P.in_w * CH) & 3) == 0 && ((uintptr_t)P.src & 3) == 0;
    const int* __restrict__ my = P.mapy + ny_lo + P.marginy;
    if (interior) {
        // all of a thread's loads in flight before its first store
        constexpr int PER = (M4_SY * WORDS + 255) / 256;
        uint32_t v[PER];
        int at[PER];
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const int i = tid + 256 * k, r = i / WORDS, w = i - r * WORDS;
            at[k] = i < M4_SY * WORDS ? r * PITCH + 4 * w : -1;
            if (at[k] >= 0) v[k] = __ldg((const uint32_t*)(P.src + (size_t)__ldg(my + r) * P.in_w * CH + ab) + w);
        }
#pragma unroll
        for (int k = 0; k < PER; k++)
            if (at[k] >= 0) *(uint32_t*)(raw + at[k]) = v[k];
    } else {
        if (tid < M4_SX) s_mx[tid] = P.mapx[nx_lo + tid + P.marginx];
        if (tid >= 128 && tid < 128 + M4_SY) s_my[tid - 128] = my[tid - 128];
        __syncthreads();
        for (int i = tid; i < M4_SY * M4_SX; i += 256) {
            const int r = i / M4_SX, c = i - r * M4_SX;
            const uint8_t* t = P.src + ((size_t)s_my[r] * P.in_w + s_mx[c]) * CH;
#pragma unroll
            for (int k = 0; k < CH; k++) raw[r * PITCH + c * CH + SH + k] = t[k];
        }
    }
    __syncthreads();
    // thread = (row y, four columns 4 xg .. 4 xg + 3): inputs 8 xg .. 8 xg + 9 of staged rows 2 y .. 2 y + 3
    const int xg = tid & 15, y = tid >> 4;
    float acc[4][CH];
#pragma unroll
    for (int o = 0; o < 4; o++)
#pragma unroll
        for (int k = 0; k < CH; k++) acc[o][k] = 0.0f;
#pragma unroll
    for (int ky = 0; ky < 4; ky++) {
        const uint8_t* row = raw + (2 * y + ky) * PITCH + 8 * xg * CH; // the window's bytes start SH further on
        uint32_t wv[CH == 1 ? 4 : 6];
        if (CH == 1) {
            const uint2 a = *(const uint2*)row, b = *(const uint2*)(row + 8);
            wv[0] = a.x; wv[1] = a.y; wv[2] = b.x; wv[3] = b.y;
        } else {
            const uint4 a = *(const uint4*)row;
            const uint2 b = *(const uint2*)(row + 16);
            wv[0] = a.x; wv[1] = a.y; wv[2] = a.z; wv[3] = a.w; wv[4] = b.x; wv[5] = b.y;
        }
        float t[10][CH];
#pragma unroll
        for (int c = 0; c < 10; c++)
#pragma unroll
            for (int k = 0; k < CH; k++) {
                const int byte = c * CH + SH + k;
                t[c][k] = lut[(wv[byte >> 2] >> (8 * (byte & 3))) & 255u];
            }
        const float wy = P.uy[ky];
#pragma unroll
        for (int o = 0; o < 4; o++)
#pragma unroll
            for (int k = 0; k < CH; k++) {
                float h = 0.0f;
#pragma unroll
                for (int kx = 0; kx < 4; kx++) h += t[2 * o + kx][k] * P.ux[kx];
                acc[o][k] += h * wy;
            }
    }
    uint8_t* d = P.dst + ((size_t)(oy0 + y) * P.out_w + ox0 + 4 * xg) * CH;
    if (CH == 1) {
        *(uchar4*)d = make_uchar4(encode_linear8(acc[0][0]), encode_linear8(acc[1][0]), encode_linear8(acc[2][0]), encode_linear8(acc[3][0]));
    } else {
        *(uint2*)d = make_uint2((uint32_t)encode_linear8(acc[0][0]) | ((uint32_t)encode_linear8(acc[0][CH - 1]) << 8) | ((uint32_t)encode_linear8(acc[1][0]) << 16) |
                                    ((uint32_t)encode_linear8(acc[1][CH - 1]) << 24),
                                (uint32_t)encode_linear8(acc[2][0]) | ((uint32_t)encode_linear8(acc[2][CH - 1]) << 8) | ((uint32_t)encode_linear8(acc[3][0]) << 16) |
                                    ((uint32_t)encode_linear8(acc[3][CH - 1]) << 24));
    }
}

cudaError_t launch_mip_resize(const ResizeParams& P, int channels, bool f32, cudaStream_t stream)
{
    if (P.uniform && P.tileOk && !f32 && (channels == 1 || channels == 2) && P.Tx == 4 && P.Ty == 4 && P.alpha < 0 && !P.srgb && P.out_w % M4_W == 0 &&
        P.out_h % M4_H == 0 && ((uintptr_t)P.dst & 7) == 0 && !getenv("HKBCN_MIP_TILE2")) {
        const dim3 grid4((unsigned)P.out_w / M4_W, (unsigned)P.out_h / M4_H);
        if (channels == 1) mip_resize_u8_t4_kernel<1><<<grid4, 256, 0, stream>>>(P);
        else mip_resize_u8_t4_kernel<2><<<grid4, 256, 0, stream>>>(P);
        return cudaGetLastError();
    }
    if (P.uniform && P.tileOk && !f32 && channels == 4 && P.Tx == 8 && P.Ty == 8 && !getenv("HKBCN_MIP_TILE2")) {
        const dim3 grid3((unsigned)(P.out_w + M3_W - 1) / M3_W, (unsigned)(P.out_h + M3_H - 1) / M3_H);
        if (P.alpha == 3) mip_resize_rgba8_t8_kernel<true><<<grid3, 256, 0, stream>>>(P);
        else if (P.alpha < 0) mip_resize_rgba8_t8_kernel<false><<<grid3, 256, 0, stream>>>(P);
        else return cudaErrorInvalidValue;
        return cudaGetLastError();
    }
    if (P.uniform && P.tileOk) {
        const dim3 grid2((unsigned)(P.out_w + 31) / 32, (unsigned)(P.out_h + 7) / 8);
        const bool t8 = P.Tx == 8 && P.Ty == 8;
#define HK_T2(CH, F)                                                                            \
    do {                                                                                        \
        if (t8) mip_resize_tile2_kernel<CH, F, 8><<<grid2, 256, 0, stream>>>(P);                \
        else mip_resize_tile2_kernel<CH, F, 0><<<grid2, 256, 0, stream>>>(P);                   \
    } while (0)
        if (f32 && channels == 3) HK_T2(3, true);
        else if (f32 && channels == 4) HK_T2(4, true);
        else if (!f32 && channels == 1) HK_T2(1, false);
        else if (!f32 && channels == 2) HK_T2(2, false);
        else if (!f32 && channels == 4) HK_T2(4, false);
#undef HK_T2
        else return cudaErrorInvalidValue;
        return cudaGetLastError();
    }
    const dim3 grid((unsigned)(P.out_w + 31) / 32, (unsigned)(P.out_h + 7) / 8);
    if (P.tileOk && P.Tx <= 8 && P.Ty <= 8) {
        if (f32 && channels == 3) mip_resize_tile_kernel<3, true><<<grid, 256, 0, stream>>>(P);
        else if (f32 && channels == 4) mip_resize_tile_kernel<4, true><<<grid, 256, 0, stream>>>(P);
        else if (!f32 && channels == 1) mip_resize_tile_kernel<1, false><<<grid, 256, 0, stream>>>(P);
        else if (!f32 && channels == 2) mip_resize_tile_kernel<2, false><<<grid, 256, 0, stream>>>(P);
        else if (!f32 && channels == 4) mip_resize_tile_kernel<4, false><<<grid, 256, 0, stream>>>(P);
        else return cudaErrorInvalidValue;
        return cudaGetLastError();
    }
    if (f32 && channels == 3) mip_resize_f32x3_kernel<<<grid, 256, 0, stream>>>(P);
    else if (f32) mip_resize_f32_kernel<<<grid, 256, 0, stream>>>(P);
    else if (channels == 1) mip_resize_u8_kernel<1><<<grid, 256, 0, stream>>>(P);
    else if (channels == 2) mip_resize_u8_kernel<2><<<grid, 256, 0, stream>>>(P);
    else if (channels == 4) mip_resize_u8_kernel<4><<<grid, 256, 0, stream>>>(P);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

} // namespace hkbcn
