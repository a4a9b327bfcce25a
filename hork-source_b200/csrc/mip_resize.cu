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
#include <vector>
#include "hkbcn_kernels.h"

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
void build_axis_taps(int in_size, int out_size, int filter, int edge, std::vector<int>& idx, std::vector<float>& w, int& T)
{
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
    for (int i = 0; i < out_size; i++)
        for (size_t k = 0; k < per_out[(size_t)i].size(); k++) {
            const int e = edge_index(edge, per_out[(size_t)i][k].first, in_size);
            idx[(size_t)i * T + k] = e < 0 ? 0 : e;
            w[(size_t)i * T + k] = e < 0 ? 0.0f : per_out[(size_t)i][k].second; // a zero texel adds nothing
        }
}

// ---------------------------------------------------------------- kernels (ResizeParams: hkbcn_kernels.h)
__device__ __forceinline__ uint8_t encode_linear8(float f)
{
    f = f < 0 ? 0 : f > 1 ? 1 : f;
    return (uint8_t)(int)((double)(f * 255.0f) + 0.5);
}
__device__ __forceinline__ uint8_t encode_srgb8(float f)
{
    f = !(f > 0) ? 0 : f > 1 ? 1 : f;
    const float v = f <= 0.0031308f ? f * 12.92f : 1.055f * powf(f, 1.0f / 2.4f) - 0.055f;
    return (uint8_t)(int)(v * 255.0f + 0.5f);
}

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

cudaError_t launch_mip_resize(const ResizeParams& P, int channels, bool f32, cudaStream_t stream)
{
    const dim3 grid((unsigned)(P.out_w + 31) / 32, (unsigned)(P.out_h + 7) / 8);
    if (f32 && channels == 3) mip_resize_f32x3_kernel<<<grid, 256, 0, stream>>>(P);
    else if (f32) mip_resize_f32_kernel<<<grid, 256, 0, stream>>>(P);
    else if (channels == 1) mip_resize_u8_kernel<1><<<grid, 256, 0, stream>>>(P);
    else if (channels == 2) mip_resize_u8_kernel<2><<<grid, 256, 0, stream>>>(P);
    else if (channels == 4) mip_resize_u8_kernel<4><<<grid, 256, 0, stream>>>(P);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

} // namespace hkbcn
