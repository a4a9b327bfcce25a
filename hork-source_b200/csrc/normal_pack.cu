// normal_pack.cu -- normal-map packers / unpackers and the roughness anti-aliasing bake on the device (sm_100a),
// the element-wise float work between load and the BC5 / BC4 / BC1 / BC3 encoders of the
// PBRImporter path, so the normal + roughness pipeline (BASELINE.json configs[1]) stays resident in HBM.
//
// Replaces (reference, scalar CPU loops):
//   PackNormals{RGBA_BC1, RG_BC5, Spheremap_BC5, Stereographic_BC5, Paraboloid_BC5, RGBA_BC3}_Compatible
//                                  Hork/Image/ImageEncoders.cpp:693-796
//   UnpackNormal*_Compatible       Hork/Image/ImageEncoders.cpp:798-878
//   NormalizeVectors               Hork/Image/Image.cpp:2251-2260  (v * 2 - 1, then TVector3::NormalizeSelf, VectorMath.h:888-899)
//   the vMF / Toksvig roughness update of CreateNormalAndRoughness, Hork/Image/Image.cpp:2526-2580
//   (== Tools/PBRImporter/Main.cpp:616-680)
//
// Arithmetic: the reference is plain IEEE single precision without contraction (Math::Round = std::round, half away from
// zero, Core/BaseMath.h:681-685; Math::Saturate = min(max(v, 0), 1), :794-798; Math::RSqrt = the 0x5f3759df bit trick with
// one Newton step, :949-962).  The library is built with --fmad=false and CUDA's sqrtf / division are correctly rounded
// by default, so every packer is bit-exact; the Toksvig branch calls sqrt() on floats, where the reference may run in
// double depending on the overload it resolves to -- contract +-1 LSB there.
// HBM-bound streaming kernels: 12 B/texel in, 2 or 4 B/texel out (packers).  The packers take four texels per thread with
// 16-byte loads when the buffers are 16-byte aligned (Float3 arrays themselves only promise 4).
#include <stdint.h>
#include "hkbcn_kernels.h"

namespace hkbcn {

namespace {

__device__ __forceinline__ float saturate_ref(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }
__device__ __forceinline__ uint8_t unorm8(float v) { return (uint8_t)roundf(saturate_ref(v) * 255.0f); }
__device__ __forceinline__ float sqrt_ref(float v) { return v > 0.0f ? sqrtf(v) : 0.0f; } // Math::Sqrt, Core/BaseMath.h:936-940
__device__ __forceinline__ float rsqrt_ref(float value)
{
    const float x2 = value * 0.5f;
    float f = __uint_as_float(0x5f3759dfu - (__float_as_uint(value) >> 1));
    f *= 1.5f - x2 * f * f;
    return f;
}

// one texel: the packed bytes in the low 16 (BC5 variants) or all 32 bits
template <int VARIANT>
__device__ __forceinline__ uint32_t pack_one(float x, float y, float z)
{
    if (VARIANT == NP_RGBA_BC1)
        return (uint32_t)unorm8(x * 0.5f + 0.5f) | ((uint32_t)unorm8(y * 0.5f + 0.5f) << 8) | ((uint32_t)unorm8(z * 0.5f + 0.5f) << 16) | 0xFF000000u;
    if (VARIANT == NP_RGBA_BC3)
        return 255u | ((uint32_t)unorm8(y * 0.5f + 0.5f) << 8) | ((uint32_t)unorm8(z * 0.5f + 0.5f) << 16) | ((uint32_t)unorm8(x * 0.5f + 0.5f) << 24);
    float u, v;
    if (VARIANT == NP_RG_BC5) {
        u = x * 0.5f + 0.5f;
        v = y * 0.5f + 0.5f;
    } else if (VARIANT == NP_SPHEREMAP_BC5) {
        const float denom = 1.0f / sqrt_ref(z * 8.0f + 8.0f);
        u = x * denom + 0.5f;
        v = y * denom + 0.5f;
    } else if (VARIANT == NP_STEREOGRAPHIC_BC5) {
        u = x / (1.0f + z) * 0.5f + 0.5f;
        v = y / (1.0f + z) * 0.5f + 0.5f;
    } else { // paraboloid
        const float a = (x * x) + (y * y), b = z, c = -1.0f;
        const float discriminant = b * b - 4.0f * a * c;
        const float t = (-b + sqrtf(discriminant)) / (2.0f * a);
        u = (x * t) * 0.5f + 0.5f;
        v = (y * t) * 0.5f + 0.5f;
    }
    return (uint32_t)unorm8(u) | ((uint32_t)unorm8(v) << 8);
}

// Four texels per thread: three 16-byte loads cover their 48 bytes of Float3 (a warp reads 1.5 KB contiguous), one 8- or
// 16-byte store takes the result.  VEC needs 16-byte aligned pointers; the tail and unaligned calls go one texel at a time.
template <int VARIANT, bool VEC>
__global__ void __launch_bounds__(256) pack_normals_kernel(const float* __restrict__ n, uint8_t* __restrict__ dst, size_t count)
{
    constexpr bool WIDE = VARIANT == NP_RGBA_BC1 || VARIANT == NP_RGBA_BC3;
    const size_t t = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (VEC) {
        const size_t i = t * 4;
        if (i + 4 <= count) {
            const float4* p = (const float4*)(n + 3 * i);
            const float4 a = __ldcs(p), b = __ldcs(p + 1), c = __ldcs(p + 2);
            const uint32_t q0 = pack_one<VARIANT>(a.x, a.y, a.z), q1 = pack_one<VARIANT>(a.w, b.x, b.y), q2 = pack_one<VARIANT>(b.z, b.w, c.x),
                           q3 = pack_one<VARIANT>(c.y, c.z, c.w);
            if (WIDE) __stcs((uint4*)(dst + 4 * i), make_uint4(q0, q1, q2, q3));
            else __stcs((uint2*)(dst + 2 * i), make_uint2(q0 | (q1 << 16), q2 | (q3 << 16)));
            return;
        }
        for (size_t k = i; k < count; k++) { // the last, partial group
            const uint32_t q = pack_one<VARIANT>(n[3 * k], n[3 * k + 1], n[3 * k + 2]);
            if (WIDE) *(uint32_t*)(dst + 4 * k) = q;
            else *(uint16_t*)(dst + 2 * k) = (uint16_t)q;
        }
    } else {
        if (t >= count) return;
        const uint32_t q = pack_one<VARIANT>(n[3 * t], n[3 * t + 1], n[3 * t + 2]);
        if (WIDE) *(uint32_t*)(dst + 4 * t) = q;
        else *(uint16_t*)(dst + 2 * t) = (uint16_t)q;
    }
}

template <int VARIANT>
__global__ void __launch_bounds__(256) unpack_normals_kernel(const uint8_t* __restrict__ src, float* __restrict__ n, size_t count)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    float x, y, z;
    if (VARIANT == NP_RGBA_BC1 || VARIANT == NP_RGBA_BC3) {
        const uchar4 d = *(const uchar4*)(src + 4 * i);
        if (VARIANT == NP_RGBA_BC1) {
            x = (d.x / 255.0f) * 2.0f - 1.0f;
            y = (d.y / 255.0f) * 2.0f - 1.0f;
            z = (d.z / 255.0f) * 2.0f - 1.0f;
        } else {
            x = d.w / 255.0f * 2.0f - 1.0f;
            y = d.y / 255.0f * 2.0f - 1.0f;
            z = d.z / 255.0f * 2.0f - 1.0f;
        }
    } else {
        const uchar2 d = *(const uchar2*)(src + 2 * i);
        if (VARIANT == NP_RG_BC5) {
            x = (d.x / 255.0f) * 2.0f - 1.0f;
            y = (d.y / 255.0f) * 2.0f - 1.0f;
            z = sqrt_ref(1.0f - (x * x + y * y));
        } else if (VARIANT == NP_SPHEREMAP_BC5) {
            const float px = (d.x / 255.0f) * 4.0f - 2.0f, py = (d.y / 255.0f) * 4.0f - 2.0f;
            const float f = px * px + py * py;
            const float s = sqrt_ref(1.0f - f / 4.0f);
            x = px * s;
            y = py * s;
            z = 1.0f - f / 2.0f;
        } else if (VARIANT == NP_STEREOGRAPHIC_BC5) {
            x = (d.x / 255.0f) * 2.0f - 1.0f;
            y = (d.y / 255.0f) * 2.0f - 1.0f;
            const float denom = 2.0f / (1 + saturate_ref(x * x + y * y));
            x *= denom;
            y *= denom;
            z = denom - 1.0f;
        } else { // paraboloid
            x = (d.x / 255.0f) * 2.0f - 1.0f;
            y = (d.y / 255.0f) * 2.0f - 1.0f;
            z = 1.0f - saturate_ref(x * x + y * y);
        }
    }
    n[3 * i] = x;
    n[3 * i + 1] = y;
    n[3 * i + 2] = z;
}

__global__ void __launch_bounds__(256) normalize_vectors_kernel(float* __restrict__ v, size_t count)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    float x = v[3 * i] * 2.0f - 1.0f, y = v[3 * i + 1] * 2.0f - 1.0f, z = v[3 * i + 2] * 2.0f - 1.0f;
    const float len = sqrtf(x * x + y * y + z * z);
    if (len != 0.0f) {
        const float inv = 1.0f / len;
        x *= inv;
        y *= inv;
        z *= inv;
    }
    v[3 * i] = x;
    v[3 * i + 1] = y;
    v[3 * i + 2] = z;
}

// one mip level: roughness[i] updated from the length of the averaged normal (Image.cpp:2546-2578)
template <int BAKE>
__global__ void __launch_bounds__(256) roughness_bake_kernel(const float* __restrict__ avg, uint8_t* __restrict__ rough, size_t count)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    const float x = avg[3 * i], y = avg[3 * i + 1], z = avg[3 * i + 2];
    const float r2 = x * x + y * y + z * z;
    if (!(r2 > 1e-8f && r2 < 1.0f)) return;
    if (BAKE == 1) { // vMF
        const float variance = 2.0f * rsqrt_ref(r2) * (1.0f - r2) / (3.0f - r2);
        const float rv = rough[i] / 255.0f;
        rough[i] = unorm8(sqrt_ref(rv * rv + variance));
    } else { // Toksvig
        const float r = sqrtf(r2);
        const uint8_t q = rough[i] < 1 ? 1 : rough[i];
        const float rr = q / 255.0f;
        const float specPower = 2.0f / (rr * rr) - 2.0f;
        const float ft = r / (specPower + r * (1.0f - specPower)); // Math::Lerp(specPower, 1, r) = a + frac * (b - a)
        rough[i] = unorm8(sqrtf(2.0f / (ft * specPower + 2.0f)));
    }
}

} // namespace

template <int VARIANT>
static void launch_pack_t(const float* d_normals, uint8_t* d_dst, size_t count, cudaStream_t st)
{
    const bool vec = !((size_t)d_normals & 15) && !((size_t)d_dst & 15);
    if (vec) pack_normals_kernel<VARIANT, true><<<(unsigned)(((count + 3) / 4 + 255) / 256), 256, 0, st>>>(d_normals, d_dst, count);
    else pack_normals_kernel<VARIANT, false><<<(unsigned)((count + 255) / 256), 256, 0, st>>>(d_normals, d_dst, count);
}

cudaError_t launch_pack_normals(int variant, const float* d_normals, uint8_t* d_dst, size_t count, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    switch (variant) {
    case NP_RGBA_BC1: launch_pack_t<NP_RGBA_BC1>(d_normals, d_dst, count, st); break;
    case NP_RG_BC5: launch_pack_t<NP_RG_BC5>(d_normals, d_dst, count, st); break;
    case NP_SPHEREMAP_BC5: launch_pack_t<NP_SPHEREMAP_BC5>(d_normals, d_dst, count, st); break;
    case NP_STEREOGRAPHIC_BC5: launch_pack_t<NP_STEREOGRAPHIC_BC5>(d_normals, d_dst, count, st); break;
    case NP_PARABOLOID_BC5: launch_pack_t<NP_PARABOLOID_BC5>(d_normals, d_dst, count, st); break;
    case NP_RGBA_BC3: launch_pack_t<NP_RGBA_BC3>(d_normals, d_dst, count, st); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_unpack_normals(int variant, const uint8_t* d_src, float* d_normals, size_t count, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    const unsigned g = (unsigned)((count + 255) / 256);
    switch (variant) {
    case NP_RGBA_BC1: unpack_normals_kernel<NP_RGBA_BC1><<<g, 256, 0, st>>>(d_src, d_normals, count); break;
    case NP_RG_BC5: unpack_normals_kernel<NP_RG_BC5><<<g, 256, 0, st>>>(d_src, d_normals, count); break;
    case NP_SPHEREMAP_BC5: unpack_normals_kernel<NP_SPHEREMAP_BC5><<<g, 256, 0, st>>>(d_src, d_normals, count); break;
    case NP_STEREOGRAPHIC_BC5: unpack_normals_kernel<NP_STEREOGRAPHIC_BC5><<<g, 256, 0, st>>>(d_src, d_normals, count); break;
    case NP_PARABOLOID_BC5: unpack_normals_kernel<NP_PARABOLOID_BC5><<<g, 256, 0, st>>>(d_src, d_normals, count); break;
    case NP_RGBA_BC3: unpack_normals_kernel<NP_RGBA_BC3><<<g, 256, 0, st>>>(d_src, d_normals, count); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_normalize_vectors(float* d_v, size_t count, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    normalize_vectors_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(d_v, count);
    return cudaGetLastError();
}

// RawImage::InvertGreen on Float3 texels (RawImage.cpp:745-760: g = 1 - g): DirectX -> OpenGL normal maps
__global__ void __launch_bounds__(256) invert_green_kernel(float* __restrict__ xyz, size_t count)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < count) xyz[3 * i + 1] = 1.0f - xyz[3 * i + 1];
}
cudaError_t launch_invert_green(float* d_xyz, size_t count, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    invert_green_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(d_xyz, count);
    return cudaGetLastError();
}

// ---- CreateImage's channel pass (Image.cpp:1152-1228): three-channel raw images gain an alpha of 255 / 1.0, BGR(A) images
// get R and B exchanged before compression.  T = uint8_t or float; one texel per thread.
template <typename T>
__global__ void __launch_bounds__(256) expand_channels_kernel(const T* __restrict__ src, T* __restrict__ dst, size_t count, int src_ch, int swap, T one)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    const T* s = src + i * (size_t)src_ch;
    T* d = dst + i * 4;
    const T r = s[swap ? 2 : 0], g = s[1], b = s[swap ? 0 : 2];
    const T a = src_ch == 4 ? s[3] : one;
    d[0] = r; d[1] = g; d[2] = b; d[3] = a;
}

cudaError_t launch_expand_channels(const void* d_src, void* d_dst, size_t count, int src_channels, bool swap_rb, bool f32, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    if (src_channels != 3 && src_channels != 4) return cudaErrorInvalidValue;
    const unsigned g = (unsigned)((count + 255) / 256);
    if (f32) expand_channels_kernel<float><<<g, 256, 0, st>>>((const float*)d_src, (float*)d_dst, count, src_channels, swap_rb ? 1 : 0, 1.0f);
    else expand_channels_kernel<uint8_t><<<g, 256, 0, st>>>((const uint8_t*)d_src, (uint8_t*)d_dst, count, src_channels, swap_rb ? 1 : 0, (uint8_t)255);
    return cudaGetLastError();
}

cudaError_t launch_roughness_bake(int bake, const float* d_avg_normals, uint8_t* d_rough, size_t count, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    const unsigned g = (unsigned)((count + 255) / 256);
    if (bake == 1) roughness_bake_kernel<1><<<g, 256, 0, st>>>(d_avg_normals, d_rough, count);
    else if (bake == 2) roughness_bake_kernel<2><<<g, 256, 0, st>>>(d_avg_normals, d_rough, count);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// ---- PBRImporter's ORMX image (Tools/PBRImporter/Main.cpp:86-190, CreateORMX): channel k of the result is channel ch[k] of
// map k (an RGBA8 image of the same size), or the default {255, 255, 0, 255} where no map was given
struct OrmxMaps { const uint8_t* map[4]; int ch[4]; };
__global__ void __launch_bounds__(256) ormx_gather_kernel(const OrmxMaps M, uchar4* __restrict__ dst, size_t count)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    uint8_t v[4] = {255, 255, 0, 255};
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (M.map[k]) v[k] = (uint8_t)(__ldg((const uint32_t*)M.map[k] + i) >> (8 * M.ch[k]));
    dst[i] = make_uchar4(v[0], v[1], v[2], v[3]);
}
cudaError_t launch_ormx_gather(const void* const d_maps[4], const int channel[4], void* d_dst, size_t count, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    OrmxMaps M;
    for (int k = 0; k < 4; k++) {
        if (channel[k] < 0 || channel[k] > 3) return cudaErrorInvalidValue;
        M.map[k] = (const uint8_t*)d_maps[k];
        M.ch[k] = channel[k];
    }
    ormx_gather_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(M, (uchar4*)d_dst, count);
    return cudaGetLastError();
}

// one channel of an RGBA8 image out to / back from a planar R8 image (the roughness channel of an ORMX chain: the bake works on R8)
__global__ void __launch_bounds__(256) channel_move_kernel(uint8_t* __restrict__ rgba, uint8_t* __restrict__ plane, size_t count, int ch, int insert)
{
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    if (insert) rgba[i * 4 + ch] = plane[i];
    else plane[i] = rgba[i * 4 + ch];
}
cudaError_t launch_channel_move(void* d_rgba, void* d_plane, size_t count, int channel, bool insert, cudaStream_t st)
{
    if (count == 0) return cudaSuccess;
    if (channel < 0 || channel > 3) return cudaErrorInvalidValue;
    channel_move_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>((uint8_t*)d_rgba, (uint8_t*)d_plane, count, channel, insert ? 1 : 0);
    return cudaGetLastError();
}

} // namespace hkbcn
