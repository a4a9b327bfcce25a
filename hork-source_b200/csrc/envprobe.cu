// envprobe.cu -- cubemap GGX prefilter ("environment probe") and irradiance convolution on the device (sm_100a),
// the "prefiltered mip chain" of BASELINE.json configs[3], which the reference produces only by
// rendering a sphere mesh into a layered cubemap through OpenGL (unavailable on this box).
//
// Replaces the shader arithmetic of
//   Hork/Embedded/Shaders/gen/envprobegen.frag:44-111  (Hammersley, ImportanceSampleGGX, PrefilterEnvMap: 1024 samples,
//                                                       NdL-weighted sum of textureLod(Envmap, L, 0))
//   Hork/Embedded/Shaders/gen/irradiancegen.frag:46-76 (hemisphere sweep, SampleDelta 0.025, weight cos * sin)
// and the drivers' geometry: EnvProbeGenerator::Generate (Hork/RenderUtils/EnvProbeGenerator.cpp:182-240: level Lod of a
// 2^maxLod cube, roughness Lod / maxLod), IrradianceGenerator::Generate (IrradianceGenerator.cpp:160-200, 32^2 faces), the
// six face views Float4x4::sGetCubeFaceMatrices (Hork/Math/VectorMath.h:3397-3408) under the 90 degree projection of
// EnvProbeGenerator.cpp:53-65.  Worked through, those matrices are the standard cube-map face frames, with (u, v) the
// texel centre in [-1, 1] and v growing with the row index (upper-left clip control):
//   +X ( 1, -v, -u)   -X (-1, -v,  u)   +Y (u,  1,  v)   -Y (u, -1, -v)   +Z (u, -v,  1)   -Z (-u, -v, -1)
// The fragment's R is normalize(interpolated sphere position) = the direction of the view ray through the texel centre.
//
// The source cube is a plain device buffer of RGBA32F faces (no texture interop); textureLod(..., 0) is restated as a
// bilinear fetch with seamless edges: a tap that falls outside the face is taken from the texel its own direction lands on.
// GL leaves filtering precision to the implementation, so the contract against the CPU restatement is relative 1e-3
// (tests/test_gpu_envprobe.py), not bits; sums run in a different order here (32 lanes x 32 samples, then a shuffle tree).
//
// Kernels: prefilter -- one warp per output texel, 32 samples per lane; irradiance -- one 256-thread CTA per output texel
// over the 252 x 63 (phi, theta) grid, angles accumulated exactly as the shader's float loops do.
// Gather-bound on L2 (a 6 x 1024^2 RGBA32F source is 100 MB of the 126 MB L2); ~134 M + 97 M bilinear fetches per probe.
#include <math.h>
#include <stdint.h>
#include "hkbcn_kernels.h"

namespace hkbcn {

namespace {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }

__device__ __forceinline__ V3 face_dir(int f, float u, float v)
{
    switch (f) {
    case 0: return v3(1.0f, -v, -u);
    case 1: return v3(-1.0f, -v, u);
    case 2: return v3(u, 1.0f, v);
    case 3: return v3(u, -1.0f, -v);
    case 4: return v3(u, -v, 1.0f);
    default: return v3(-u, -v, -1.0f);
    }
}
// major-axis face selection and its (u, v) in [-1, 1] (the inverse of face_dir)
__device__ __forceinline__ int dir_face(V3 d, float& u, float& v)
{
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    if (ax >= ay && ax >= az) {
        const float m = 1.0f / ax;
        if (d.x > 0) { u = -d.z * m; v = -d.y * m; return 0; }
        u = d.z * m; v = -d.y * m; return 1;
    }
    if (ay >= az) {
        const float m = 1.0f / ay;
        if (d.y > 0) { u = d.x * m; v = d.z * m; return 2; }
        u = d.x * m; v = -d.z * m; return 3;
    }
    const float m = 1.0f / az;
    if (d.z > 0) { u = d.x * m; v = -d.y * m; return 4; }
    u = -d.x * m; v = -d.y * m; return 5;
}

struct Cube {
    const float4* texels; // [6][n][n]
    int n;
};
__device__ __forceinline__ float4 texel(const Cube& c, int f, int x, int y) { return __ldg(c.texels + ((size_t)f * c.n + y) * c.n + x); }
// a tap of face f at integer texel (x, y); outside the face it is the texel the tap's own direction lands on
__device__ __forceinline__ float4 tap(const Cube& c, int f, int x, int y)
{
    if ((unsigned)x < (unsigned)c.n && (unsigned)y < (unsigned)c.n) return texel(c, f, x, y);
    const float inv = 2.0f / (float)c.n;
    float u, v;
    const int g = dir_face(face_dir(f, ((float)x + 0.5f) * inv - 1.0f, ((float)y + 0.5f) * inv - 1.0f), u, v);
    int gx = (int)floorf((u * 0.5f + 0.5f) * (float)c.n), gy = (int)floorf((v * 0.5f + 0.5f) * (float)c.n);
    gx = min(max(gx, 0), c.n - 1);
    gy = min(max(gy, 0), c.n - 1);
    return texel(c, g, gx, gy);
}
// textureLod(Envmap, d, 0).rgb
__device__ __forceinline__ V3 sample_cube(const Cube& c, V3 d)
{
    float u, v;
    const int f = dir_face(d, u, v);
    const float x = (u * 0.5f + 0.5f) * (float)c.n - 0.5f, y = (v * 0.5f + 0.5f) * (float)c.n - 0.5f;
    const float x0f = floorf(x), y0f = floorf(y);
    const float fx = x - x0f, fy = y - y0f;
    const int x0 = (int)x0f, y0 = (int)y0f;
    const float4 a = tap(c, f, x0, y0), b = tap(c, f, x0 + 1, y0), p = tap(c, f, x0, y0 + 1), q = tap(c, f, x0 + 1, y0 + 1);
    const float w00 = (1.0f - fx) * (1.0f - fy), w10 = fx * (1.0f - fy), w01 = (1.0f - fx) * fy, w11 = fx * fy;
    return v3(a.x * w00 + b.x * w10 + p.x * w01 + q.x * w11, a.y * w00 + b.y * w10 + p.y * w01 + q.y * w11,
              a.z * w00 + b.z * w10 + p.z * w01 + q.z * w11);
}

constexpr float kPI = 3.1415926f; // the shaders' own constant

// ---- GGX prefilter: all levels of the probe in one launch, one warp per output texel
struct ProbeParams {
    Cube     src;
    float4*  dst;        // levels back to back, each [6][size][size]
    int      max_lod;
    uint32_t total;      // output texels over all levels
};

__global__ void __launch_bounds__(256) envprobe_prefilter_kernel(const ProbeParams P)
{
    const uint32_t warp = (blockIdx.x * 256u + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (warp >= P.total) return;
    // which level / face / texel
    uint32_t rest = warp, base = 0;
    int lod = 0, size = 1 << P.max_lod;
    for (;; lod++, size >>= 1) {
        const uint32_t n = 6u * (uint32_t)size * (uint32_t)size;
        if (rest < n) break;
        rest -= n;
        base += n;
    }
    const int f = (int)(rest / (uint32_t)(size * size)), t = (int)(rest % (uint32_t)(size * size)), py = t / size, px = t % size;
    const float roughness = (float)lod / (float)P.max_lod;
    const float inv = 2.0f / (float)size;
    const V3 R = normalize(face_dir(f, ((float)px + 0.5f) * inv - 1.0f, ((float)py + 0.5f) * inv - 1.0f));
    // ImportanceSampleGGX's tangent frame depends on N = R only
    const V3 up = fabsf(R.z) < 0.99f ? v3(0, 0, 1) : v3(1, 0, 0);
    const V3 T = normalize(cross(up, R)), B = cross(R, T);
    const float a = roughness * roughness;
    V3 sum = v3(0, 0, 0);
    float wsum = 0;
    const int NumSamples = 1024;
    for (int i = (int)lane; i < NumSamples; i += 32) {
        // Hammersley(i, 1024): the radical inverse of i (exact in binary), (i + 0.5) / n
        const float xi_x = (float)__brev((uint32_t)i) * 2.3283064365386963e-10f, xi_y = ((float)i + 0.5f) / (float)NumSamples;
        const float Phi = 2 * kPI * xi_x;
        const float CosTheta = sqrtf((1 - xi_y) / (1 + (a * a - 1) * xi_y));
        const float SinTheta = sqrtf(1 - CosTheta * CosTheta);
        float sp, cp;
        sincosf(Phi, &sp, &cp);
        const V3 H = normalize(T * (SinTheta * cp) + B * (SinTheta * sp) + R * CosTheta);
        const V3 L = normalize(H * (2 * dot(R, H)) - R);
        const float NdL = fminf(fmaxf(dot(R, L), 0.0f), 1.0f);
        if (NdL > 0) {
            const V3 c = sample_cube(P.src, L);
            sum = sum + c * NdL;
            wsum += NdL;
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        sum.x += __shfl_xor_sync(0xFFFFFFFFu, sum.x, o);
        sum.y += __shfl_xor_sync(0xFFFFFFFFu, sum.y, o);
        sum.z += __shfl_xor_sync(0xFFFFFFFFu, sum.z, o);
        wsum += __shfl_xor_sync(0xFFFFFFFFu, wsum, o);
    }
    if (lane == 0) P.dst[base + rest] = make_float4(sum.x / wsum, sum.y / wsum, sum.z / wsum, 1.0f);
}

// ---- irradiance: one CTA per output texel
struct IrrParams {
    Cube    src;
    float4* dst; // [6][size][size]
    int     size;
};
constexpr int IRR_MAX_PHI = 256, IRR_MAX_THETA = 64;

__global__ void __launch_bounds__(256) irradiance_kernel(const IrrParams P)
{
    __shared__ float s_phi[IRR_MAX_PHI], s_theta[IRR_MAX_THETA];
    __shared__ int s_np, s_nt;
    __shared__ float s_red[8][3];
    const float _2PI = 2.0f * kPI, HalfPI = 0.5f * kPI, SampleDelta = 0.025f;
    if (threadIdx.x == 0) { // the shader's `for (float Phi = 0; Phi < 2 PI; Phi += 0.025)`: the same float accumulation
        int n = 0;
        for (float Phi = 0.0f; Phi < _2PI && n < IRR_MAX_PHI; Phi += SampleDelta) s_phi[n++] = Phi;
        s_np = n;
    }
    if (threadIdx.x == 32) {
        int n = 0;
        for (float Theta = 0.0f; Theta < HalfPI && n < IRR_MAX_THETA; Theta += SampleDelta) s_theta[n++] = Theta;
        s_nt = n;
    }
    __syncthreads();
    const int tix = blockIdx.x, f = tix / (P.size * P.size), t = tix % (P.size * P.size), py = t / P.size, px = t % P.size;
    const float inv = 2.0f / (float)P.size;
    const V3 N = normalize(face_dir(f, ((float)px + 0.5f) * inv - 1.0f, ((float)py + 0.5f) * inv - 1.0f));
    V3 Up = v3(0.0f, 1.0f, 0.0f);
    const V3 Right = cross(Up, N);
    Up = cross(N, Right);
    const int np = s_np, nt = s_nt, total = np * nt;
    V3 sum = v3(0, 0, 0);
    for (int k = (int)threadIdx.x; k < total; k += 256) {
        const int ip = k / nt, it = k - ip * nt;
        float sp, cp, st, ct;
        sincosf(s_phi[ip], &sp, &cp);
        sincosf(s_theta[it], &st, &ct);
        const V3 d = Right * (st * cp) + Up * (st * sp) + N * ct;
        const V3 c = sample_cube(P.src, d);
        sum = sum + c * (ct * st);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        sum.x += __shfl_xor_sync(0xFFFFFFFFu, sum.x, o);
        sum.y += __shfl_xor_sync(0xFFFFFFFFu, sum.y, o);
        sum.z += __shfl_xor_sync(0xFFFFFFFFu, sum.z, o);
    }
    if ((threadIdx.x & 31) == 0) { s_red[threadIdx.x >> 5][0] = sum.x; s_red[threadIdx.x >> 5][1] = sum.y; s_red[threadIdx.x >> 5][2] = sum.z; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float r = 0, g = 0, b = 0;
        for (int w = 0; w < 8; w++) { r += s_red[w][0]; g += s_red[w][1]; b += s_red[w][2]; }
        const float k = kPI / (float)total;
        P.dst[tix] = make_float4(r * k, g * k, b * k, 1.0f);
    }
}

} // namespace

uint32_t envprobe_texels(int max_lod)
{
    uint32_t n = 0;
    for (int s = 1 << max_lod; s >= 1; s >>= 1) n += 6u * (uint32_t)s * (uint32_t)s;
    return n;
}

cudaError_t launch_envprobe_prefilter(const float* d_cube, int src_size, int max_lod, float* d_dst, cudaStream_t st)
{
    ProbeParams P;
    P.src.texels = (const float4*)d_cube;
    P.src.n = src_size;
    P.dst = (float4*)d_dst;
    P.max_lod = max_lod;
    P.total = envprobe_texels(max_lod);
    envprobe_prefilter_kernel<<<(P.total + 7) / 8, 256, 0, st>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_irradiance(const float* d_cube, int src_size, int out_size, float* d_dst, cudaStream_t st)
{
    IrrParams P;
    P.src.texels = (const float4*)d_cube;
    P.src.n = src_size;
    P.dst = (float4*)d_dst;
    P.size = out_size;
    irradiance_kernel<<<6 * out_size * out_size, 256, 0, st>>>(P);
    return cudaGetLastError();
}

} // namespace hkbcn
