/* TEST INFRASTRUCTURE -- CPU restatement of the reference's environment-probe shaders.  Never linked
 * into the product; only tests/ and bench.py's cpu_baseline leg load it (through liboracle.so).
 *
 * Parity unpinned: the reference runs this arithmetic only as GLSL through OpenGL (geometry-shader layered rendering with
 * hardware seamless cube filtering), which is unavailable on this box, and it ships no images or vectors for it.  The
 * functions below are the shader bodies written out in C, statement by statement, in single precision and in the shaders'
 * own loop order:
 *   Hammersley, ImportanceSampleGGX, PrefilterEnvMap   Hork/Embedded/Shaders/gen/envprobegen.frag:44-111
 *   CalcIrradiance                                     Hork/Embedded/Shaders/gen/irradiancegen.frag:46-76
 *   level sizes / roughness per level                  Hork/RenderUtils/EnvProbeGenerator.cpp:182-240 (size 2^maxLod, Lod / maxLod)
 *   face frames                                        Float4x4::sGetCubeFaceMatrices, Hork/Math/VectorMath.h:3397-3408, under
 *                                                      the projection of EnvProbeGenerator.cpp:53-65 (worked out in DESIGN.md 8d)
 * textureLod(Envmap, d, 0) is a bilinear fetch whose out-of-face taps are taken from the texel their own direction lands
 * on (seamless edges); GL leaves the filter's precision to the implementation. */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

typedef struct { float x, y, z; } v3;
static v3 mk(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static v3 add(v3 a, v3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 sub(v3 a, v3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 mul(v3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
static float dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static v3 cross3(v3 a, v3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static v3 norm3(v3 a) { return mul(a, 1.0f / sqrtf(dot3(a, a))); }

static v3 face_dir(int f, float u, float v)
{
    switch (f) {
    case 0: return mk(1.0f, -v, -u);
    case 1: return mk(-1.0f, -v, u);
    case 2: return mk(u, 1.0f, v);
    case 3: return mk(u, -1.0f, -v);
    case 4: return mk(u, -v, 1.0f);
    default: return mk(-u, -v, -1.0f);
    }
}
static int dir_face(v3 d, float* u, float* v)
{
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    if (ax >= ay && ax >= az) {
        const float m = 1.0f / ax;
        if (d.x > 0) { *u = -d.z * m; *v = -d.y * m; return 0; }
        *u = d.z * m; *v = -d.y * m; return 1;
    }
    if (ay >= az) {
        const float m = 1.0f / ay;
        if (d.y > 0) { *u = d.x * m; *v = d.z * m; return 2; }
        *u = d.x * m; *v = -d.z * m; return 3;
    }
    {
        const float m = 1.0f / az;
        if (d.z > 0) { *u = d.x * m; *v = -d.y * m; return 4; }
        *u = -d.x * m; *v = -d.y * m; return 5;
    }
}

typedef struct { const float* t; int n; } cube;
static const float* texel(const cube* c, int f, int x, int y) { return c->t + (((size_t)f * c->n + y) * c->n + x) * 4; }
static const float* tap(const cube* c, int f, int x, int y)
{
    if (x >= 0 && x < c->n && y >= 0 && y < c->n) return texel(c, f, x, y);
    const float inv = 2.0f / (float)c->n;
    float u, v;
    const int g = dir_face(face_dir(f, ((float)x + 0.5f) * inv - 1.0f, ((float)y + 0.5f) * inv - 1.0f), &u, &v);
    int gx = (int)floorf((u * 0.5f + 0.5f) * (float)c->n), gy = (int)floorf((v * 0.5f + 0.5f) * (float)c->n);
    gx = gx < 0 ? 0 : gx > c->n - 1 ? c->n - 1 : gx;
    gy = gy < 0 ? 0 : gy > c->n - 1 ? c->n - 1 : gy;
    return texel(c, g, gx, gy);
}
static v3 sample_cube(const cube* c, v3 d)
{
    float u, v;
    const int f = dir_face(d, &u, &v);
    const float x = (u * 0.5f + 0.5f) * (float)c->n - 0.5f, y = (v * 0.5f + 0.5f) * (float)c->n - 0.5f;
    const float x0f = floorf(x), y0f = floorf(y);
    const float fx = x - x0f, fy = y - y0f;
    const int x0 = (int)x0f, y0 = (int)y0f;
    const float *a = tap(c, f, x0, y0), *b = tap(c, f, x0 + 1, y0), *p = tap(c, f, x0, y0 + 1), *q = tap(c, f, x0 + 1, y0 + 1);
    const float w00 = (1.0f - fx) * (1.0f - fy), w10 = fx * (1.0f - fy), w01 = (1.0f - fx) * fy, w11 = fx * fy;
    return mk(a[0] * w00 + b[0] * w10 + p[0] * w01 + q[0] * w11, a[1] * w00 + b[1] * w10 + p[1] * w01 + q[1] * w11,
              a[2] * w00 + b[2] * w10 + p[2] * w01 + q[2] * w11);
}

static const float PI = 3.1415926f;

static void hammersley(int k, int n, float* x, float* y)
{
    float u = 0, p = 0.5f;
    for (int kk = k; kk; p *= 0.5f, kk /= 2)
        if (kk & 1) u += p;
    *x = u;
    *y = (k + 0.5f) / n;
}
static v3 importance_sample_ggx(float xi_x, float xi_y, float Roughness, v3 N)
{
    const float a = Roughness * Roughness;
    const float Phi = 2 * PI * xi_x;
    const float CosTheta = sqrtf((1 - xi_y) / (1 + (a * a - 1) * xi_y));
    const float SinTheta = sqrtf(1 - CosTheta * CosTheta);
    const v3 H = mk(SinTheta * cosf(Phi), SinTheta * sinf(Phi), CosTheta);
    const v3 UpVector = fabsf(N.z) < 0.99f ? mk(0, 0, 1) : mk(1, 0, 0);
    const v3 Tangent = norm3(cross3(UpVector, N));
    const v3 Bitangent = cross3(N, Tangent);
    return norm3(add(add(mul(Tangent, H.x), mul(Bitangent, H.y)), mul(N, H.z)));
}
static v3 prefilter_env_map(const cube* c, float Roughness, v3 R)
{
    v3 PrefilteredColor = mk(0, 0, 0);
    float TotalWeight = 0.0f;
    const int NumSamples = 1024;
    for (int i = 0; i < NumSamples; i++) {
        float xx, xy;
        hammersley(i, NumSamples, &xx, &xy);
        const v3 H = importance_sample_ggx(xx, xy, Roughness, R);
        const v3 L = norm3(sub(mul(H, 2 * dot3(R, H)), R));
        float NdL = dot3(R, L);
        NdL = NdL < 0.0f ? 0.0f : NdL > 1.0f ? 1.0f : NdL;
        if (NdL > 0) {
            PrefilteredColor = add(PrefilteredColor, mul(sample_cube(c, L), NdL));
            TotalWeight += NdL;
        }
    }
    return mk(PrefilteredColor.x / TotalWeight, PrefilteredColor.y / TotalWeight, PrefilteredColor.z / TotalWeight);
}

/* dst: levels back to back, each [6][size][size] RGBA32F, size = 2^max_lod >> lod; returns texels written */
size_t oracle_envprobe_prefilter(const float* cube_rgba, int src_size, int max_lod, float* dst)
{
    const cube c = {cube_rgba, src_size};
    size_t o = 0;
    int lod = 0;
    for (int size = 1 << max_lod; size >= 1; size >>= 1, lod++) {
        const float roughness = (float)lod / (float)max_lod;
        const float inv = 2.0f / (float)size;
        for (int f = 0; f < 6; f++)
            for (int y = 0; y < size; y++)
                for (int x = 0; x < size; x++) {
                    const v3 R = norm3(face_dir(f, ((float)x + 0.5f) * inv - 1.0f, ((float)y + 0.5f) * inv - 1.0f));
                    const v3 col = prefilter_env_map(&c, roughness, R);
                    dst[4 * o] = col.x; dst[4 * o + 1] = col.y; dst[4 * o + 2] = col.z; dst[4 * o + 3] = 1.0f;
                    o++;
                }
    }
    return o;
}

void oracle_irradiance(const float* cube_rgba, int src_size, int out_size, float* dst)
{
    const cube c = {cube_rgba, src_size};
    const float _2PI = 2.0f * PI, HalfPI = 0.5f * PI, SampleDelta = 0.025f;
    const float inv = 2.0f / (float)out_size;
    size_t o = 0;
    for (int f = 0; f < 6; f++)
        for (int y = 0; y < out_size; y++)
            for (int x = 0; x < out_size; x++) {
                const v3 Normal = norm3(face_dir(f, ((float)x + 0.5f) * inv - 1.0f, ((float)y + 0.5f) * inv - 1.0f));
                v3 Irradiance = mk(0, 0, 0);
                v3 Up = mk(0.0f, 1.0f, 0.0f);
                const v3 Right = cross3(Up, Normal);
                Up = cross3(Normal, Right);
                float NumSamples = 0.0f;
                for (float Phi = 0.0f; Phi < _2PI; Phi += SampleDelta) {
                    const float sp = sinf(Phi), cp = cosf(Phi);
                    for (float Theta = 0.0f; Theta < HalfPI; Theta += SampleDelta) {
                        const float st = sinf(Theta), ct = cosf(Theta);
                        const v3 sampleVec = add(add(mul(Right, st * cp), mul(Up, st * sp)), mul(Normal, ct));
                        Irradiance = add(Irradiance, mul(sample_cube(&c, sampleVec), ct * st));
                        NumSamples += 1.0f;
                    }
                }
                Irradiance = mul(Irradiance, PI / NumSamples);
                dst[4 * o] = Irradiance.x; dst[4 * o + 1] = Irradiance.y; dst[4 * o + 2] = Irradiance.z; dst[4 * o + 3] = 1.0f;
                o++;
            }
}
