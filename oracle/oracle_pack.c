/* TEST INFRASTRUCTURE -- plain-C restatement of the reference's normal-map packers, unpackers, NormalizeVectors and the
 * roughness bake.  Never linked into the product; only tests/ load it (through liboracle.so).
 *
 * Parity unpinned: the reference holds no golden vectors or tests for these functions and they sit behind the engine's core
 * headers (RawImage, Float3), so they cannot be compiled on their own here; each function below follows the cited reference
 * lines statement by statement in single precision without contraction (-ffp-contract=off, Makefile).
 *
 *   packers     Hork/Image/ImageEncoders.cpp:693-796     unpackers  :798-878
 *   NormalizeVectors  Hork/Image/Image.cpp:2251-2260 (TVector3::NormalizeSelf, Hork/Math/VectorMath.h:888-899)
 *   bake        Hork/Image/Image.cpp:2546-2578
 *   Math::Round = std::round (Core/BaseMath.h:681-685), Math::Saturate = SSE max then min (:794-798; NaN -> 0),
 *   Math::Sqrt = v > 0 ? sqrt(v) : 0 (:936-940), Math::RSqrt (:949-962), Math::Lerp (:675-679). */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

static float sat(float v)
{
    v = v > 0.0f ? v : 0.0f; /* _mm_max_ss(val, 0): a NaN gives the second operand */
    return v < 1.0f ? v : 1.0f;
}
static uint8_t un8(float v) { return (uint8_t)roundf(sat(v) * 255.0f); }
static float msqrt(float v) { return v > 0.0f ? sqrtf(v) : 0.0f; }
static float mrsqrt(float value)
{
    const float x2 = value * 0.5f;
    uint32_t i;
    float f;
    memcpy(&i, &value, 4);
    i = 0x5f3759dfu - (i >> 1);
    memcpy(&f, &i, 4);
    f *= 1.5f - x2 * f * f;
    return f;
}

/* variant: 0 RGBA_BC1, 1 RG_BC5, 2 Spheremap, 3 Stereographic, 4 Paraboloid, 5 RGBA_BC3 */
int oracle_pack_normals(int variant, const float* n, uint8_t* d, size_t count)
{
    for (size_t k = 0; k < count; k++) {
        const float x = n[3 * k], y = n[3 * k + 1], z = n[3 * k + 2];
        switch (variant) {
        case 0:
            d[4 * k + 0] = un8(x * 0.5f + 0.5f);
            d[4 * k + 1] = un8(y * 0.5f + 0.5f);
            d[4 * k + 2] = un8(z * 0.5f + 0.5f);
            d[4 * k + 3] = 255;
            break;
        case 1:
            d[2 * k + 0] = un8(x * 0.5f + 0.5f);
            d[2 * k + 1] = un8(y * 0.5f + 0.5f);
            break;
        case 2: {
            const float denom = 1.0f / msqrt(z * 8.0f + 8.0f);
            d[2 * k + 0] = un8(x * denom + 0.5f);
            d[2 * k + 1] = un8(y * denom + 0.5f);
            break;
        }
        case 3:
            d[2 * k + 0] = un8(x / (1.0f + z) * 0.5f + 0.5f);
            d[2 * k + 1] = un8(y / (1.0f + z) * 0.5f + 0.5f);
            break;
        case 4: {
            float nx = x, ny = y;
            const float a = (nx * nx) + (ny * ny), b = z, c = -1.0f;
            const float discriminant = b * b - 4.0f * a * c;
            const float t = (-b + sqrtf(discriminant)) / (2.0f * a);
            nx = nx * t;
            ny = ny * t;
            d[2 * k + 0] = un8(nx * 0.5f + 0.5f);
            d[2 * k + 1] = un8(ny * 0.5f + 0.5f);
            break;
        }
        case 5:
            d[4 * k + 0] = 255;
            d[4 * k + 1] = un8(y * 0.5f + 0.5f);
            d[4 * k + 2] = un8(z * 0.5f + 0.5f);
            d[4 * k + 3] = un8(x * 0.5f + 0.5f);
            break;
        default: return -1;
        }
    }
    return 0;
}

int oracle_unpack_normals(int variant, const uint8_t* s, float* n, size_t count)
{
    for (size_t k = 0; k < count; k++) {
        float x, y, z;
        if (variant == 0) {
            const uint8_t* d = s + 4 * k;
            x = (d[0] / 255.0f) * 2.0f - 1.0f;
            y = (d[1] / 255.0f) * 2.0f - 1.0f;
            z = (d[2] / 255.0f) * 2.0f - 1.0f;
        } else if (variant == 5) {
            const uint8_t* d = s + 4 * k;
            x = d[3] / 255.0f * 2.0f - 1.0f;
            y = d[1] / 255.0f * 2.0f - 1.0f;
            z = d[2] / 255.0f * 2.0f - 1.0f;
        } else {
            const uint8_t* d = s + 2 * k;
            if (variant == 1) {
                x = (d[0] / 255.0f) * 2.0f - 1.0f;
                y = (d[1] / 255.0f) * 2.0f - 1.0f;
                z = msqrt(1.0f - (x * x + y * y));
            } else if (variant == 2) {
                const float px = (d[0] / 255.0f) * 4.0f - 2.0f, py = (d[1] / 255.0f) * 4.0f - 2.0f;
                const float f = px * px + py * py;
                const float sq = msqrt(1.0f - f / 4.0f);
                x = px * sq;
                y = py * sq;
                z = 1.0f - f / 2.0f;
            } else if (variant == 3) {
                x = (d[0] / 255.0f) * 2.0f - 1.0f;
                y = (d[1] / 255.0f) * 2.0f - 1.0f;
                const float denom = 2.0f / (1 + sat(x * x + y * y));
                x *= denom;
                y *= denom;
                z = denom - 1.0f;
            } else if (variant == 4) {
                x = (d[0] / 255.0f) * 2.0f - 1.0f;
                y = (d[1] / 255.0f) * 2.0f - 1.0f;
                z = 1.0f - sat(x * x + y * y);
            } else return -1;
        }
        n[3 * k] = x;
        n[3 * k + 1] = y;
        n[3 * k + 2] = z;
    }
    return 0;
}

void oracle_normalize_vectors(float* v, size_t count)
{
    for (size_t k = 0; k < count; k++) {
        float x = v[3 * k] * 2.0f - 1.0f, y = v[3 * k + 1] * 2.0f - 1.0f, z = v[3 * k + 2] * 2.0f - 1.0f;
        const float len = sqrtf(x * x + y * y + z * z);
        if (len != 0.0f) {
            const float inv = 1.0f / len;
            x *= inv;
            y *= inv;
            z *= inv;
        }
        v[3 * k] = x;
        v[3 * k + 1] = y;
        v[3 * k + 2] = z;
    }
}

/* one mip level of the bake; bake: 1 vMF, 2 Toksvig */
void oracle_bake_roughness_level(int bake, const float* avg, uint8_t* rough, size_t count)
{
    for (size_t i = 0; i < count; i++) {
        const float x = avg[3 * i], y = avg[3 * i + 1], z = avg[3 * i + 2];
        const float r2 = x * x + y * y + z * z;
        if (r2 > 1e-8f && r2 < 1.0f) {
            if (bake == 1) {
                const float variance = 2.0f * mrsqrt(r2) * (1.0f - r2) / (3.0f - r2);
                const float roughnessVal = rough[i] / 255.0f;
                rough[i] = un8(msqrt(roughnessVal * roughnessVal + variance));
            } else {
                const float r = sqrtf(r2);
                const uint8_t q = rough[i] > 1 ? rough[i] : 1;
                const float rr = q / 255.0f;
                const float specPower = 2.0f / (rr * rr) - 2.0f;
                const float ft = r / (specPower + r * (1.0f - specPower));
                rough[i] = un8(sqrtf(2.0f / (ft * specPower + 2.0f)));
            }
        }
    }
}
