// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// Pins row f4 to the reference's own object code: the normal-map packers / unpackers
// (Hork/Image/ImageEncoders.cpp:692-878), NormalizeVectors (Hork/Image/Image.cpp:2251-2260) and the per-texel roughness
// update of CreateNormalAndRoughness (Image.cpp:2541-2579) are compiled FROM THE REFERENCE TREE: oracle/Makefile cuts those
// line ranges out of the files where they lie into oracle/_ref/gen_*.inc (generated, git-ignored, never committed) and this
// file includes them between the few declarations they need.  Float3, Math::Round / Saturate / Sqrt / RSqrt / Lerp / Max and
// NormalizeSelf / LengthSqr are the reference's own too (Hork/Math/VectorMath.h, Hork/Core/BaseMath.h compile standalone).
// What is restated here is data plumbing only: a RawImage that owns bytes, the settings enum, and the loop variables the
// roughness excerpt reads.
#include <Hork/Math/VectorMath.h>

#include <stdint.h>
#include <vector>

HK_NAMESPACE_BEGIN

enum RAW_IMAGE_FORMAT { RAW_IMAGE_FORMAT_RGBA8, RAW_IMAGE_FORMAT_R8_ALPHA };
struct RawImage {
    std::vector<uint8_t> bytes;
    RawImage() {}
    RawImage(uint32_t w, uint32_t h, RAW_IMAGE_FORMAT f) : bytes((size_t)w * h * (f == RAW_IMAGE_FORMAT_RGBA8 ? 4 : 2)) {}
    void* GetData() { return bytes.data(); }
};

#include "_ref/gen_packers.inc" // ImageEncoders.cpp:692-878

#include "_ref/gen_normalize.inc" // Image.cpp:2251-2260

struct NormalRoughnessImportSettings {
    enum ROUGHNESS_BAKE { ROUGHNESS_BAKE_vMF, ROUGHNESS_BAKE_TOKSVIG }; // Image.h:593-597
    ROUGHNESS_BAKE RoughnessBake;
};
struct PlainSubresource {
    uint8_t* data;
    uint32_t w, h;
    uint32_t GetWidth() const { return w; }
    uint32_t GetHeight() const { return h; }
    void*    GetData() { return data; }
};
struct PlainNormals {
    float* data;
    void*  GetData() { return data; }
};
static void BakeLevel(NormalRoughnessImportSettings const& Settings, PlainNormals averageNormals, PlainSubresource roughness)
{
#include "_ref/gen_bake.inc" // Image.cpp:2541-2579: pixCount and the per-texel vMF / Toksvig update
}

HK_NAMESPACE_END

extern "C" {

// variant: 0 RGBA_BC1, 1 RG_BC5, 2 Spheremap, 3 Stereographic, 4 Paraboloid, 5 RGBA_BC3 (hkbcn_normal_pack)
int ref_pack_normals(int variant, const float* n, uint8_t* dst, uint32_t count)
{
    using namespace Hk;
    RawImage img;
    Float3 const* N = (Float3 const*)n;
    switch (variant) {
    case 0: img = PackNormalsRGBA_BC1_Compatible(N, count, 1); break;
    case 1: img = PackNormalsRG_BC5_Compatible(N, count, 1); break;
    case 2: img = PackNormalsSpheremap_BC5_Compatible(N, count, 1); break;
    case 3: img = PackNormalsStereographic_BC5_Compatible(N, count, 1); break;
    case 4: img = PackNormalsParaboloid_BC5_Compatible(N, count, 1); break;
    case 5: img = PackNormalsRGBA_BC3_Compatible(N, count, 1); break;
    default: return -1;
    }
    memcpy(dst, img.bytes.data(), img.bytes.size());
    return 0;
}

int ref_unpack_normals(int variant, const uint8_t* src, float* out, uint32_t count)
{
    using namespace Hk;
    for (uint32_t i = 0; i < count; i++) {
        Float3 v;
        switch (variant) {
        case 0: v = UnpackNormalRGBA_BC1_Compatible(src, i); break;
        case 1: v = UnpackNormalRG_BC5_Compatible(src, i); break;
        case 2: v = UnpackNormalSpheremap_BC5_Compatible(src, i); break;
        case 3: v = UnpackNormalStereographic_BC5_Compatible(src, i); break;
        case 4: v = UnpackNormalParaboloid_BC5_Compatible(src, i); break;
        case 5: v = UnpackNormalRGBA_BC3_Compatible(src, i); break;
        default: return -1;
        }
        out[3 * i] = v.X; out[3 * i + 1] = v.Y; out[3 * i + 2] = v.Z;
    }
    return 0;
}

void ref_normalize_vectors(float* v, size_t count) { Hk::NormalizeVectors((Hk::Float3*)v, count); }

// bake: 1 vMF, 2 Toksvig (HKBCN_ROUGHNESS_BAKE_*); avg_normals: the averaged normals of the level, rough: its roughness texels
void ref_bake_roughness_level(int bake, const float* avg_normals, uint8_t* rough, uint32_t count)
{
    Hk::NormalRoughnessImportSettings S;
    S.RoughnessBake = bake == 1 ? Hk::NormalRoughnessImportSettings::ROUGHNESS_BAKE_vMF : Hk::NormalRoughnessImportSettings::ROUGHNESS_BAKE_TOKSVIG;
    Hk::BakeLevel(S, Hk::PlainNormals{(float*)avg_normals}, Hk::PlainSubresource{rough, count, 1});
}

} // extern "C"
