// HkTextureBlockCompression.cpp -- forwards Hk::TextureBlockCompression (reference: Hork/Image/ImageEncoders.cpp:77-557)
// to the C ABI of libhkbcn.so.  Link this file + -lhkbcn in place of ImageEncoders.cpp's TextureBlockCompression part.
#include "HkTextureBlockCompression.h"
#include "../../include/hkbcn.h"

#include <stdio.h>
#include <stdlib.h>

HK_NAMESPACE_BEGIN

namespace TextureBlockCompression
{

namespace
{
void Check(int rc, const char* what)
{
    if (rc == HKBCN_OK) return;
    fprintf(stderr, "Hk::TextureBlockCompression::%s failed (%d): %s\n", what, rc, hkbcn_last_error());
    abort(); // the reference API has no error channel and this path has no CPU fallback
}
} // namespace

void Decode_BC1(void const* s, void* d, size_t RowStride) { Check(hkbcn_decode_blocks(HKBCN_BC1_UNORM, s, d, 1, RowStride, 0), "Decode_BC1"); }
void Decode_BC2(void const* s, void* d, size_t RowStride) { Check(hkbcn_decode_blocks(HKBCN_BC2_UNORM, s, d, 1, RowStride, 0), "Decode_BC2"); }
void Decode_BC3(void const* s, void* d, size_t RowStride) { Check(hkbcn_decode_blocks(HKBCN_BC3_UNORM, s, d, 1, RowStride, 0), "Decode_BC3"); }
void Decode_BC4(void const* s, void* d, size_t RowStride) { Check(hkbcn_decode_blocks(HKBCN_BC4_UNORM, s, d, 1, RowStride, 0), "Decode_BC4"); }
void Decode_BC5(void const* s, void* d, size_t RowStride) { Check(hkbcn_decode_blocks(HKBCN_BC5_UNORM, s, d, 1, RowStride, 0), "Decode_BC5"); }
void Decode_BC6h_f16(void const* s, void* d, size_t RowStride, bool bSigned)
{
    Check(hkbcn_decode_blocks(bSigned ? HKBCN_BC6H_SFLOAT : HKBCN_BC6H_UFLOAT, s, d, 1, RowStride, 0), "Decode_BC6h_f16");
}
void Decode_BC6h_f32(void const* s, void* d, size_t RowStride, bool bSigned)
{
    Check(hkbcn_decode_blocks(bSigned ? HKBCN_BC6H_SFLOAT : HKBCN_BC6H_UFLOAT, s, d, 1, RowStride, HKBCN_FLAG_DST_F32), "Decode_BC6h_f32");
}
void Decode_BC7(void const* s, void* d, size_t RowStride) { Check(hkbcn_decode_blocks(HKBCN_BC7_UNORM, s, d, 1, RowStride, 0), "Decode_BC7"); }

// Levels 0..18 of rgbcx and both 3-colour options (read at levels >= 5, as in rgbcx::encode_bc1).  A library built without
// rgbcx's total-ordering tables answers HKBCN_ERR_UNSUPPORTED above level 5, which Check() turns into an abort.
void Encode_BC1(void const* s, void* d, uint32_t Level, bool b3ColorMode, bool bTransparentPixelsForBlack)
{
    const uint32_t opts = (b3ColorMode ? (uint32_t)HKBCN_FLAG_BC1_3COLOR : 0u) | (bTransparentPixelsForBlack ? (uint32_t)HKBCN_FLAG_BC1_TRANSPARENT_BLACK : 0u);
    Check(hkbcn_encode_blocks(HKBCN_BC1_UNORM, s, d, 1, HKBCN_FLAG_LEVEL(Level) | opts), "Encode_BC1");
}
void Encode_BC2(void const* s, void* d, uint32_t Level) { Check(hkbcn_encode_blocks(HKBCN_BC2_UNORM, s, d, 1, HKBCN_FLAG_LEVEL(Level)), "Encode_BC2"); }
void Encode_BC3(void const* s, void* d, uint32_t Level, bool bMaxQuality)
{
    Check(hkbcn_encode_blocks(HKBCN_BC3_UNORM, s, d, 1, HKBCN_FLAG_LEVEL(Level) | (bMaxQuality ? 0u : (uint32_t)HKBCN_FLAG_FAST)), "Encode_BC3");
}
void Encode_BC4(void const* s, void* d, bool bMaxQuality) { Check(hkbcn_encode_blocks(HKBCN_BC4_UNORM, s, d, 1, bMaxQuality ? 0 : HKBCN_FLAG_FAST), "Encode_BC4"); }
void Encode_BC5(void const* s, void* d, bool bMaxQuality) { Check(hkbcn_encode_blocks(HKBCN_BC5_UNORM, s, d, 1, bMaxQuality ? 0 : HKBCN_FLAG_FAST), "Encode_BC5"); }
void Encode_BC6h_f16(void const* s, void* d, bool bSigned)
{
    Check(hkbcn_encode_blocks(bSigned ? HKBCN_BC6H_SFLOAT : HKBCN_BC6H_UFLOAT, s, d, 1, HKBCN_FLAG_SRC_F16), "Encode_BC6h_f16");
}
void Encode_BC6h_f32(void const* s, void* d, bool bSigned) { Check(hkbcn_encode_blocks(bSigned ? HKBCN_BC6H_SFLOAT : HKBCN_BC6H_UFLOAT, s, d, 1, 0), "Encode_BC6h_f32"); }
void Encode_BC7(void const* s, void* d, uint32_t Level) { Check(hkbcn_encode_blocks(HKBCN_BC7_UNORM, s, d, 1, HKBCN_FLAG_LEVEL(Level)), "Encode_BC7"); }

void CompressBC1(void const* s, void* d, uint32_t w, uint32_t h) { Check(hkbcn_compress(HKBCN_BC1_UNORM, s, d, w, h, 0), "CompressBC1"); }
void CompressBC2(void const* s, void* d, uint32_t w, uint32_t h) { Check(hkbcn_compress(HKBCN_BC2_UNORM, s, d, w, h, 0), "CompressBC2"); }
void CompressBC3(void const* s, void* d, uint32_t w, uint32_t h) { Check(hkbcn_compress(HKBCN_BC3_UNORM, s, d, w, h, 0), "CompressBC3"); }
void CompressBC4(void const* s, void* d, uint32_t w, uint32_t h) { Check(hkbcn_compress(HKBCN_BC4_UNORM, s, d, w, h, 0), "CompressBC4"); }
void CompressBC5(void const* s, void* d, uint32_t w, uint32_t h) { Check(hkbcn_compress(HKBCN_BC5_UNORM, s, d, w, h, 0), "CompressBC5"); }
void CompressBC6h(void const* s, void* d, uint32_t w, uint32_t h, bool bSigned)
{
    Check(hkbcn_compress(bSigned ? HKBCN_BC6H_SFLOAT : HKBCN_BC6H_UFLOAT, s, d, w, h, 0), "CompressBC6h");
}
void CompressBC7(void const* s, void* d, uint32_t w, uint32_t h) { Check(hkbcn_compress(HKBCN_BC7_UNORM, s, d, w, h, 0), "CompressBC7"); }

void DecompressSurface(int Format, void const* s, void* d, uint32_t w, uint32_t h, size_t RowStride, bool bFloat32)
{
    Check(hkbcn_decompress(Format, s, d, w, h, RowStride, bFloat32 ? HKBCN_FLAG_DST_F32 : 0), "DecompressSurface");
}
void CompressChain(int Format, void const* s, void* d, uint32_t w, uint32_t h, uint32_t NumMipmaps)
{
    Check(hkbcn_compress_chain(Format, s, d, w, h, NumMipmaps, 0), "CompressChain");
}

} // namespace TextureBlockCompression

HK_NAMESPACE_END
