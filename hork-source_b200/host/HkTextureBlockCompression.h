// HkTextureBlockCompression.h -- header-compatible replacement for the TextureBlockCompression part of
// Hork/Image/ImageEncoders.h (reference: Hork/Image/ImageEncoders.h:37-87).
//
// Same namespace, names, argument order and meaning as the reference, so Image.cpp (:1718-1751 takes these
// as `void (*)(void const*, void*, uint32_t, uint32_t)`), Tools/PBRImporter/Main.cpp:721-727 and
// RenderUtils/Utilites.cpp:256,263 compile and link unchanged.  Every function forwards to the C ABI
// (include/hkbcn.h, libhkbcn.so); there is no CPU implementation behind it.  The reference functions return
// void and cannot fail; this path can (no device, out of memory): a failure is logged to stderr and the
// process aborts, the closest equivalent of the reference's "cannot happen".
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifndef HK_NAMESPACE_BEGIN // inside the engine these come from Hork/Core/BaseTypes.h
#define HK_NAMESPACE_BEGIN namespace Hk {
#define HK_NAMESPACE_END }
#endif

HK_NAMESPACE_BEGIN

namespace TextureBlockCompression
{

const uint32_t BC1_ENCODE_MAX_LEVEL = 18;
const uint32_t BC2_ENCODE_MAX_LEVEL = 18;
const uint32_t BC3_ENCODE_MAX_LEVEL = 18;
const uint32_t BC7_ENCODE_MAX_LEVEL = 4;

// per 4x4 block (ImageEncoders.cpp:77-136): pDest receives 4 rows of 4 texels at RowStride bytes per row
void Decode_BC1(void const* pSrc, void* pDest, size_t RowStride);
void Decode_BC2(void const* pSrc, void* pDest, size_t RowStride);
void Decode_BC3(void const* pSrc, void* pDest, size_t RowStride);
void Decode_BC4(void const* pSrc, void* pDest, size_t RowStride);
void Decode_BC5(void const* pSrc, void* pDest, size_t RowStride);
void Decode_BC6h_f16(void const* pSrc, void* pDest, size_t RowStride, bool bSigned);
void Decode_BC6h_f32(void const* pSrc, void* pDest, size_t RowStride, bool bSigned);
void Decode_BC7(void const* pSrc, void* pDest, size_t RowStride);

// per 4x4 block (ImageEncoders.cpp:138-212).  Built on the device: BC1/2/3 levels 0, 1, 2, 3, 5 (the rgbcx levels without
// error tracking; 5 is what the whole-surface functions bake in), BC3 with and without bMaxQuality, BC4/BC5 hq and non-hq,
// BC7 levels 0..4, BC6H signed / unsigned from f16 / f32.  BC1/2/3 levels 4 and 6..18 and the 3-colour flags of Encode_BC1
// abort with a message rather than silently encoding something else.
void Encode_BC1(void const* pSrc, void* pDest, uint32_t Level, bool b3ColorMode, bool bTransparentPixelsForBlack);
void Encode_BC2(void const* pSrc, void* pDest, uint32_t Level);
void Encode_BC3(void const* pSrc, void* pDest, uint32_t Level, bool bMaxQuality);
void Encode_BC4(void const* pSrc, void* pDest, bool bMaxQuality);
void Encode_BC5(void const* pSrc, void* pDest, bool bMaxQuality);
void Encode_BC6h_f16(void const* pSrc, void* pDest, bool bSigned);
void Encode_BC6h_f32(void const* pSrc, void* pDest, bool bSigned);
void Encode_BC7(void const* pSrc, void* pDest, uint32_t Level);

// whole surface (ImageEncoders.cpp:215-557): tightly packed source, tightly packed blocks out
void CompressBC1(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height);  // RGBA8 in
void CompressBC2(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height);  // RGBA8 in
void CompressBC3(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height);  // RGBA8 in
void CompressBC4(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height);  // R8 in
void CompressBC5(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height);  // RG8 in
void CompressBC6h(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height, bool bSigned); // RGBA32_FLOAT in
void CompressBC7(void const* pSrc, void* pDest, uint32_t Width, uint32_t Height);  // RGBA8 in

// Additions of this path (no counterpart in the reference, which only decodes per block):
// whole-surface decode, and compression of a pre-built mip chain in one launch (ImageStorage layout, Image.cpp:498-574).
void DecompressSurface(int Format /* Hk::TEXTURE_FORMAT */, void const* pSrc, void* pDest, uint32_t Width, uint32_t Height, size_t RowStride, bool bFloat32 = false);
void CompressChain(int Format /* Hk::TEXTURE_FORMAT */, void const* pSrc, void* pDest, uint32_t Width, uint32_t Height, uint32_t NumMipmaps);

} // namespace TextureBlockCompression

HK_NAMESPACE_END
