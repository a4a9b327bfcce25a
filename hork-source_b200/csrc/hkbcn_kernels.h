// hkbcn_kernels.h -- internal launcher interface between the C-ABI layer and the .cu files.
#pragma once
#include "hkbcn_common.cuh"

namespace hkbcn {

// codec ids used internally (format family, independent of sRGB/SNORM tagging)
enum Codec { C_BC1 = 0, C_BC2, C_BC3, C_BC4, C_BC5, C_BC6H, C_BC7 };

// ---- encode -----------------------------------------------------------------
uint32_t bc45_tile_blocks(int channels);
cudaError_t launch_bc45_encode(int channels, const EncodeParams& P, int sm_count, cudaStream_t stream);

uint32_t bc123_tile_blocks(int codec);
cudaError_t launch_bc123_encode(int codec, const EncodeParams& P, int sm_count, cudaStream_t stream);
cudaError_t bc1_upload_tables(); // a12: single-colour match + 565 midpoint tables; the orderings tables of levels 6..18
bool bc1_hq_available();           // the library was built with rgbcx's total-ordering tables (build/bc1_orderings.inc)
bool bc123_needs_hq(int codec, uint32_t flags); // this level / option set takes the error-tracking branch (bc1_hq.cuh)

uint32_t bc7_tile_blocks();
// P.aux: bc7_aux_bytes(blocks of the launch) bytes of device scratch whose first 16 bytes are zero (the kernels leave them zero)
size_t bc7_aux_bytes(uint32_t total_blocks);
cudaError_t launch_bc7_encode(const EncodeParams& P, int sm_count, cudaStream_t stream);
cudaError_t bc7_upload_tables();

uint32_t bc6h_tile_blocks();
cudaError_t launch_bc6h_encode(bool is_signed, bool src_f16, const EncodeParams& P, int sm_count, cudaStream_t stream);
cudaError_t bc6h_upload_tables(); // header bit layouts of the 14 modes

// ---- decode -----------------------------------------------------------------
struct DecodeParams {
    const uint8_t* src;       // blocks
    uint8_t*       dst;       // texels
    uint32_t       nbx;       // blocks per row
    uint32_t       nblocks;
    size_t         dstRowStride; // bytes
    int            codec;
    uint32_t       is_signed; // BC6H
    uint32_t       f32out;    // BC6H
    uint32_t*      sched;     // two zeroed words of device memory: tile counter and finished-CTA counter (BC6H / BC7 kernels)
};
cudaError_t launch_decode(const DecodeParams& P, int sm_count, cudaStream_t stream);
cudaError_t decode_upload_tables(); // BC6H header layout table

// ---- mip cascade (mip_resize.cu) ---------------------------------------------
struct ResizeParams {
    const uint8_t* src;
    uint8_t*       dst;
    const int*     ix;  // taps of the x axis: output column i reads ix[i*Tx + k] with weight wx[i*Tx + k]
    const float*   wx;
    const int*     iy;
    const float*   wy;
    const float*   lut; // 256 floats: decode of a uint8 colour channel (sRGB table or v / 255)
    int            in_w, in_h, out_w, out_h, Tx, Ty;
    int            alpha;  // alpha channel index (decoded / encoded linearly) or -1
    int            weight; // 1: colour is weighted by alpha before filtering and un-weighted after (not premultiplied input)
    int            srgb;
    // the same taps as runs of consecutive inputs (the separable tile kernel): output i reads the cnt[i] inputs
    // first[i] .. first[i] + cnt[i] - 1 BEFORE edge mapping, input n lives at map[n + margin] (any texel where the weight is 0)
    const int*     firstx; const int* cntx; const int* mapx; int marginx;
    const int*     firsty; const int* cnty; const int* mapy; int marginy;
    int            tileOk; // every 32 x 8 output tile needs at most 72 x 24 inputs: the tile kernel applies
    // both axes are exact 2 : 1 reductions whose taps do not depend on the output (power-of-two chains: every output reads the
    // Tx inputs 2 * i + x0 .. with the same weights ux[]): the register-blocked tile kernel applies
    int            uniform, x0, y0;
    float          ux[8], uy[8];
    // every weight twice (the 64-bit operand of a packed f32x2 instruction) and a +0.0 pair that ptxas cannot fold
    float          ux2[16], uy2[16], zero2[2];
};
cudaError_t launch_mip_resize(const ResizeParams& P, int channels, bool f32, cudaStream_t stream);
bool mip_srgb_is_stbir(); // sRGB channels are encoded with stbir's own segment table (build/stbir_srgb.inc): *_SRGB chains are bit-identical
// any size to any size, 1..4 channels, alpha weighting for float images too (ResampleRawImage): one thread per output texel
cudaError_t launch_resample_general(const ResizeParams& P, int channels, bool f32, cudaStream_t stream);

// ---- environment probes (envprobe.cu) ----------------------------------------
uint32_t envprobe_texels(int max_lod); // 6 * sum of (2^max_lod >> lod)^2
cudaError_t launch_envprobe_prefilter(const float* d_cube, int src_size, int max_lod, float* d_dst, cudaStream_t st);
cudaError_t launch_irradiance(const float* d_cube, int src_size, int out_size, float* d_dst, cudaStream_t st);

// ---- normal-map packers and roughness bake (normal_pack.cu) ------------------
enum { NP_RGBA_BC1 = 0, NP_RG_BC5 = 1, NP_SPHEREMAP_BC5 = 2, NP_STEREOGRAPHIC_BC5 = 3, NP_PARABOLOID_BC5 = 4, NP_RGBA_BC3 = 5 };
cudaError_t launch_pack_normals(int variant, const float* d_normals, uint8_t* d_dst, size_t count, cudaStream_t st);
cudaError_t launch_unpack_normals(int variant, const uint8_t* d_src, float* d_normals, size_t count, cudaStream_t st);
cudaError_t launch_normalize_vectors(float* d_v, size_t count, cudaStream_t st);
cudaError_t launch_roughness_bake(int bake, const float* d_avg_normals, uint8_t* d_rough, size_t count, cudaStream_t st);
cudaError_t launch_invert_green(float* d_xyz, size_t count, cudaStream_t st); // RawImage::InvertGreen on Float3 texels
// RGB -> RGBA (alpha 255 / 1.0) and / or R <-> B: CreateImage's channel pass (Image.cpp:1152-1228)
// PBRImporter's CreateORMX gather (Tools/PBRImporter/Main.cpp:86-190) and one channel of an RGBA8 image <-> a planar R8 image
cudaError_t launch_ormx_gather(const void* const d_maps[4], const int channel[4], void* d_dst, size_t count, cudaStream_t st);
cudaError_t launch_channel_move(void* d_rgba, void* d_plane, size_t count, int channel, bool insert, cudaStream_t st);
cudaError_t launch_expand_channels(const void* d_src, void* d_dst, size_t count, int src_channels, bool swap_rb, bool f32, cudaStream_t st);

} // namespace hkbcn

#include <vector>
namespace hkbcn {
struct AxisTaps {
    std::vector<int>   idx, first, cnt, map;
    std::vector<float> w;
    int                T = 0, margin = 0;
    bool               uniform = false; // cnt[i] == T, first[i] == 2 * i + first[0], identical weights for every output
};
void build_axis_taps(int in_size, int out_size, int filter, int edge, AxisTaps& t);
} // namespace hkbcn
