// bc6h_encode.cu -- BC6H surface encoder (sm_100a).
//
// Replaces TextureBlockCompression::CompressBC6h (Hork/Image/ImageEncoders.cpp:395-521) ->
// Encode_BC6h_f32 / Encode_BC6h_f16 (:187-205) -> bc6h_enc::EncodeBC6HU / EncodeBC6HS
// (ThirdParty/bc6h_enc/bc6h_enc.h:3650-3660).  See bc6h_thread.cuh for the per-block search and the parity contract.
// Unlike the reference's 16-thread driver (ImageEncoders.cpp:426-427) no tail block is dropped.
//
// Kernel shape: persistent 64-thread CTAs.  A CTA stages a tile of 64 consecutive blocks (16 KB of RGBA32F, 8 KB of
// RGBA16F) into shared memory with TMA bulk copies, converts every texel once to its half bit pattern
// (half_from_float, Hork/Core/half.c:802) and to the encoder's integer domain, stores both transposed
// (component-major, thread-minor: conflict-free), and every thread then encodes one block.
// Algorithmic HBM bytes: 17 B/px (f32 source) / 9 B/px (f16 source); ALU-bound by four to five orders of magnitude.
#include <cuda_fp16.h>
#include "hkbcn_common.cuh"
#include "hkbcn_kernels.h"
#include "bc6h_thread.cuh"
#include "bc6h_layout.h"

namespace hkbcn {

__device__ bc6h::Layout g_bc6h_enc_layout;

#ifndef BC6E_CTAS_PER_SM
#define BC6E_CTAS_PER_SM 7 // resident CTAs per SM (shared memory: 29.5 KB each); ptxas settles on 128 registers by itself --
                           // asking for it through a min-blocks launch bound changed the allocation and cost 7 % (106.7 vs 99.2 ms)
#endif
constexpr int BC6_NT = 64;
constexpr int BC6_TB = 64;

template <bool F16SRC>
__global__ void __launch_bounds__(BC6_NT) bc6h_encode_kernel(const __grid_constant__ EncodeParams P, const uint32_t isSigned)
{
    constexpr int BPP = F16SRC ? 8 : 16;
    constexpr int ROWB = BC6_TB * 4 * BPP;
    // The staged texels and the integer-domain copy of the tile share one buffer: every thread converts its block into
    // registers, a barrier retires the stage, then the integers are written over it.  30 KB per CTA instead of 42 KB, so
    // seven CTAs (14 warps) fit an SM instead of five -- with 65,536 blocks = 2048 warp-tiles per 1024^2 surface that is the
    // difference between one round of tiles and two, the second of them three-quarters empty.
    constexpr int BUFB = 4 * ROWB > 48 * BC6_NT * 4 ? 4 * ROWB : 48 * BC6_NT * 4;
    __shared__ __align__(128) uint8_t buf[BUFB];
    __shared__ __align__(16) bc6h::Layout L6;
    __shared__ float tf[48 * BC6_NT];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t nextTile;
    uint8_t* const stage = buf;
    int* const ti = (int*)buf;

    {
        const uint32_t* src = (const uint32_t*)&g_bc6h_enc_layout;
        uint32_t* dst = (uint32_t*)&L6;
        for (uint32_t i = threadIdx.x; i < sizeof(bc6h::Layout) / 4; i += BC6_NT) dst[i] = src[i];
    }
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    uint32_t phase = 0;
    bc6h::Block blk;
    blk.f = tf + threadIdx.x;
    blk.ip = ti + threadIdx.x;
    blk.stride = BC6_NT;

    uint32_t tile = blockIdx.x;
    while (tile < P.totalTiles) {
        const Level& L = P.lv[find_level(P, tile)];
        const uint32_t b0 = (tile - L.tileBase) * BC6_TB;
        const uint32_t nb = min((uint32_t)BC6_TB, L.nblocks - b0);
        stage_tile<BPP, BC6_TB, BC6_NT>(L, b0, nb, stage, &bar, phase);
        __syncthreads();
        const uint32_t k = threadIdx.x;
        int q[48];
        bool nan_in = false;
        if (k < nb) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int x = 0; x < 4; x++) {
                    float v[3];
                    if (F16SRC) {
                        // Encode_BC6h_f16 widens the 64 halfs with f16tof32 first (ImageEncoders.cpp:187-197)
                        const uint2 h = *(const uint2*)(stage + r * ROWB + k * 32 + x * 8);
                        v[0] = __half2float(__ushort_as_half((unsigned short)(h.x & 0xFFFFu)));
                        v[1] = __half2float(__ushort_as_half((unsigned short)(h.x >> 16)));
                        v[2] = __half2float(__ushort_as_half((unsigned short)(h.y & 0xFFFFu)));
                    } else {
                        const float4 t = *(const float4*)(stage + r * ROWB + k * 64 + x * 16);
                        v[0] = t.x; v[1] = t.y; v[2] = t.z;
                    }
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        const int i = 3 * (4 * r + x) + c;
                        tf[i * BC6_NT + k] = v[c];
                        nan_in = nan_in || v[c] != v[c];
                        q[i] = bc6h::f16_to_int(bc6h::half_from_float_bits(__float_as_uint(v[c])), isSigned != 0);
                    }
                }
        }
        __syncthreads(); // every thread has read its texels: the stage becomes the integer copy
        if (k < nb) {
#pragma unroll
            for (int i = 0; i < 48; i++) ti[i * BC6_NT + k] = q[i];
            blk.nan_sign = nan_in ? 0u : 0x80000000u;
            const uint4 out = bc6h::encode_block(blk, L6, isSigned != 0);
            *(uint4*)(L.dst + (size_t)(b0 + k) * 16) = out;
        }
        // tiles are claimed from the launch's counter, not strided: the search exits early on exact fits, so a tile's
        // cost varies by an order of magnitude with content
        if (threadIdx.x == 0) nextTile = gridDim.x + atomicAdd(P.sched, 1u);
        __syncthreads(); // the buffer is staged into again by the next tile
        tile = nextTile;
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(P.sched + 1, 1u) == gridDim.x - 1u) {
            P.sched[0] = 0;
            P.sched[1] = 0;
        }
    }
}

uint32_t bc6h_tile_blocks() { return BC6_TB; }

cudaError_t launch_bc6h_encode(bool is_signed, bool src_f16, const EncodeParams& P, int sm_count, cudaStream_t stream)
{
    if (P.totalTiles == 0) return cudaSuccess;
    // one tile per CTA iteration; the grid is one resident wave (or the tile count if smaller) and CTAs claim further tiles
    const uint32_t g = (uint32_t)sm_count * BC6E_CTAS_PER_SM;
    const int grid = (int)(P.totalTiles < g ? P.totalTiles : g);
    if (src_f16)
        bc6h_encode_kernel<true><<<grid, BC6_NT, 0, stream>>>(P, is_signed ? 1u : 0u);
    else
        bc6h_encode_kernel<false><<<grid, BC6_NT, 0, stream>>>(P, is_signed ? 1u : 0u);
    return cudaGetLastError();
}

cudaError_t bc6h_upload_tables()
{
    static bc6h::Layout H;
    bc6h_parse_layout(H.tok);
    return cudaMemcpyToSymbol(g_bc6h_enc_layout, &H, sizeof(H));
}

} // namespace hkbcn
