// hkbcn_common.cuh -- shared device plumbing for the BCn encode kernels (sm_100a).
//
// Work decomposition (all encoders): a surface (or a whole mip chain) is cut into
// "tiles" of TB consecutive 4x4 blocks in row-major block order, the order the
// reference's Compress* loops emit them (Hork/Image/ImageEncoders.cpp:228-246).
// One CTA iteration stages the texels of one tile into shared memory -- with the
// TMA bulk-copy engine (cp.async.bulk global->shared + mbarrier complete_tx) when
// the row segments are 16-byte aligned, else with plain loads -- and its warps then
// encode the tile's blocks from shared memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define HKBCN_MAX_LEVELS 16

namespace hkbcn {

// One surface (= one mip level of one slice) inside a launch.
struct Level {
    const uint8_t* src;      // device, tightly packed texels (rowStride = width * bpp)
    uint8_t*       dst;      // device, tightly packed blocks
    uint32_t       nbx;      // blocks per row  (Width / 4)
    uint32_t       nblocks;  // nbx * nby
    uint32_t       rowStride;// bytes
    uint32_t       tileBase; // index of this level's first tile in the launch
    uint32_t       bulkOk;   // 1: src base and row segments are 16-B aligned -> TMA bulk path
    uint32_t       pad;
};

struct EncodeParams {
    Level    lv[HKBCN_MAX_LEVELS];
    uint32_t nLevels;
    uint32_t totalTiles;
    uint32_t flags;
    uint32_t pad;
    uint32_t* sched; // {next tile, finished CTAs} of this launch (persistent encoders); both zero between launches
    uint32_t* aux;   // BC7: {blocks with alpha, opaque blocks, 2 unused} + one work-list entry per block of the launch (bc7_encode.cu)
};

// ---- mbarrier / bulk-copy (TMA 1-D) wrappers --------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy, completion counted on `bar` (SASS: UBLKCP).
// dst, src 16-B aligned; bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Find the level a tile belongs to (levels are few; linear scan over uniform data).
__device__ __forceinline__ int find_level(const EncodeParams& P, uint32_t tile)
{
    int l = 0;
#pragma unroll 1
    for (int i = 1; i < (int)P.nLevels; i++)
        if (tile >= P.lv[i].tileBase) l = i;
    return l;
}

// Stage one tile (blocks [b0, b0+nb) of level L) into shared memory.
// Layout: stage[r * ROWB + k * (4*BPP) + ...] = texel row r (0..3) of tile-local block k.
// ROWB = TB * 4 * BPP.  Must be called by all threads of the CTA; returns after the
// data is visible to all threads.  `phase` is the mbarrier phase bit for this use.
template <int BPP, int TB, int NTHREADS>
__device__ __forceinline__ void stage_tile(const Level& L, uint32_t b0, uint32_t nb, uint8_t* stage, uint64_t* bar, uint32_t& phase)
{
    constexpr uint32_t BW = 4 * BPP;          // bytes of one block row
    constexpr uint32_t ROWB = TB * BW;
    if (L.bulkOk) {
        if (threadIdx.x == 0) {
            // order earlier generic-proxy accesses of `stage` before the async-proxy writes
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive_expect_tx(bar, nb * BW * 4);
            uint32_t k = 0;
            while (k < nb) {
                uint32_t b = b0 + k;
                uint32_t by = b / L.nbx, bx = b - by * L.nbx;
                uint32_t run = min(nb - k, L.nbx - bx);
                const uint8_t* g = L.src + (size_t)by * 4 * L.rowStride + (size_t)bx * BW;
#pragma unroll
                for (int r = 0; r < 4; r++) bulk_g2s(stage + r * ROWB + k * BW, g + (size_t)r * L.rowStride, run * BW, bar);
                k += run;
            }
        }
        mbar_wait(bar, phase);
        phase ^= 1;
    } else {
        // unaligned / narrow surfaces (small mips): plain loads, one (block,row) per step
        for (uint32_t i = threadIdx.x; i < nb * 4; i += NTHREADS) {
            uint32_t k = i >> 2, r = i & 3;
            uint32_t b = b0 + k;
            uint32_t by = b / L.nbx, bx = b - by * L.nbx;
            const uint8_t* g = L.src + ((size_t)by * 4 + r) * L.rowStride + (size_t)bx * BW;
            uint8_t* s = stage + r * ROWB + k * BW;
            if constexpr (BW % 4 == 0) {
                // rowStride is a multiple of BW/… only for bpp>=1 with width%4==0 -> 4-B aligned
                if (((size_t)g & 3) == 0) {
#pragma unroll
                    for (uint32_t j = 0; j < BW / 4; j++) ((uint32_t*)s)[j] = ((const uint32_t*)g)[j];
                } else {
#pragma unroll
                    for (uint32_t j = 0; j < BW; j++) s[j] = g[j];
                }
            } else {
#pragma unroll
                for (uint32_t j = 0; j < BW; j++) s[j] = g[j];
            }
        }
        __syncthreads();
    }
}

} // namespace hkbcn
