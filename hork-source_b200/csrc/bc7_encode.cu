// bc7_encode.cu -- BC7 surface encoder (sm_100a).
//
// Replaces TextureBlockCompression::CompressBC7 (Hork/Image/ImageEncoders.cpp:524-557) ->
// Encode_BC7(.., BC7_ENCODE_MAX_LEVEL = 4) (:207-212) -> bc7enc_compress_block with
// compressionParams.bc7_params[4] (:59-63; ThirdParty/bc7enc_rdo/bc7enc.cpp:2402-2438).
// See bc7_thread.cuh for the per-block algorithm and the parity contract.
//
// Kernel shape: persistent 512-thread CTAs, one per SM (128 registers per thread), tiles claimed from the launch's counter
// (tile cost varies several-fold with content).  A CTA stages a tile of 512 consecutive blocks (32 KB of RGBA8) into
// shared memory with TMA bulk copies, sorts the tile's blocks into opaque ones and ones with alpha (the two classes run
// different mode sets, bc7enc.cpp:2428-2437; sorted so that at most one warp per tile holds both), and every thread then
// encodes one block from a transposed copy of its 16 texels; the warps of a CTA are re-aligned between search stages so
// that they fetch the same instructions.
// Algorithmic HBM bytes: 5 B/px; the kernel is ALU-bound by about three orders of magnitude.
#include <atomic>
#include "hkbcn_common.cuh"
#include "hkbcn_kernels.h"
#include "bc7_thread.cuh"
#include "bc7_host_tables.h"

namespace hkbcn {

__device__ bc7::Tables g_bc7_tables;

#ifndef BC7_DYNAMIC
#define BC7_DYNAMIC 1
#endif
#ifndef BC7_TILE_BLOCKS
#define BC7_TILE_BLOCKS 640
#endif
#ifndef BC7_PHASE_SYNC
#define BC7_PHASE_SYNC 1
#endif
#ifndef BC7_THREADS
#define BC7_THREADS 640
#endif
constexpr int BC7_NT = BC7_THREADS;
constexpr int BC7_OCC = BC7_NT >= 512 ? 1 : 512 / BC7_NT; // 128 registers per thread at 512 threads
constexpr int BC7_TB = BC7_TILE_BLOCKS;
constexpr size_t BC7_SMEM = 4 * BC7_TB * 16 + 64 * BC7_NT + sizeof(bc7::Tables); // stage, transposed texels, tables

__global__ void __launch_bounds__(BC7_NT, BC7_OCC) bc7_encode_kernel(const __grid_constant__ EncodeParams P)
{
    constexpr int ROWB = BC7_TB * 16;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const stage = smem;                                   // 4 * ROWB
    uint32_t* const tpx = (uint32_t*)(smem + 4 * ROWB);            // 16 * BC7_NT words
    bc7::Tables& T = *(bc7::Tables*)(smem + 4 * ROWB + 64 * BC7_NT);
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint16_t order[BC7_TB];
    __shared__ uint32_t count[2];
    __shared__ uint32_t nextTile;

    {
        const uint32_t* src = (const uint32_t*)&g_bc7_tables;
        uint32_t* dst = (uint32_t*)&T;
        for (uint32_t i = threadIdx.x; i < sizeof(bc7::Tables) / 4; i += BC7_NT) dst[i] = src[i];
    }
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    uint32_t phase = 0;
    // HKBCN_FLAG_LEVEL: uber level + 1 in bits 4..8 of the flags (0 = the CompressBC7 preset, level 4)
    const int level = ((P.flags >> 4) & 31u) ? min(4, (int)((P.flags >> 4) & 31u) - 1) : 4;
    bc7::Block blk;
    blk.px = tpx + threadIdx.x;
    blk.stride = BC7_NT;

    uint32_t tile = blockIdx.x;
    while (tile < P.totalTiles) {
        const Level& L = P.lv[find_level(P, tile)];
        const uint32_t b0 = (tile - L.tileBase) * BC7_TB;
        const uint32_t nb = min((uint32_t)BC7_TB, L.nblocks - b0);
        if (threadIdx.x < 2) count[threadIdx.x] = 0;
        stage_tile<4, BC7_TB, BC7_NT>(L, b0, nb, stage, &bar, phase);
        __syncthreads();

        // class 1 = some texel has alpha < 255 (bc7enc.cpp:2428-2436); count, then place opaque blocks first
        uint32_t cls[BC7_TB / BC7_NT], slot[BC7_TB / BC7_NT];
#pragma unroll
        for (int q = 0; q < BC7_TB / BC7_NT; q++) {
            const uint32_t k = threadIdx.x + q * BC7_NT;
            cls[q] = 2;
            if (k < nb) {
                uint32_t m = 0xFFFFFFFFu;
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    const uint4 v = *(const uint4*)(stage + r * ROWB + k * 16);
                    m &= v.x & v.y & v.z & v.w;
                }
                cls[q] = (m >> 24) != 255u;
                slot[q] = atomicAdd(&count[cls[q]], 1u);
            }
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < BC7_TB / BC7_NT; q++)
            if (cls[q] < 2) order[slot[q] + (cls[q] ? count[0] : 0)] = (uint16_t)(threadIdx.x + q * BC7_NT);
        __syncthreads();

#if BC7_PHASE_SYNC
#pragma unroll 1
        for (uint32_t base = 0; base < nb; base += BC7_NT) {
            const uint32_t item = base + threadIdx.x;
            const int c = item < nb ? (item >= count[0] ? 1 : 0) : 2;
            const uint32_t k = c < 2 ? order[item] : 0;
            if (c < 2) {
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    const uint4 v = *(const uint4*)(stage + r * ROWB + k * 16);
                    tpx[(4 * r + 0) * BC7_NT + threadIdx.x] = v.x;
                    tpx[(4 * r + 1) * BC7_NT + threadIdx.x] = v.y;
                    tpx[(4 * r + 2) * BC7_NT + threadIdx.x] = v.z;
                    tpx[(4 * r + 3) * BC7_NT + threadIdx.x] = v.w;
                }
            }
            const uint4 out = bc7::encode_phased(blk, T, c, level);
            if (c < 2) *(uint4*)(L.dst + (size_t)(b0 + k) * 16) = out;
        }
#else
#pragma unroll 1
        for (uint32_t item = threadIdx.x; item < nb; item += BC7_NT) {
            const uint32_t k = order[item];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint4 v = *(const uint4*)(stage + r * ROWB + k * 16);
                tpx[(4 * r + 0) * BC7_NT + threadIdx.x] = v.x;
                tpx[(4 * r + 1) * BC7_NT + threadIdx.x] = v.y;
                tpx[(4 * r + 2) * BC7_NT + threadIdx.x] = v.z;
                tpx[(4 * r + 3) * BC7_NT + threadIdx.x] = v.w;
            }
            const uint4 out = (item >= count[0]) ? bc7::encode_alpha(blk, T, level) : bc7::encode_opaque(blk, T, level);
            *(uint4*)(L.dst + (size_t)(b0 + k) * 16) = out;
        }
#endif
        // tile cost varies several-fold with content (opaque vs alpha, flat vs busy), so tiles are claimed, not strided
        if (threadIdx.x == 0) nextTile = BC7_DYNAMIC ? gridDim.x + atomicAdd(P.sched, 1u) : tile + gridDim.x;
        __syncthreads(); // stage / order are reused by the next tile
        tile = nextTile;
    }
    if (BC7_DYNAMIC && threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(P.sched + 1, 1u) == gridDim.x - 1u) {
            P.sched[0] = 0;
            P.sched[1] = 0;
        }
    }
}

uint32_t bc7_tile_blocks() { return BC7_TB; }

cudaError_t launch_bc7_encode(const EncodeParams& P, int sm_count, cudaStream_t stream)
{
    if (P.totalTiles == 0) return cudaSuccess;
    const uint32_t g = (uint32_t)sm_count * BC7_OCC;
    const int grid = (int)(P.totalTiles < g ? P.totalTiles : g);
    // function attributes live in the device's context: once per DEVICE, not per process (a second GPU launching with the
    // first one's flag set fails with "invalid argument").  Same value every time, so a race only repeats the call.
    static std::atomic<uint64_t> done{0};
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
        cudaError_t e = cudaFuncSetAttribute(bc7_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BC7_SMEM);
        if (e != cudaSuccess) return e;
        done.fetch_or(bit, std::memory_order_release);
    }
    bc7_encode_kernel<<<grid, BC7_NT, BC7_SMEM, stream>>>(P);
    return cudaGetLastError();
}

// ---- a12: lookup tables (bc7_host_tables.h) ----------------------------------------------------
cudaError_t bc7_upload_tables()
{
    static bc7::Tables* H = nullptr;
    if (!H) {
        H = new bc7::Tables();
        if (!bc7::fill_tables(H)) return cudaErrorInvalidValue;
    }
    return cudaMemcpyToSymbol(g_bc7_tables, H, sizeof(bc7::Tables));
}

} // namespace hkbcn
