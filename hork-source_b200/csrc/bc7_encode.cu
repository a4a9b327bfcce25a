// bc7_encode.cu -- BC7 surface encoder (sm_100a).
//
// Replaces TextureBlockCompression::CompressBC7 (Hork/Image/ImageEncoders.cpp:524-557) ->
// Encode_BC7(.., BC7_ENCODE_MAX_LEVEL = 4) (:207-212) -> bc7enc_compress_block with
// compressionParams.bc7_params[4] (:59-63; ThirdParty/bc7enc_rdo/bc7enc.cpp:2402-2438).
// See bc7_thread.cuh for the per-block algorithm and the parity contract.
//
// Kernel shape: persistent 640-thread CTAs, one per SM (96 registers per thread), every thread encodes one block per tile;
// the warps of a CTA are re-aligned between search stages so that they fetch the same instructions (the encoder is ~200 KB
// of SASS against a 32 KB instruction cache).
//
// Blocks with alpha and opaque blocks run different mode sets (bc7enc.cpp:2428-2437: 6 + 5 + 7 against 6 + 1) and an alpha
// block costs about three times an opaque one.  A tile that holds both makes the opaque warps wait for the alpha warps at
// every stage barrier and at the end of the tile (ncu on a texture with a soft alpha mask: 2.8 barrier stall cycles per issue,
// issue slots 55 % busy against 73 % on opaque content).  So the launch is two kernels:
//   bc7_classify_kernel   one thread per block: alpha test of its 16 texels, then a warp-aggregated append of the block's
//                         (level, index) to a work list -- blocks with alpha from the front, opaque blocks from the back;
//   bc7_encode_kernel     tiles of 640 consecutive LIST entries, claimed from the launch's counter front to back: every tile
//                         but the one that holds the boundary is of one class, and the expensive (alpha) tiles go first, so
//                         the tail of the launch is made of cheap ones.  A thread loads the four 16-byte rows of its block
//                         straight from the surface (5 B/px against ~11,000 warp instructions per block: the loads do not
//                         show) into a transposed shared-memory copy, encodes, and stores the block at its final offset.
// The list lives in scratch memory of the launching stream (EncodeParams::aux, 4 bytes per block; hkbcn_api.cu).
// Algorithmic HBM bytes: 5 B/px; the kernel is ALU-bound by about three orders of magnitude.
#include <atomic>
#include "hkbcn_common.cuh"
#include "hkbcn_kernels.h"
#include "bc7_thread.cuh"
#include "bc7_host_tables.h"

namespace hkbcn {

__device__ bc7::Tables g_bc7_tables;

#ifndef BC7_THREADS
#define BC7_THREADS 640
#endif
constexpr int BC7_NT = BC7_THREADS;
#ifdef BC7_CTAS_PER_SM
constexpr int BC7_OCC = BC7_CTAS_PER_SM;
#else
constexpr int BC7_OCC = BC7_NT >= 512 ? 1 : 512 / BC7_NT; // 128 registers per thread at 512 threads
#endif
constexpr int BC7_TB = BC7_NT;                            // one block per thread and tile
constexpr size_t BC7_SMEM = 64 * BC7_NT + sizeof(bc7::Tables); // transposed texels, tables

// first block of the tile-th group of BC7_TB blocks in the launch's (level, block) numbering
__device__ __forceinline__ uint32_t total_blocks(const EncodeParams& P)
{
    uint32_t n = 0;
#pragma unroll 1
    for (int i = 0; i < (int)P.nLevels; i++) n += P.lv[i].nblocks;
    return n;
}

// 16 bytes of one block row from an address of any alignment
__device__ __forceinline__ uint4 load_row16(const uint8_t* g, uint32_t align)
{
    if (align >= 16) return *(const uint4*)g;
    if (align >= 4) {
        const uint32_t* w = (const uint32_t*)g;
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; j++) w[j] = (uint32_t)g[4 * j] | ((uint32_t)g[4 * j + 1] << 8) | ((uint32_t)g[4 * j + 2] << 16) | ((uint32_t)g[4 * j + 3] << 24);
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ uint32_t level_align(const Level& L)
{
    if (L.bulkOk) return 16;
    return ((((size_t)L.src) | L.rowStride) & 3) == 0 ? 4 : 1;
}

// ---- pass 1: the class-sorted work list.  aux[0] = blocks with alpha, aux[1] = opaque blocks (both zero at launch, left
// zero again by the encode kernel), aux[4 + i] = entry i: level << 28 | block index within the level.
// Appends are aggregated per CTA and tile (one atomic per class and 640 blocks: 44 k warp-level atomics on one address took
// longer than reading the surface).
__global__ void __launch_bounds__(BC7_NT) bc7_classify_kernel(const __grid_constant__ EncodeParams P, const uint32_t nTotal)
{
    uint32_t* const cnt = P.aux;
    uint32_t* const list = P.aux + 4;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    __shared__ uint32_t wcnt[2][BC7_NT / 32]; // per warp: blocks with alpha, opaque blocks -> exclusive prefix + the CTA's base
    for (uint32_t tile = blockIdx.x; tile < P.totalTiles; tile += gridDim.x) { // a tile of BC7_TB blocks of one level
        const int li = find_level(P, tile);
        const Level& L = P.lv[li];
        const uint32_t b0 = (tile - L.tileBase) * BC7_TB;
        const uint32_t nb = min((uint32_t)BC7_TB, L.nblocks - b0);
        const uint32_t align = level_align(L);
        const uint32_t k = threadIdx.x;
        const bool valid = k < nb;
        bool alpha = false;
        if (valid) {
            const uint32_t b = b0 + k, by = b / L.nbx, bx = b - by * L.nbx;
            const uint8_t* g = L.src + (size_t)by * 4 * L.rowStride + (size_t)bx * 16;
            uint32_t m = 0xFFFFFFFFu;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint4 v = load_row16(g + (size_t)r * L.rowStride, align);
                m &= v.x & v.y & v.z & v.w;
            }
            alpha = (m >> 24) != 255u; // some texel has alpha < 255 (bc7enc.cpp:2428-2436)
        }
        const uint32_t ma = __ballot_sync(0xFFFFFFFFu, valid && alpha), mo = __ballot_sync(0xFFFFFFFFu, valid && !alpha);
        if (lane == 0) {
            wcnt[0][warp] = (uint32_t)__popc(ma);
            wcnt[1][warp] = (uint32_t)__popc(mo);
        }
        __syncthreads();
        if (threadIdx.x < 2) { // thread c: class c
            uint32_t run = 0;
            for (int w = 0; w < BC7_NT / 32; w++) {
                const uint32_t c = wcnt[threadIdx.x][w];
                wcnt[threadIdx.x][w] = run;
                run += c;
            }
            const uint32_t base = run ? atomicAdd(&cnt[threadIdx.x], run) : 0u;
            for (int w = 0; w < BC7_NT / 32; w++) wcnt[threadIdx.x][w] += base;
        }
        __syncthreads();
        if (valid) {
            const uint32_t below = (1u << lane) - 1u;
            const uint32_t pos = alpha ? wcnt[0][warp] + (uint32_t)__popc(ma & below) : nTotal - 1u - (wcnt[1][warp] + (uint32_t)__popc(mo & below));
            list[pos] = ((uint32_t)li << 28) | (b0 + k);
        }
        __syncthreads(); // wcnt is reused by the next tile
    }
}

// ---- pass 2: the encoder over the list
__global__ void __launch_bounds__(BC7_NT, BC7_OCC) bc7_encode_kernel(const __grid_constant__ EncodeParams P, const uint32_t nTotal)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint32_t* const tpx = (uint32_t*)smem;                          // 16 * BC7_NT words
    bc7::Tables& T = *(bc7::Tables*)(smem + 64 * BC7_NT);
    __shared__ uint32_t nextTile;
    {
        const uint32_t* src = (const uint32_t*)&g_bc7_tables;
        uint32_t* dst = (uint32_t*)&T;
        for (uint32_t i = threadIdx.x; i < sizeof(bc7::Tables) / 4; i += BC7_NT) dst[i] = src[i];
    }
    __syncthreads();
    // HKBCN_FLAG_LEVEL: uber level + 1 in bits 4..8 of the flags (0 = the CompressBC7 preset, level 4)
    const int level = ((P.flags >> 4) & 31u) ? min(4, (int)((P.flags >> 4) & 31u) - 1) : 4;
    bc7::Block blk;
    blk.px = tpx + threadIdx.x;
    blk.stride = BC7_NT;
    const uint32_t nAlpha = P.aux[0];
    const uint32_t* const list = P.aux + 4;
    const uint32_t nTiles = (nTotal + BC7_TB - 1) / BC7_TB;

    uint32_t tile = blockIdx.x;
    while (tile < nTiles) {
        const uint32_t item = tile * BC7_TB + threadIdx.x;
        const int c = item < nTotal ? (item < nAlpha ? 1 : 0) : 2;
        uint8_t* out_at = nullptr;
        if (c < 2) {
            const uint32_t ent = list[item];
            const Level& L = P.lv[ent >> 28];
            const uint32_t b = ent & 0x0FFFFFFFu, by = b / L.nbx, bx = b - by * L.nbx;
            const uint8_t* g = L.src + (size_t)by * 4 * L.rowStride + (size_t)bx * 16;
            const uint32_t align = level_align(L);
            out_at = L.dst + (size_t)b * 16;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint4 v = load_row16(g + (size_t)r * L.rowStride, align);
                tpx[(4 * r + 0) * BC7_NT + threadIdx.x] = v.x;
                tpx[(4 * r + 1) * BC7_NT + threadIdx.x] = v.y;
                tpx[(4 * r + 2) * BC7_NT + threadIdx.x] = v.z;
                tpx[(4 * r + 3) * BC7_NT + threadIdx.x] = v.w;
            }
        }
        const uint4 out = bc7::encode_phased(blk, T, c, level);
        if (c < 2) *(uint4*)out_at = out;
        // tile cost varies several-fold with content, so tiles are claimed, not strided
        if (threadIdx.x == 0) nextTile = gridDim.x + atomicAdd(P.sched, 1u);
        __syncthreads();
        tile = nextTile;
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(P.sched + 1, 1u) == gridDim.x - 1u) { // last CTA out: leave the counters as the next launch expects them
            P.sched[0] = 0;
            P.sched[1] = 0;
            P.aux[0] = 0;
            P.aux[1] = 0;
        }
    }
}

uint32_t bc7_tile_blocks() { return BC7_TB; }

size_t bc7_aux_bytes(uint32_t total_blocks) { return 16 + 4 * (size_t)total_blocks; }

cudaError_t launch_bc7_encode(const EncodeParams& P, int sm_count, cudaStream_t stream)
{
    if (P.totalTiles == 0) return cudaSuccess;
    if (!P.aux) return cudaErrorInvalidValue;
    uint32_t nTotal = 0;
    for (uint32_t i = 0; i < P.nLevels; i++) nTotal += P.lv[i].nblocks;
    if (nTotal == 0) return cudaSuccess;
    const uint32_t nTiles = (nTotal + BC7_TB - 1) / BC7_TB;
    const uint32_t g = (uint32_t)sm_count * BC7_OCC;
    const int grid = (int)(nTiles < g ? nTiles : g);
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
    const uint32_t cgrid = P.totalTiles < (uint32_t)sm_count * 3u ? P.totalTiles : (uint32_t)sm_count * 3u;
    bc7_classify_kernel<<<cgrid, BC7_NT, 0, stream>>>(P, nTotal);
    bc7_encode_kernel<<<grid, BC7_NT, BC7_SMEM, stream>>>(P, nTotal);
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
