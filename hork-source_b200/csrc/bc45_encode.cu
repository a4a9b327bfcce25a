// bc45_encode.cu -- BC4 / BC5 surface encoders (sm_100a).
//
// Replaces TextureBlockCompression::CompressBC4 / CompressBC5
// (Hork/Image/ImageEncoders.cpp:323-392) -> Encode_BC4/5(..., bMaxQuality = true)
// (:171-185) -> rgbcx::encode_bc4_hq / encode_bc5_hq (rgbcx.cpp:2730-2884, 2921-2927).
// Bit-exact contract; see bc4_search.cuh for the tie-break argument.
//
// Mapping: a sub-warp of 2 lanes per channel-block (lane parity = BC4 mode), i.e.
// 2 lanes per BC4 block, 4 lanes per BC5 block; a 256-thread CTA encodes a tile of
// 128 (BC4) / 64 (BC5) consecutive blocks staged in shared memory by the TMA bulk
// engine.  Algorithmic HBM bytes: 1.5 B/px (BC4), 3 B/px (BC5); the kernel is
// ALU-bound by more than two orders of magnitude.
#include "hkbcn_common.cuh"
#include "bc4_search.cuh"
#include "hkbcn_kernels.h"

namespace hkbcn {

template <int CH> // CH = channels = bytes per texel (1: BC4, 2: BC5)
__global__ void __launch_bounds__(256) bc45_encode_kernel(const __grid_constant__ EncodeParams P)
{
    constexpr int NT = 256;
    constexpr int G = 2 * CH;       // lanes per block
    constexpr int TB = NT / G;      // blocks per tile
    constexpr int BW = 4 * CH;      // bytes per block row
    constexpr int ROWB = TB * BW;
    __shared__ __align__(128) uint8_t stage[4 * ROWB];
    __shared__ __align__(8) uint64_t bar;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    uint32_t phase = 0;

    const int k = threadIdx.x / G;
    const int sub = threadIdx.x % G;
    const int ch = sub >> 1;
    const int mode = sub & 1;
    const bool fast = (P.flags & 2u) != 0; // HKBCN_FLAG_FAST

    for (uint32_t tile = blockIdx.x; tile < P.totalTiles; tile += gridDim.x) {
        const Level& L = P.lv[find_level(P, tile)];
        const uint32_t b0 = (tile - L.tileBase) * TB;
        const uint32_t nb = min((uint32_t)TB, L.nblocks - b0);

        stage_tile<CH, TB, NT>(L, b0, nb, stage, &bar, phase);

        const bool valid = (uint32_t)k < nb;
        uint32_t px2[8];
        int minv = 255, maxv = 0;
        if (valid) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint8_t* s = stage + r * ROWB + k * BW + ch;
                int p0 = s[0], p1 = s[CH], p2 = s[2 * CH], p3 = s[3 * CH];
                px2[r * 2 + 0] = p0 | (p1 << 16);
                px2[r * 2 + 1] = p2 | (p3 << 16);
                minv = min(min(minv, min(p0, p1)), min(p2, p3));
                maxv = max(max(maxv, max(p0, p1)), max(p2, p3));
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) px2[i] = 0;
            minv = maxv = 0;
        }

        uint2 out;
        if (fast) {
            out = bc4_fast(px2, minv, maxv);
        } else {
            uint32_t key = 0xFFFFFFFFu;
            if (minv != maxv) key = bc4hq_search_mode(px2, minv, maxv, mode);
            key = min(key, __shfl_xor_sync(0xFFFFFFFFu, key, 1));
            if (mode == 0) {
                if (minv == maxv) out = bc4_solid((uint32_t)minv);
                else out = bc4hq_emit(px2, minv, maxv, key & 127u);
            }
        }
        if (valid && mode == 0) {
            uint2* d = (uint2*)(L.dst + ((size_t)(b0 + k) * CH + ch) * 8);
            *d = out;
        }
        __syncthreads(); // stage is reused by the next tile
    }
}

static int grid_for(uint32_t tiles, int sm_count, int ctas_per_sm)
{
    uint32_t g = (uint32_t)sm_count * ctas_per_sm;
    return (int)(tiles < g ? (tiles ? tiles : 1) : g);
}

cudaError_t launch_bc45_encode(int channels, const EncodeParams& P, int sm_count, cudaStream_t stream)
{
    if (P.totalTiles == 0) return cudaSuccess;
    int grid = grid_for(P.totalTiles, sm_count, 8);
    if (channels == 1) bc45_encode_kernel<1><<<grid, 256, 0, stream>>>(P);
    else bc45_encode_kernel<2><<<grid, 256, 0, stream>>>(P);
    return cudaGetLastError();
}

uint32_t bc45_tile_blocks(int channels) { return 256 / (2 * channels); }

} // namespace hkbcn
