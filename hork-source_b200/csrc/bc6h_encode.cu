// bc6h_encode.cu -- BC6H surface encoder (sm_100a).
//
// Replaces TextureBlockCompression::CompressBC6h (Hork/Image/ImageEncoders.cpp:395-521) ->
// Encode_BC6h_f32 / Encode_BC6h_f16 (:187-205) -> bc6h_enc::EncodeBC6HU / EncodeBC6HS
// (ThirdParty/bc6h_enc/bc6h_enc.h:3650-3660).  See bc6h_thread.cuh for the per-block search and the parity contract,
// bc6h_coop.cuh for how the search is cut into tasks.
// Unlike the reference's 16-thread driver (ImageEncoders.cpp:426-427) no tail block is dropped.
//
// Kernel shape: persistent one-warp CTAs, tiles of 8 consecutive blocks claimed from the launch's counter.  The warp loads
// the tile's texels (2 KB of RGBA32F, 1 KB of RGBA16F: coalesced 16-byte loads, block rows of one texel row are adjacent),
// converts each once to its half bit pattern (half_from_float, Hork/Core/half.c:802) and to the encoder's integer domain,
// and then runs the phases of bc6h_coop.cuh with lane = shape / candidate / (candidate, region); the only cross-lane
// traffic is a ballot per phase step (task lists) and the best-candidate commit.
// Algorithmic HBM bytes: 17 B/px (f32 source) / 9 B/px (f16 source); ALU-bound by four to five orders of magnitude.
#include <cuda_fp16.h>
#include "hkbcn_common.cuh"
#include "hkbcn_kernels.h"
#include "bc6h_thread.cuh"
#include "bc6h_coop.cuh"
#include "bc6h_layout.h"

namespace hkbcn {

__device__ bc6h::Layout g_bc6h_enc_layout;

#ifndef BC6E_CTAS_PER_SM
#define BC6E_CTAS_PER_SM 12 // resident one-warp CTAs per SM
#endif
#ifndef BC6E_MIN_CTAS
#define BC6E_MIN_CTAS 12    // launch bound: 12 one-warp CTAs per SM = 168 registers per thread, which is also what 17.8 KB of shared
                            // memory per CTA allows.  Measured on a 1024^2 face: 128 registers 57 ms (spills, still 12 resident),
                            // 168 registers 43 ms, 255 registers / 8 warps 55 ms.  Tiles of 4 blocks (9.3 KB, -DBC6_COOP_TB=4) make
                            // room for more: 12 / 14 / 16 warps 46.7 / 45.7 / 43.6 ms -- the smaller tile's worse balance of the
                            // optimise items costs what the four extra warps bring
#endif
constexpr int BC6_NT = 32;
constexpr int BC6_TB = bc6h::COOP_TB;

struct Bc6Smem {
    alignas(16) int tq[BC6_TB][64];                      // texels in the encoder's integer domain, 4 ints per texel
    alignas(16) float tf[BC6_TB][64];                    // ... as floats (|v| < 2^15: exact), what the optimise passes read
    union {                                              // rough phase / rounds
        struct {
            alignas(16) float th[BC6_TB][64];            // HDR texels, 4 floats per texel
            float rough[BC6_TB][32];
        } r;
        bc6h::TaskRec rec[bc6h::COOP_MAX_TASKS];
    } u;
    int16_t  unq8[BC6_TB][8][2][6];                      // unquantised endpoints of the 8 ranked shapes (A[3], B[3] per region)
    int16_t  unq1[BC6_TB][6];                            // ... of the one-region fit
    uint8_t  order[BC6_TB][32];
    float    best_err[BC6_TB];
    int      best_ord[BC6_TB];
    alignas(16) uint4 best_blk[BC6_TB];
    uint32_t nan_sign[BC6_TB];
    alignas(16) bc6h::Layout L6;
    uint32_t nextTile;
};

__device__ __forceinline__ void load_unq(const int16_t* s, bc6h::Ep& e)
{
#pragma unroll
    for (int c = 0; c < 3; c++) {
        e.A[c] = s[c];
        e.B[c] = s[3 + c];
    }
}
__device__ __forceinline__ void store_unq(int16_t* s, const bc6h::Ep& e)
{
#pragma unroll
    for (int c = 0; c < 3; c++) {
        s[c] = (int16_t)e.A[c];
        s[3 + c] = (int16_t)e.B[c];
    }
}

// The phases around the optimise passes run once per candidate, not once per texel: real functions, so that their register
// needs (assign_indices keeps two palettes of eight entries, the rough fit a dozen accumulators) are theirs alone and the
// passes' loop gets an allocation of its own.
template <int N>
__device__ __noinline__ float rough_region_fn(const float* th, const int* tq, uint32_t nan_sign, uint32_t mask, bool isSigned, int16_t* unq_out)
{
    bc6h::Block blk{th, tq, 4, 1};
    blk.nan_sign = nan_sign;
    bc6h::Ep e;
    const float err = bc6h::rough_region<N>(blk, mask, isSigned, e);
    if (unq_out) store_unq(unq_out, e);
    return err;
}
template <int N, int NR>
__device__ __noinline__ uint32_t setup_fn(const int* tq, int k, int mode0, int nmodes, int rank, bool isSigned, uint32_t shapeIdx, const int16_t* unq_s,
                                          bc6h::TaskRec* R)
{
    bc6h::Block blk{nullptr, tq, 4, 1};
    bc6h::Ep unq[NR];
    load_unq(unq_s, unq[0]);
    if (NR == 2) load_unq(unq_s + 6, unq[NR - 1]);
    return bc6h::coop_setup<N, NR>(blk, k, mode0, nmodes, rank, isSigned, shapeIdx, unq, *R);
}
template <int N, int NR>
__device__ __noinline__ bool finish_fn(const int* tq, const bc6h::Layout* L, const bc6h::TaskRec* R, bool isSigned, const int16_t* unq_s, float* err, int* ord,
                                       uint4* out)
{
    bc6h::Block blk{nullptr, tq, 4, 1};
    bc6h::Ep unq[NR];
    load_unq(unq_s, unq[0]);
    if (NR == 2) load_unq(unq_s + 6, unq[NR - 1]);
    return bc6h::coop_finish<N, NR>(blk, *L, *R, isSigned, unq, *err, *ord, *out);
}

// one pass (channel CH) of the optimise phase of one round: lanes pull (candidate, region) items and run OptChan to completion,
// one texel per iteration
template <int N, int NR, bool SIGNED>
__device__ __noinline__ void optimise_pass(Bc6Smem& S, int nSurv, const int CH, const unsigned long long zero2)
{
    const uint32_t lane = threadIdx.x;
    const uint32_t lt = (1u << lane) - 1u;
    const int nItems = nSurv * NR;
    bc6h::OptChan<N, SIGNED> M;
    M.ch = CH;
    M.zero2 = zero2;
    M.busy = false;
    M.in_trial = false;
    int next = 0, mine = -1;
#pragma unroll 1
    for (;;) {
        if (__any_sync(0xFFFFFFFFu, !M.busy)) { // some lane is between items (an item ends every few hundred iterations)
            const bool idle = !M.busy;
            if (idle && mine >= 0) { // the item just finished: channel CH of the region's endpoints and the region's error go back
                bc6h::TaskRec& R = S.u.rec[NR == 2 ? mine >> 1 : mine];
                const int p = NR == 2 ? mine & 1 : 0;
                if (NR == 2) {
                    R.ep16[p][CH] = (int16_t)M.a;
                    R.ep16[p][3 + CH] = (int16_t)M.b;
                } else {
                    R.ep32[CH] = M.a;
                    R.ep32[3 + CH] = M.b;
                }
                R.cur_err[p] = M.opt_err;
                mine = -1;
            }
            if (next < nItems) {
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, idle);
                const int my = next + __popc(m & lt);
                next += __popc(m);
                if (idle && my < nItems) {
                    const bc6h::TaskRec& R = S.u.rec[NR == 2 ? my >> 1 : my];
                    const int p = NR == 2 ? my & 1 : 0;
                    const int mode0 = (R.meta >> 8) & 15;
                    const uint32_t shapeIdx = (R.meta >> 20) & 31u;
                    bc6h::Ep e;
                    bc6h::rec_load_ep<NR>(R, p, e);
                    // OptimizeEndPoints (bc6h_enc.h:3271-3293): region 0 over g_aPartitionTable[0] = every texel, region 1 over its own
                    M.start(S.tf[R.meta & 15u], p ? kShape2[shapeIdx] : 0xFFFFu, bc6h::kModes[mode0].prec[0][0], e, R.cur_err[p]);
                    mine = my;
                }
            }
            if (!__any_sync(0xFFFFFFFFu, M.busy)) {
                if (next >= nItems && !__any_sync(0xFFFFFFFFu, mine >= 0)) break;
                continue;
            }
        }
        // a lane whose trial has ended moves to its next candidate
        if (M.busy && !M.in_trial) M.advance();
        if (M.in_trial) M.eval_texel();
    }
    __syncwarp(); // the next pass reads what this one wrote
}

template <bool F16SRC, bool SIGNED>
__global__ void __launch_bounds__(BC6_NT, BC6E_MIN_CTAS) bc6h_encode_kernel(const __grid_constant__ EncodeParams P)
{
    constexpr int BPP = F16SRC ? 8 : 16;
    __shared__ Bc6Smem S;
    const uint32_t lane = threadIdx.x;
    const uint32_t lt = (1u << lane) - 1u;
    constexpr bool isSigned = SIGNED;
    // a +0.0 pair from a kernel parameter (EncodeParams::pad is always 0): see f2_sq in bc6h_coop.cuh
    const unsigned long long zero2 = ((unsigned long long)P.pad << 32) | P.pad;
    {
        const uint32_t* src = (const uint32_t*)&g_bc6h_enc_layout;
        uint32_t* dst = (uint32_t*)&S.L6;
        for (uint32_t i = lane; i < sizeof(bc6h::Layout) / 4; i += BC6_NT) dst[i] = src[i];
    }
    __syncwarp();

    uint32_t tile = blockIdx.x;
    while (tile < P.totalTiles) {
        const Level& L = P.lv[find_level(P, tile)];
        const uint32_t b0 = (tile - L.tileBase) * BC6_TB;
        const int nb = (int)min((uint32_t)BC6_TB, L.nblocks - b0);
        // ---- load and convert the tile's texels: lane -> (block, texel); 4 consecutive lanes read one 64-byte block row
        if (lane < (uint32_t)nb) S.nan_sign[lane] = 0x80000000u;
        __syncwarp();
#pragma unroll 1
        for (int idx = (int)lane; idx < nb * 16; idx += BC6_NT) {
            const int k = idx >> 4, t = idx & 15;
            const uint32_t b = b0 + (uint32_t)k;
            const uint32_t by = b / L.nbx, bx = b - by * L.nbx;
            const uint8_t* g = L.src + ((size_t)by * 4 + (uint32_t)(t >> 2)) * L.rowStride + ((size_t)bx * 4 + (uint32_t)(t & 3)) * BPP;
            float v[3];
            if (F16SRC) {
                // Encode_BC6h_f16 widens the 64 halfs with f16tof32 first (ImageEncoders.cpp:187-197)
                uint32_t lo, hi;
                if (L.bulkOk) {
                    const uint2 h = __ldg((const uint2*)g);
                    lo = h.x;
                    hi = h.y;
                } else {
                    const uint16_t* q = (const uint16_t*)g;
                    lo = (uint32_t)q[0] | ((uint32_t)q[1] << 16);
                    hi = q[2];
                }
                v[0] = __half2float(__ushort_as_half((unsigned short)(lo & 0xFFFFu)));
                v[1] = __half2float(__ushort_as_half((unsigned short)(lo >> 16)));
                v[2] = __half2float(__ushort_as_half((unsigned short)(hi & 0xFFFFu)));
            } else if (L.bulkOk) {
                const float4 f = __ldg((const float4*)g);
                v[0] = f.x; v[1] = f.y; v[2] = f.z;
            } else {
                const float* q = (const float*)g;
                v[0] = q[0]; v[1] = q[1]; v[2] = q[2];
            }
            bool nan_in = false;
#pragma unroll
            for (int c = 0; c < 3; c++) {
                S.u.r.th[k][4 * t + c] = v[c];
                const int q = bc6h::f16_to_int(bc6h::half_from_float_bits(__float_as_uint(v[c])), isSigned);
                S.tq[k][4 * t + c] = q;
                S.tf[k][4 * t + c] = (float)q;
                nan_in = nan_in || v[c] != v[c];
            }
            S.u.r.th[k][4 * t + 3] = 0.0f;
            S.tq[k][4 * t + 3] = 0;
            S.tf[k][4 * t + 3] = 0.0f;
            if (nan_in) S.nan_sign[k] = 0u; // see Block::nan_sign
        }
        __syncwarp();
        // ---- rough: lane = shape (RoughMSE, bc6h_enc.h:3565-3640)
#pragma unroll 1
        for (int k = 0; k < nb; k++) {
            const uint32_t m1 = kShape2[lane];
            float err = rough_region_fn<8>(S.u.r.th[k], S.tq[k], S.nan_sign[k], ~m1 & 0xFFFFu, isSigned, nullptr);
            err += rough_region_fn<8>(S.u.r.th[k], S.tq[k], S.nan_sign[k], m1, isSigned, nullptr);
            S.u.r.rough[k][lane] = err;
            S.order[k][lane] = (uint8_t)lane;
        }
        __syncwarp();
        // ---- rank: lane = block; the one-region fit on the same lanes
        if (lane < (uint32_t)nb) {
            bc6h::rank_shapes(S.u.r.rough[lane], S.order[lane]);
            rough_region_fn<16>(S.u.r.th[lane], S.tq[lane], S.nan_sign[lane], 0xFFFFu, isSigned, S.unq1[lane]);
            S.best_err[lane] = FLT_MAX;
            S.best_ord[lane] = 0x7FFFFFFF;
            S.best_blk[lane] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
        // ---- unquantised endpoints of the ranked shapes: lane = (block, rank)
#pragma unroll 1
        for (int it = (int)lane; it < nb * 8; it += BC6_NT) {
            const int k = it >> 3, rank = it & 7;
            const uint32_t m1 = kShape2[S.order[k][rank]];
            rough_region_fn<8>(S.u.r.th[k], S.tq[k], S.nan_sign[k], ~m1 & 0xFFFFu, isSigned, S.unq8[k][rank][0]);
            rough_region_fn<8>(S.u.r.th[k], S.tq[k], S.nan_sign[k], m1, isSigned, S.unq8[k][rank][1]);
        }
        __syncwarp(); // the HDR texels and the rough errors are dead from here on: their storage becomes the task records

        // ---- the three rounds: setup -> optimise -> finish
#pragma unroll 1
#ifndef BC6_ROUND_LO
#define BC6_ROUND_LO 0
#define BC6_ROUND_HI 3
#endif
        for (int round = BC6_ROUND_LO; round < BC6_ROUND_HI; round++) { // (timing experiments cut rounds out)
            const int ntask = round < 2 ? nb * 24 : nb * 4;
            int nSurv = 0;
#pragma unroll 1
            for (int t0 = 0; t0 < ntask; t0 += BC6_NT) {
                const int t = t0 + (int)lane;
                bc6h::TaskRec R;
                uint32_t fits = 0;
                if (t < ntask) {
                    if (round < 2) {
                        // group-major, so that the lanes of the optimise phase mostly search at one precision
                        const int gi = t / (nb * 8), rem = t - gi * nb * 8;
                        const int k = rem >> 3, rank = rem & 7, g = round * 3 + gi;
                        fits = setup_fn<8, 2>(S.tq[k], k, bc6h::kGroupMode0[g], bc6h::kGroupModes[g], rank, isSigned, S.order[k][rank], S.unq8[k][rank][0], &R);
                    } else {
                        // longest searches first (mode 14 has 16-bit endpoints: 16 steps per perturbation)
                        const int mi = t / nb, k = t - mi * nb;
                        fits = setup_fn<16, 1>(S.tq[k], k, 13 - mi, 1, 0, isSigned, 0, S.unq1[k], &R);
                    }
                }
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, fits != 0u);
                if (fits) S.u.rec[nSurv + __popc(m & lt)] = R;
                nSurv += __popc(m);
            }
            __syncwarp();
            // OptimizeOne's R, G, B searches as three passes over the round's candidates
#pragma unroll 1
            for (int ch = 0; ch < 3; ch++) {
                if (round < 2) optimise_pass<8, 2, SIGNED>(S, nSurv, ch, zero2);
                else optimise_pass<16, 1, SIGNED>(S, nSurv, ch, zero2);
            }
#pragma unroll 1
            for (int s0 = 0; s0 < nSurv; s0 += BC6_NT) {
                const int s = s0 + (int)lane;
                bool won = false;
                int k = 0, ord = 0;
                float err = 0.0f;
                uint4 out = make_uint4(0, 0, 0, 0);
                if (s < nSurv) {
                    const bc6h::TaskRec& R = S.u.rec[s];
                    k = (int)(R.meta & 15u);
                    const int rank = (R.meta >> 4) & 15;
                    err = S.best_err[k];
                    ord = S.best_ord[k];
                    if (round < 2) won = finish_fn<8, 2>(S.tq[k], &S.L6, &R, isSigned, S.unq8[k][rank][0], &err, &ord, &out);
                    else won = finish_fn<16, 1>(S.tq[k], &S.L6, &R, isSigned, S.unq1[k], &err, &ord, &out);
                }
                // commit: the candidates that beat the block's best as it stood are applied one lane at a time, each re-checked
                // against the best as it stands now -- (error, visiting order) ascending, the order the reference's strict `<` keeps
                uint32_t m = __ballot_sync(0xFFFFFFFFu, won);
                while (m) {
                    const uint32_t l = (uint32_t)__ffs((int)m) - 1u;
                    m &= m - 1u;
                    if (lane == l && (err < S.best_err[k] || (err == S.best_err[k] && ord < S.best_ord[k]))) {
                        S.best_err[k] = err;
                        S.best_ord[k] = ord;
                        S.best_blk[k] = out;
                    }
                    __syncwarp();
                }
            }
            __syncwarp();
        }
        if (lane < (uint32_t)nb) *(uint4*)(L.dst + (size_t)(b0 + lane) * 16) = S.best_blk[lane];
        // tiles are claimed from the launch's counter, not strided: a tile's cost varies several-fold with content
        if (lane == 0) S.nextTile = gridDim.x + atomicAdd(P.sched, 1u);
        __syncwarp();
        tile = S.nextTile;
        __syncwarp();
    }
    if (lane == 0) {
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
    // one tile per warp iteration; the grid is one resident wave (or the tile count if smaller) and warps claim further tiles
    const uint32_t g = (uint32_t)sm_count * BC6E_CTAS_PER_SM;
    const int grid = (int)(P.totalTiles < g ? P.totalTiles : g);
    if (src_f16) {
        if (is_signed) bc6h_encode_kernel<true, true><<<grid, BC6_NT, 0, stream>>>(P);
        else bc6h_encode_kernel<true, false><<<grid, BC6_NT, 0, stream>>>(P);
    } else {
        if (is_signed) bc6h_encode_kernel<false, true><<<grid, BC6_NT, 0, stream>>>(P);
        else bc6h_encode_kernel<false, false><<<grid, BC6_NT, 0, stream>>>(P);
    }
    return cudaGetLastError();
}

cudaError_t bc6h_upload_tables()
{
    static bc6h::Layout H;
    bc6h_parse_layout(H.tok);
    return cudaMemcpyToSymbol(g_bc6h_enc_layout, &H, sizeof(H));
}

} // namespace hkbcn
