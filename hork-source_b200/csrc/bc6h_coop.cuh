// bc6h_coop.cuh -- BC6H encoder, task-parallel form (sm_100a): the search of bc6h_thread.cuh cut into tasks that the lanes of
// a warp execute independently, so that no lane waits for a neighbour's longer search.
//
// Same reference, same arithmetic, same result as bc6h_thread.cuh (D3DX_BC6H::Encode, ThirdParty/bc6h_enc/bc6h_enc.h:
// 2910-2952 and callees); what changes is who does what when.  A warp owns a tile of TB consecutive blocks and runs
//
//   rough     lane = shape: RoughMSE of the 32 two-region shapes of each block (:3565-3640), then the reference's partial
//             sort of the 8 best (:2925-2937) and the one-region fit
//   rounds    the 14 modes in three rounds -- {1, 2, 3-5}, {6, 7-9, 10}, {11, 12, 13, 14} -- each made of
//     setup     lane = (block, precision group, rank): quantise, assign indices, fix-up swap, fit test (first half of Refine,
//               :3420-3440); candidates that fit no mode of their group end here (about a third do)
//     optimise  three passes (R, G, B), lane = (surviving candidate, region): one channel of OptimizeOne (:3233-3267) as a
//               STATE MACHINE that advances one texel of one perturbation trial per iteration.  A trial ends at the texel
//               where its running error reaches the error to beat (see map_colors_quantized_body), which happens after
//               three texels on average -- in lock step a warp would run every trial to the longest of its 32 lanes; here a
//               lane whose trial has ended sets up its next trial in the same iteration, and a lane whose item has ended
//               pulls the next item of the pass
//     finish    lane = surviving candidate: indices of the optimised endpoints, per-mode choice between optimised and
//               original endpoints, emission; the block's best candidate is kept by (error, visiting order), which picks
//               what the reference's sequential search picks (see refine in bc6h_thread.cuh)
//
// Everything here is per-lane code without warp intrinsics (the kernel in bc6h_encode.cu adds those), so
// tools/bc6h_host_check.cpp can run the same functions serially on the CPU against the reference.
#pragma once
#include "bc6h_thread.cuh"

namespace hkbcn {
namespace bc6h {

#ifndef BC6_COOP_TB
#define BC6_COOP_TB 8
#endif
constexpr int COOP_TB = BC6_COOP_TB;           // blocks per warp tile
constexpr int COOP_MAX_TASKS = COOP_TB * 24;  // candidates per round: TB blocks x 3 precision groups x 8 ranks

// precision groups of the two-region modes (modes of one group share RGBAPrec[0][0] and the transform flag)
#ifdef HKBCN_HOST_EMU
static const uint8_t kGroupMode0[6] = {0, 1, 2, 5, 6, 9};
static const uint8_t kGroupModes[6] = {1, 1, 3, 1, 3, 1};
#else
__device__ const uint8_t kGroupMode0[6] = {0, 1, 2, 5, 6, 9};
__device__ const uint8_t kGroupModes[6] = {1, 1, 3, 1, 3, 1};
#endif

// One candidate that passed the fit test.  ep: the quantised endpoints of its regions (absolute values, after the fix-up
// swap) -- `org` when written by setup, replaced region by region with `opt` by the optimise phase.  Two-region modes hold at
// most 11-bit values (int16 x 12); the one-region modes reach 16 bits plus sign (int32 x 6): same 24 bytes.
struct TaskRec {
    uint64_t org_idx;
    float    org_err[2];
    union {
        int16_t ep16[2][6];
        int32_t ep32[6];
    };
    uint32_t meta; // block | rank << 4 | mode0 << 8 | nmodes << 12 | fits << 16 | shape << 20 | regions swapped by the fix-up << 28
    float    cur_err[2]; // error of ep per region as the optimise passes leave it (starts as org_err)
    uint32_t pad;
};
static_assert(sizeof(TaskRec) == 56, "TaskRec layout");

__device__ __forceinline__ uint32_t task_meta(int block, int rank, int mode0, int nmodes, uint32_t fits, uint32_t shape)
{
    return (uint32_t)block | ((uint32_t)rank << 4) | ((uint32_t)mode0 << 8) | ((uint32_t)nmodes << 12) | (fits << 16) | (shape << 20);
}

template <int NR> __device__ __forceinline__ void rec_store_ep(TaskRec& R, int p, const Ep& e)
{
#pragma unroll
    for (int c = 0; c < 3; c++) {
        if (NR == 2) {
            R.ep16[p][c] = (int16_t)e.A[c];
            R.ep16[p][3 + c] = (int16_t)e.B[c];
        } else {
            R.ep32[c] = e.A[c];
            R.ep32[3 + c] = e.B[c];
        }
    }
}
template <int NR> __device__ __forceinline__ void rec_load_ep(const TaskRec& R, int p, Ep& e)
{
#pragma unroll
    for (int c = 0; c < 3; c++) {
        if (NR == 2) {
            e.A[c] = R.ep16[p][c];
            e.B[c] = R.ep16[p][3 + c];
        } else {
            e.A[c] = R.ep32[c];
            e.B[c] = R.ep32[3 + c];
        }
    }
}

// The reference's partial bubble sort of the shapes by rough error (bc6h_enc.h:2925-2937): positions 0..7 end up holding
// the 8 best in ascending order; ties fall as the sequential swaps make them fall, so it is run as written.
__device__ __forceinline__ void rank_shapes(float* rough, uint8_t* order)
{
#pragma unroll 1
    for (int i = 0; i < 8; i++)
#pragma unroll 1
        for (int j = i + 1; j < 32; j++)
            if (rough[i] > rough[j]) {
                const float tf = rough[i]; rough[i] = rough[j]; rough[j] = tf;
                const uint8_t ts = order[i]; order[i] = order[j]; order[j] = ts;
            }
}

// ---- setup: first half of Refine for one candidate; returns the fit mask and fills the record ---------------------------------
template <int N, int NR>
__device__ __forceinline__ uint32_t coop_setup(const Block& blk, int block, int mode0, int nmodes, int rank, bool isSigned, uint32_t shapeIdx, const Ep unq[NR],
                                               TaskRec& R)
{
    Ep org[NR];
    float err[NR];
    uint64_t idx;
    const uint32_t fits = refine_setup<N, NR>(blk, mode0, nmodes, isSigned, shapeIdx, unq, org, idx, err);
    if (fits) {
        R.org_idx = idx;
        R.org_err[0] = err[0];
        R.org_err[1] = NR == 2 ? err[NR - 1] : 0.0f;
        R.cur_err[0] = R.org_err[0];
        R.cur_err[1] = R.org_err[1];
        // which regions the fix-up swap exchanged (SwapIndices, bc6h_enc.h:3295-3318): finish rebuilds `org` from the rough
        // endpoints and this flag instead of keeping a second copy of the endpoints in the record
        const uint8_t* prec = kModes[mode0].prec[0];
        uint32_t swapped = 0;
#pragma unroll
        for (int p = 0; p < NR; p++) {
            rec_store_ep<NR>(R, p, org[p]);
            bool differs = false;
#pragma unroll
            for (int c = 0; c < 3; c++) differs = differs || org[p].A[c] != quantize(unq[p].A[c], prec[c], isSigned);
            swapped |= differs ? 1u << p : 0u;
        }
        R.meta = task_meta(block, rank, mode0, nmodes, fits, shapeIdx) | (swapped << 28);
    }
    return fits;
}

// ---- packed single precision (sm_100: add / sub / fma .f32x2, SASS FADD2 / FFMA2): two IEEE-rounded operations per issue slot.
// Each half is rounded exactly as the scalar instruction rounds it, so results do not change; what changes is the issue count
// of the palette scan, which is what bounds the optimise passes.  ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2
// (a single rounding -- not what the reference computes) whatever the -fmad setting, so squares are taken as fma(d, d, z)
// with z a +0.0 pair the compiler cannot see through (it comes from a kernel parameter): exact, and not contractible.
#ifdef HKBCN_HOST_EMU
struct f2 { float x, y; };
static inline f2 f2_pack(float a, float b) { return f2{a, b}; }
static inline f2 f2_sub(f2 a, f2 b) { return f2{a.x - b.x, a.y - b.y}; }
static inline f2 f2_add(f2 a, f2 b) { return f2{a.x + b.x, a.y + b.y}; }
static inline f2 f2_sq(f2 a, f2 z) { return f2{a.x * a.x + z.x, a.y * a.y + z.y}; } // z = +0: the product, rounded once
static inline float f2_lo(f2 a) { return a.x; }
static inline float f2_hi(f2 a) { return a.y; }
#else
typedef unsigned long long f2;
__device__ __forceinline__ f2 f2_pack(float a, float b) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ f2 f2_sub(f2 a, f2 b) { f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) { f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 f2_sq(f2 a, f2 z) { f2 r; asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(a), "l"(z)); return r; }
__device__ __forceinline__ float f2_lo(f2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return x; }
__device__ __forceinline__ float f2_hi(f2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return y; }
#endif

// ---- optimise: one channel of OptimizeOne (bc6h_enc.h:3233-3267) for one region as a state machine -----------------------------
// Control flow restated from optimize_channel / perturb_one (bc6h_thread.cuh):
//     err0 = perturb(A[CH]) and err1 = perturb(B[CH]), both from the current endpoints           (phase 0, phase 1)
//     apply the better one if it beats opt_err, else done; then alternate B / A / B ...            (phase 2)
//     until a perturbation no longer improves
//   perturb(x): for step = 2^(prec-1) .. 1: try x - step, then x + step (skipping values outside [0, 2^prec)); a trial's error
//     is the region's texels against the palette of the moved endpoints; x moves by the better improving step.
// The optimise phase runs the R, G and B searches of all candidates as three passes (the reference does R, G, B in sequence per
// region, each starting from the previous one's result, which the record carries from pass to pass), so that every lane of the
// warp perturbs the same channel.
// The channel is a run-time value, the same for every lane of a pass: it selects which texel component moves and how the three
// squared differences are associated (see eval_texel) -- one instantiation of the pass loop per palette size instead of three.
// (A compile-time channel saves two selects per palette entry, but the B-channel instantiation then spilled its state machine
// under any register cap: 66 ms per 1024^2 face against 48 ms for this form.)
template <int N, bool SIGNED> struct OptChan {
    int          ch;      // 0, 1, 2 = R, G, B
    const float* tf;      // the block's texels in the encoder's integer domain as floats (exact), 4 per texel
    uint32_t     mask;    // texels of the region (region 0 of a two-region mode: all 16, bc6h_enc.h:3284)
    int          prec;    // endpoint precision (the same for R, G, B in every mode)
    int          a, b;    // committed endpoints of channel ch
    int          ua, ub;  // ... unquantised
    float        opt_err;
    int          phase, do_b;
    float        err0;
    int          val_a;
    int          cur, cand;      // the perturbed endpoint as it stands at this step size, the candidate on trial
    int          cur_u, cand_u;  // ... unquantised: carried along so that no step of the search unquantises a value twice
    int          best_c, best_u; // the best candidate of this step size so far (cur itself while neither sign has improved)
    int          val_a_u;
    int          t;       // index of the current trial within the perturbation: step = 2^(prec-1-t/2), sign = t odd ? + : -
    bool         trial_valid;
    float        min_err;
    f2           pal1[N / 2], pal2[N / 2]; // palettes of the channels at rest (committed endpoints), entries (2q, 2q + 1) as float
                                           // pairs: integers < 2^16, exact
    f2           zero2;                    // a +0.0 pair the compiler cannot fold (see f2_sq)
    int          base, delta;      // the trial's palette of channel ch: entry j = finish((base + delta * w_j) >> 6)
    float        tot;
    uint32_t     tm;
    bool         busy;      // an item is loaded and not finished
    bool         in_trial;  // texels of the current trial remain to be evaluated

    // GeneratePaletteQuantized (bc6h_enc.h:3073-3124): (a * (64 - w) + b * w + 32) >> 6 = (a * 64 + 32 + (b - a) * w) >> 6
    static __device__ __forceinline__ float entry(int base_, int delta_, int j)
    {
        return (float)finish_unquantize((base_ + delta_ * weight_of<N>(j)) >> 6, SIGNED);
    }
    // Unquantize (bc6h_enc.h:2977-3017) without branches
    __device__ __forceinline__ int unq(int comp) const
    {
        if (SIGNED) return unquantize(comp, prec, true);
        const int v = ((comp << 16) + 0x8000) >> prec;
        const int r = comp == 0 ? 0 : comp == ((1 << prec) - 1) ? 0xFFFF : v;
        return prec >= 15 ? comp : r;
    }
    __device__ __forceinline__ void start_perturb(int db)
    {
        do_b = db;
        cur = db ? b : a;
        cur_u = db ? ub : ua;
        best_c = cur;
        best_u = cur_u;
        min_err = opt_err;
        t = -1;
        trial_valid = false;
    }
    // e: the region's committed endpoints (all three channels), err: their error
    __device__ __forceinline__ void start(const float* texels, uint32_t region_mask, int precision, const Ep& e, float err)
    {
        tf = texels;
        mask = region_mask & 0xFFFFu;
        prec = precision;
        a = ch == 0 ? e.A[0] : ch == 1 ? e.A[1] : e.A[2];
        b = ch == 0 ? e.B[0] : ch == 1 ? e.B[1] : e.B[2];
        opt_err = err;
        in_trial = false;
        // no trial can get below an error of zero: the endpoints stay as they are (perturb_one would return old_err every time)
        busy = opt_err > 0.0f;
        if (!busy) return;
        {
            const int c1 = ch == 0 ? 1 : 0; // the two channels at rest, in R, G, B order
            const int u0 = unq(c1 ? e.A[1] : e.A[0]), u1 = unq(c1 ? e.B[1] : e.B[0]);
#pragma unroll
            for (int q = 0; q < N / 2; q++) pal1[q] = f2_pack(entry(u0 * 64 + 32, u1 - u0, 2 * q), entry(u0 * 64 + 32, u1 - u0, 2 * q + 1));
        }
        {
            const int u0 = unq(ch == 2 ? e.A[1] : e.A[2]), u1 = unq(ch == 2 ? e.B[1] : e.B[2]);
#pragma unroll
            for (int q = 0; q < N / 2; q++) pal2[q] = f2_pack(entry(u0 * 64 + 32, u1 - u0, 2 * q), entry(u0 * 64 + 32, u1 - u0, 2 * q + 1));
        }
        ua = unq(a);
        ub = unq(b);
        phase = 0;
        start_perturb(0);
    }
    __device__ __forceinline__ void commit(int db, int v, int vu, float err)
    {
        a = db ? a : v;
        ua = db ? ua : vu;
        b = db ? v : b;
        ub = db ? vu : ub;
        opt_err = err;
    }
    __device__ __forceinline__ void perturb_end()
    {
        const float err = min_err;
        if (phase == 0) {
            err0 = err;
            val_a = cur;
            val_a_u = cur_u;
            phase = 1;
            start_perturb(1);
        } else if (phase == 1) {
            if (err0 < err) {
                if (err0 >= opt_err) { busy = false; return; }
                commit(0, val_a, val_a_u, err0);
                phase = 2;
                start_perturb(1);
            } else {
                if (err >= opt_err) { busy = false; return; }
                commit(1, cur, cur_u, err);
                phase = 2;
                start_perturb(0);
            }
        } else {
            if (err >= opt_err) { busy = false; return; }
            commit(do_b, cur, cur_u, err);
            start_perturb(1 - do_b);
        }
    }
    // The previous trial (if any) is over: account for it and move to the next candidate -- ONE transition per call, written
    // as straight-line selects (the warp runs this body every iteration for the few lanes that are between trials, so every
    // instruction here is paid by all 32): only the end of a perturbation (every 2 * prec calls) branches.  A "-" candidate
    // below zero is skipped on the spot (its "+" twin shares the step and the state); a "+" candidate above the range leaves the
    // lane without a trial until the next call.
    __device__ __forceinline__ void advance()
    {
        const bool better = trial_valid && tot < min_err;
        min_err = better ? tot : min_err;
        best_c = better ? cand : best_c;
        best_u = better ? cand_u : best_u;
        trial_valid = false;
        t++;
        if (!(t & 1)) { // both signs of a step have been tried: the endpoint moves to the better one (or stays)
            cur = best_c;
            cur_u = best_u;
        }
        if (t == 2 * prec) {
            perturb_end();
            return;
        }
        const int step = 1 << (prec - 1 - (t >> 1));
        const bool skip = !(t & 1) && cur < step; // cur - step < 0: straight to cur + step
        t += skip ? 1 : 0;
        cand = cur + ((t & 1) ? step : -step);
        // an out-of-range candidate is skipped; an empty trial (error to beat already 0) would total 0, not below min_err
        const bool ok = (unsigned)cand < (1u << prec) && min_err > 0.0f;
        cand_u = unq(cand);
        const int u0 = do_b ? ua : cand_u, u1 = do_b ? cand_u : ub;
        base = u0 * 64 + 32;
        delta = u1 - u0;
        tot = 0.0f;
        tm = mask;
        trial_valid = ok;
        in_trial = ok;
    }
    // one texel of the current trial: nearest palette entry by the reference's ordered scan (nearest_entry), added to the total.
    // The moving channel's palette entries are computed on the spot (five integer operations each): a trial lasts four texels
    // on average and a third of them one, and a separate palette set-up would run with the few lanes that need it.
    __device__ __forceinline__ void eval_texel()
    {
        const int i = __ffs((int)tm) - 1;
        tm &= tm - 1u;
#ifdef HKBCN_HOST_EMU
        const float t0 = tf[4 * i], t1 = tf[4 * i + 1], t2 = tf[4 * i + 2];
#else
        const float4 tx = *(const float4*)(tf + 4 * i);
        const float t0 = tx.x, t1 = tx.y, t2 = tx.z;
#endif
        const float tm_ = ch == 0 ? t0 : ch == 1 ? t1 : t2; // the moving channel's component, then the two at rest in R, G, B order
        const float tc1 = ch == 0 ? t1 : t0, tc2 = ch == 2 ? t1 : t2;
        HK_COUNT(N);
        const f2 bm = f2_pack(tm_, tm_), b1 = f2_pack(tc1, tc1), b2 = f2_pack(tc2, tc2);
        const bool chB = ch == 2; // the same for every lane of the pass: a uniform branch, not a select per entry
        float best = 0.0f, prev = 0.0f;
        bool open = true;
#pragma unroll
        for (int q = 0; q < N / 2; q++) {
            const f2 dm = f2_sub(bm, f2_pack(entry(base, delta, 2 * q), entry(base, delta, 2 * q + 1)));
            const f2 sm = f2_sq(dm, zero2), s1 = f2_sq(f2_sub(b1, pal1[q]), zero2), s2 = f2_sq(f2_sub(b2, pal2[q]), zero2);
            // (dR^2 + dG^2) + dB^2 in exactly this association: with m the moving channel and c1 < c2 the two at rest, R and G
            // give (c1^2 + m^2) + c2^2 (the inner sum commutes) and B gives (c1^2 + c2^2) + m^2
            const f2 e2 = chB ? f2_add(f2_add(s1, s2), sm) : f2_add(f2_add(s1, sm), s2);
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const float e = h ? f2_hi(e2) : f2_lo(e2);
                // nearest_entry without the index: stay on the entries while the error does not increase.  (Its `best > 0` stop
                // changes nothing here: from 0 the error can only stay 0 or close the scan.)  While the scan is open `best` is
                // the previous entry's error, so the test is written against that: the comparisons then depend on each other
                // only through the predicate, not through a chain of selects.
                if (q == 0 && h == 0) best = e;
                else {
                    open = open && e <= prev;
                    best = open ? e : best;
                }
                prev = e;
            }
        }
        tot += best;
        if (tm == 0u || !(tot < min_err)) in_trial = false;
    }
};

// ---- finish: second half of Refine for one candidate ---------------------------------------------------------------------------
// best_err / best_ord: the block's best so far (a stale value only costs an emission that loses the commit later)
template <int N, int NR>
__device__ __forceinline__ bool coop_finish(const Block& blk, const Layout& L, const TaskRec& R, bool isSigned, const Ep unq[NR], float& best_err, int& best_ord,
                                            uint4& out)
{
    const int rank = (R.meta >> 4) & 15, mode0 = (R.meta >> 8) & 15, nmodes = (R.meta >> 12) & 15;
    const uint32_t fits = (R.meta >> 16) & 15u, shapeIdx = (R.meta >> 20) & 31u;
    const uint8_t* prec = kModes[mode0].prec[0];
    // org again: the quantised rough endpoints, exchanged where setup's fix-up swap exchanged them
    Ep org[NR], opt[NR];
#pragma unroll
    for (int p = 0; p < NR; p++) {
        rec_load_ep<NR>(R, p, opt[p]);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            org[p].A[c] = quantize(unq[p].A[c], prec[c], isSigned);
            org[p].B[c] = quantize(unq[p].B[c], prec[c], isSigned);
        }
    }
    // which regions setup swapped: bit p of the meta word's top bits
    const uint32_t swapped = R.meta >> 28;
#pragma unroll
    for (int p = 0; p < NR; p++)
        if (swapped >> p & 1u) {
#pragma unroll
            for (int c = 0; c < 3; c++) {
                const int t = org[p].A[c];
                org[p].A[c] = org[p].B[c];
                org[p].B[c] = t;
            }
        }
    const int ord0 = best_ord;
    const float err0 = best_err;
    float oe[NR];
    oe[0] = R.org_err[0];
    if (NR == 2) oe[NR - 1] = R.org_err[1];
    refine_finish<N, NR>(blk, L, mode0, nmodes, fits, rank, isSigned, shapeIdx, org, R.org_idx, oe, opt, best_err, best_ord, out);
    return best_ord != ord0 || best_err != err0;
}

} // namespace bc6h
} // namespace hkbcn
