// hkbcn_api.cu -- the C-ABI layer (include/hkbcn.h) over the sm_100a kernels.
//
// Host side of the drop-in boundary for Hk::TextureBlockCompression
// (Hork/Image/ImageEncoders.h:37-87).  Owns per-device state (stream, lazily grown
// device scratch), builds the launch descriptors, moves host buffers, and reports
// errors as status codes -- the reference's functions return void and cannot fail;
// this path can (no device, out of memory) and never falls back to the CPU.
#include "../../include/hkbcn.h"
#include "hkbcn_kernels.h"

#include <nvtx3/nvToolsExt.h> // header-only; ranges cost nothing unless a tool (nsys / ncu --nvtx) is attached
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>
#include <map>
#include <math.h>
#include <string>
#include <algorithm>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

using namespace hkbcn;

namespace {

thread_local char t_err[512] = "";
void set_err(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
}

// NVTX range over one C-ABI call (tracing): the entry points show up by name on a timeline
struct Range {
    explicit Range(const char* name) { nvtxRangePushA(name); }
    ~Range() { nvtxRangePop(); }
};

struct Device {
    int          id = -1;
    int          sm_count = 0;
    cudaStream_t stream = nullptr;
    void*        d_src = nullptr; size_t src_cap = 0;
    void*        d_dst = nullptr; size_t dst_cap = 0;
    std::mutex   mu; // serialises host-pointer calls that share the scratch buffers
    // pageable callers (Image.cpp hands over HeapBlob memory): upload ring of two pinned pieces, a pinned copy of the output,
    // a copy stream and events -- see compress_host_pipelined
    cudaStream_t copy_stream = nullptr;
    cudaStream_t stream2 = nullptr; // second encode stream of the piece pipeline (compress_host_pipelined)
    void*        h_in[2] = {nullptr, nullptr}; size_t h_in_cap[2] = {0, 0};
    void*        h_out = nullptr; size_t h_out_cap = 0;
    cudaEvent_t  ev_in[2] = {nullptr, nullptr};
    cudaEvent_t  ev_piece[12] = {};
    // mip cascade: decode tables and tap tables (one pair per in/out size, filter and edge mode), built on first use
    float*       lut_srgb = nullptr;
    float*       lut_lin = nullptr;
    struct Taps { int* idx; float* w; int T; int *first, *cnt, *map; int margin; int spanOk; int uniform; int first0; float u[8]; };
    std::map<uint64_t, Taps> taps;
    std::mutex   taps_mu;
    // tile counters of the persistent kernels: one zeroed {next tile, finished CTAs} pair PER STREAM.  Launches on one stream
    // run one after another and a kernel leaves its pair zeroed again, so two kernels never share a live pair however many
    // launches are outstanding; streams get their pair on first use (chunks of SCHED_CHUNK pairs, never moved or freed
    // before shutdown because kernels in flight hold pointers into them).
    static constexpr uint32_t SCHED_CHUNK = 64;
    std::vector<uint32_t*> sched_chunks;
    std::map<cudaStream_t, uint32_t*> sched_of_stream;
    uint32_t     sched_used = 0;
    std::mutex   sched_mu;
    // BC7 work lists (bc7_encode.cu): one buffer per stream, grown on demand; a buffer that was outgrown may still be read by a
    // kernel in flight, so it is parked until shutdown instead of freed
    struct Aux { uint32_t* p = nullptr; size_t cap = 0; };
    std::map<cudaStream_t, Aux> aux_of_stream;
    std::vector<uint32_t*> aux_retired;
    // hkbcn_import_batch: two pipeline slots, each with its own stream, device buffers and pinned staging, so the upload of
    // texture i + 1 and the download of texture i - 1 overlap the kernels of texture i
    struct Slot {
        cudaStream_t st = nullptr;
        void *d_src = nullptr, *d_dst = nullptr, *h_src = nullptr, *h_dst = nullptr;
        size_t dsc = 0, ddc = 0, hsc = 0, hdc = 0;
    } slots[2];
    std::mutex   batch_mu;
};

std::mutex            g_mu;
std::atomic<bool>     g_init{false}; // published (release) only after g_devs is complete; readers load with acquire
std::vector<Device*>  g_devs;      // selected devices
std::atomic<uint64_t> g_launches{0};
std::vector<double>   g_batch_ms;  // per-device worker time of the last batch call (hkbcn_batch_stats)
std::mutex            g_batch_mu;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            set_err("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);   \
            return HKBCN_ERR_CUDA;                                                                 \
        }                                                                                          \
    } while (0)

int codec_of(int fmt)
{
    switch (fmt) {
    case HKBCN_BC1_UNORM: case HKBCN_BC1_UNORM_SRGB: return C_BC1;
    case HKBCN_BC2_UNORM: case HKBCN_BC2_UNORM_SRGB: return C_BC2;
    case HKBCN_BC3_UNORM: case HKBCN_BC3_UNORM_SRGB: return C_BC3;
    case HKBCN_BC4_UNORM: case HKBCN_BC4_SNORM: return C_BC4;
    case HKBCN_BC5_UNORM: case HKBCN_BC5_SNORM: return C_BC5;
    case HKBCN_BC6H_UFLOAT: case HKBCN_BC6H_SFLOAT: return C_BC6H;
    case HKBCN_BC7_UNORM: case HKBCN_BC7_UNORM_SRGB: return C_BC7;
    }
    return -1;
}
int blk_bytes(int codec) { return (codec == C_BC1 || codec == C_BC4) ? 8 : 16; }
int src_bpp(int codec, uint32_t flags)
{
    switch (codec) {
    case C_BC4: return 1;
    case C_BC5: return 2;
    case C_BC6H: return (flags & HKBCN_FLAG_SRC_F16) ? 8 : 16;
    default: return 4;
    }
}
int dst_bpp(int codec, uint32_t flags)
{
    switch (codec) {
    case C_BC4: return 1;
    case C_BC5: return 2;
    case C_BC6H: return (flags & HKBCN_FLAG_DST_F32) ? 12 : 6;
    default: return 4;
    }
}
uint32_t tile_blocks(int codec)
{
    switch (codec) {
    case C_BC4: return bc45_tile_blocks(1);
    case C_BC5: return bc45_tile_blocks(2);
    case C_BC1: case C_BC2: case C_BC3: return bc123_tile_blocks(codec);
    case C_BC7: return bc7_tile_blocks();
    case C_BC6H: return bc6h_tile_blocks();
    }
    return 64;
}

int ensure_init()
{
    if (g_init.load(std::memory_order_acquire)) return HKBCN_OK;
    return hkbcn_init(0);
}

// the stream's own {next tile, finished CTAs} pair (see Device::sched_of_stream); nullptr + error text on failure
uint32_t* sched_slot(Device* D, cudaStream_t st)
{
    std::lock_guard<std::mutex> lk(D->sched_mu);
    auto it = D->sched_of_stream.find(st);
    if (it != D->sched_of_stream.end()) return it->second;
    if (D->sched_used % Device::SCHED_CHUNK == 0) {
        uint32_t* c = nullptr;
        // cudaMalloc / cudaMemset act on the CURRENT device: callers have done cudaSetDevice(D->id)
        if (cudaMalloc((void**)&c, Device::SCHED_CHUNK * 32) != cudaSuccess || cudaMemset(c, 0, Device::SCHED_CHUNK * 32) != cudaSuccess) {
            set_err("cannot allocate tile counters for a new stream: %s", cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        D->sched_chunks.push_back(c);
    }
    uint32_t* p = D->sched_chunks.back() + 8 * (D->sched_used % Device::SCHED_CHUNK); // one 32-byte sector each
    D->sched_used++;
    D->sched_of_stream[st] = p;
    return p;
}

// the stream's BC7 work-list scratch, at least `bytes` large, first 16 bytes zero; nullptr + error text on failure
uint32_t* bc7_aux(Device* D, cudaStream_t st, size_t bytes)
{
    std::lock_guard<std::mutex> lk(D->sched_mu);
    Device::Aux& a = D->aux_of_stream[st];
    if (a.cap >= bytes) return a.p;
    const size_t want = std::max(bytes + bytes / 2, (size_t)1 << 20);
    uint32_t* p = nullptr;
    if (cudaMalloc((void**)&p, want) != cudaSuccess || cudaMemset(p, 0, 16) != cudaSuccess) {
        set_err("cannot allocate %zu bytes of BC7 work-list scratch: %s", want, cudaGetErrorString(cudaGetLastError()));
        if (p) cudaFree(p);
        return nullptr;
    }
    if (a.p) D->aux_retired.push_back(a.p);
    a.p = p;
    a.cap = want;
    return p;
}

Device* dev_at(int i) { return (i >= 0 && i < (int)g_devs.size()) ? g_devs[i] : nullptr; }

int grow(void** p, size_t* cap, size_t need)
{
    if (*cap >= need) return HKBCN_OK;
    if (*p) { cudaFree(*p); *p = nullptr; *cap = 0; }
    size_t want = std::max(need, (size_t)1 << 20);
    CK(cudaMalloc(p, want));
    *cap = want;
    return HKBCN_OK;
}

int grow_pinned(void** p, size_t* cap, size_t need)
{
    if (*cap >= need) return HKBCN_OK;
    if (*p) { cudaFreeHost(*p); *p = nullptr; *cap = 0; }
    size_t want = std::max(need, (size_t)1 << 20);
    CK(cudaHostAlloc(p, want, cudaHostAllocPortable));
    *cap = want;
    return HKBCN_OK;
}
bool is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// Append one surface to a launch descriptor.
// src_w: texels per source row when it differs from w (a surface whose Width is not a multiple of 4: the reference
// truncates the block count, numBlocksX = Width / 4, but keeps rowStride = Width * bpp, ImageEncoders.cpp:224-226)
void add_level(EncodeParams& P, int codec, uint32_t flags, const void* src, void* dst, uint32_t w, uint32_t h, uint32_t src_w = 0)
{
    Level& L = P.lv[P.nLevels++];
    const uint32_t bpp = src_bpp(codec, flags);
    const uint32_t tb = tile_blocks(codec);
    L.src = (const uint8_t*)src;
    L.dst = (uint8_t*)dst;
    L.nbx = w / 4;
    L.nblocks = (w / 4) * (h / 4);
    L.rowStride = (src_w ? src_w : w) * bpp;
    L.tileBase = P.totalTiles;
    // TMA bulk path: base 16-B aligned and every block-row segment a multiple of 16 B
    L.bulkOk = (((size_t)src & 15) == 0 && ((L.nbx * 4 * bpp) & 15) == 0 && (L.rowStride & 15) == 0) ? 1u : 0u;
    L.pad = 0;
    P.totalTiles += (L.nblocks + tb - 1) / tb;
}

// the encoders store whole blocks (8 / 16 bytes at once): the destination of a *_dev call must be block-aligned.  Sources
// may sit at any address (unaligned ones take the plain-load staging path instead of the bulk copy).
int check_dev_ptrs(int codec, const void* d_src, const void* d_dst)
{
    if (!d_src || !d_dst) { set_err("null device pointer"); return HKBCN_ERR_BAD_ARG; }
    if ((size_t)d_dst % (size_t)blk_bytes(codec)) { set_err("d_dst must be %d-byte aligned (one block)", blk_bytes(codec)); return HKBCN_ERR_BAD_ARG; }
    return HKBCN_OK;
}

int launch_encode(Device* D, int fmt, int codec, const EncodeParams& P0, cudaStream_t st)
{
    if (P0.totalTiles == 0) return HKBCN_OK;
    if (P0.flags & HKBCN_FLAG_LEVEL_MASK) {
        const int level = (int)((P0.flags & HKBCN_FLAG_LEVEL_MASK) >> 4) - 1;
        const bool ok = (codec == C_BC1 || codec == C_BC2 || codec == C_BC3) ? level <= 18 : codec == C_BC7 ? level <= 4 : false;
        if (!ok) {
            set_err("quality level %d of format %d does not exist (BC1/2/3: 0..18; BC7: 0..4; see hkbcn.h)", level, fmt);
            return HKBCN_ERR_UNSUPPORTED;
        }
    }
    if ((codec == C_BC1 || codec == C_BC2 || codec == C_BC3) && bc123_needs_hq(codec, P0.flags) && !bc1_hq_available()) {
        set_err("BC1/2/3 levels 6..18 and the 3-colour options need rgbcx's total-ordering tables; this libhkbcn.so was built without them (csrc/Makefile: RGBCX_DIR)");
        return HKBCN_ERR_UNSUPPORTED;
    }
    EncodeParams P = P0;
    if (!(P.sched = sched_slot(D, st))) return HKBCN_ERR_CUDA;
    if (codec == C_BC7) {
        uint32_t nb = 0;
        for (uint32_t i = 0; i < P.nLevels; i++) {
            // a work-list entry is level << 28 | block: 2^28 blocks = a 65536 x 65536 level
            if (P.lv[i].nblocks >= (1u << 28)) { set_err("BC7: a level of %u blocks is beyond the 2^28 the work list addresses", P.lv[i].nblocks); return HKBCN_ERR_BAD_ARG; }
            nb += P.lv[i].nblocks;
        }
        if (!(P.aux = bc7_aux(D, st, bc7_aux_bytes(nb)))) return HKBCN_ERR_CUDA;
    }
    cudaError_t e = cudaSuccess;
    switch (codec) {
    case C_BC4: e = launch_bc45_encode(1, P, D->sm_count, st); break;
    case C_BC5: e = launch_bc45_encode(2, P, D->sm_count, st); break;
    case C_BC1: case C_BC2: case C_BC3: e = launch_bc123_encode(codec, P, D->sm_count, st); break;
    case C_BC7: e = launch_bc7_encode(P, D->sm_count, st); break;
    case C_BC6H: e = launch_bc6h_encode(fmt == HKBCN_BC6H_SFLOAT, (P.flags & HKBCN_FLAG_SRC_F16) != 0, P, D->sm_count, st); break;
    default: set_err("bad codec"); return HKBCN_ERR_BAD_FORMAT;
    }
    if (e == cudaErrorNotSupported) { set_err("encoder for format %d not built yet", fmt); return HKBCN_ERR_UNSUPPORTED; }
    if (e != cudaSuccess) { set_err("kernel launch failed: %s", cudaGetErrorString(e)); return HKBCN_ERR_CUDA; }
    g_launches.fetch_add(codec == C_BC7 ? 2 : 1, std::memory_order_relaxed); // BC7: classify + encode
    return HKBCN_OK;
}

// level i of a chain: max(w>>i,4) x max(h>>i,4) texels (compressed dims clamp to >= 4, Image.cpp:412-416)
inline uint32_t mip_dim(uint32_t d, uint32_t i) { uint32_t v = d >> i; return v < 4 ? 4 : v; }

// What ImageStorage::Reset accepts for block-compressed storage (Image.cpp:412-431): dimensions that are multiples of the
// block size; powers of two when the storage is mipmapped (else level i = dim >> i stops being a whole number of blocks and
// the reference refuses to load the result); at most CalcNumMips levels (the reference wants 1 or exactly that many -- a
// prefix of a valid chain is accepted here because every level of it is still whole blocks).
int check_chain_dims(uint32_t w, uint32_t h, uint32_t mips)
{
    if (!w || !h || (w & 3) || (h & 3)) { set_err("chain dims must be non-zero multiples of 4 (Image.cpp:412-413)"); return HKBCN_ERR_BAD_ARG; }
    if (mips == 0 || mips > HKBCN_MAX_LEVELS) { set_err("num_mips %u out of range 1..%d", mips, HKBCN_MAX_LEVELS); return HKBCN_ERR_BAD_ARG; }
    if (mips > 1) {
        if ((w & (w - 1)) || (h & (h - 1))) { set_err("mipmapped compressed storage needs power-of-two dimensions (Image.cpp:415-416), got %u x %u", w, h); return HKBCN_ERR_BAD_ARG; }
        if (mips > hkbcn_calc_num_mips(w, h)) { set_err("num_mips %u exceeds CalcNumMips(%u, %u) = %u", mips, w, h, hkbcn_calc_num_mips(w, h)); return HKBCN_ERR_BAD_ARG; }
    }
    return HKBCN_OK;
}

int build_chain(EncodeParams& P, int codec, uint32_t flags, const void* src, void* dst, uint32_t w, uint32_t h, uint32_t mips,
                uint32_t slices = 1)
{
    // ImageStorage keeps all slices of a level together (Image.cpp:498-574): `slices` tightly packed w x h surfaces are one
    // w x (slices*h) surface whose row-major block order is slice 0's blocks, then slice 1's, ... -- no block straddles two
    // slices because h is a multiple of 4 -- so a sliced level needs no Level entry of its own.
    if (slices == 0 || (uint64_t)mip_dim(h, 0) * slices > 0x40000000ull) { set_err("slice count %u out of range", slices); return HKBCN_ERR_BAD_ARG; }
    if (mips == 0 || mips > HKBCN_MAX_LEVELS) { set_err("num_mips %u out of range 1..%d", mips, HKBCN_MAX_LEVELS); return HKBCN_ERR_BAD_ARG; }
    memset(&P, 0, sizeof(P));
    P.flags = flags;
    const uint8_t* s = (const uint8_t*)src;
    uint8_t* d = (uint8_t*)dst;
    for (uint32_t i = 0; i < mips; i++) {
        uint32_t lw = mip_dim(w, i), lh = mip_dim(h, i) * slices;
        add_level(P, codec, flags, s, d, lw, lh);
        s += (size_t)lw * lh * src_bpp(codec, flags);
        d += (size_t)(lw / 4) * (lh / 4) * blk_bytes(codec);
    }
    return HKBCN_OK;
}

size_t chain_src_bytes(int codec, uint32_t flags, uint32_t w, uint32_t h, uint32_t mips, uint32_t slices = 1)
{
    size_t n = 0;
    for (uint32_t i = 0; i < mips; i++) n += (size_t)mip_dim(w, i) * mip_dim(h, i) * src_bpp(codec, flags) * slices;
    return n;
}

// the device cascade (generate_mips_on, defined with the mip entry points below), handed in by the import form
typedef int (*MipCascadeFn)(Device* D, int fmt, void* d_chain, uint32_t w, uint32_t h, uint32_t mips, const hkbcn_mip_config* cfg, uint32_t flags,
                            cudaStream_t st, uint32_t slices);

// Host-pointer compress of a large surface or chain, cut into pieces -- bands of block rows of level 0, then levels 1..
// together -- each uploaded on a copy stream and encoded by its own launch as soon as it has arrived, so the upload of piece
// i + 1 runs under the kernel of piece i and the blocks of piece i go back under the kernel of piece i + 1.
//   pinned source / destination: the copy engine reads and writes the caller's buffers directly;
//   PAGEABLE ones: the copy engine cannot touch them -- a plain cudaMemcpyAsync is staged by the driver and serialises with
//     everything else (27.6 ms for the 4096^2 BC7 chain against 21.0 ms from pinned memory) -- so this thread copies each
//     piece through one of two pinned buffers (outputs through a third), still under the kernels;
//   gen (the import form: src holds level 0 only): after the last band the device cascade builds levels 1.. in place and the
//     last piece encodes them -- nothing but level 0 crosses PCIe.
// Same kernels, same bytes out.
int compress_host_pipelined(Device* D, int fmt, int codec, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, uint32_t mips, uint32_t flags,
                            MipCascadeFn gen = nullptr, const hkbcn_mip_config* gen_cfg = nullptr)
{
    const uint32_t bpp = src_bpp(codec, flags), bb = blk_bytes(codec);
    struct Piece { size_t soff, sbytes, doff, dbytes; uint32_t lvl0, lvl1, row0, rows; };
    std::vector<Piece> pieces;
    const uint32_t nby = h / 4;
    const size_t l0 = (size_t)w * h * bpp;
    // the slow encoders (BC7, BC6H) want pieces of 8 MB or more (their tile rounds, below); for the fast ones the copies are
    // the larger share of the call and a launch costs little: pieces down to 2 MB
    const bool slow = codec == C_BC7 || codec == C_BC6H;
    const size_t min_piece = slow ? ((size_t)8 << 20) : ((size_t)2 << 20);
    int nb = (int)std::min<size_t>(6, std::max<size_t>(1, l0 / (slow ? (12u << 20) : (3u << 20))));
    if (nb > 1) {
        // Every band is a launch of its own, and a persistent encoder works in rounds of one tile per SM: a band of 2.2 rounds
        // costs three (measured, BC7 4096^2 chain from pageable memory: 5 bands 25.5 ms, 4 bands 22.3, 6 bands 21.9).  Among
        // the band counts that keep a piece at 8 MB or more, take the one whose last round is fullest; ties go to more bands.
        const uint32_t tb = tile_blocks(codec), sms = (uint32_t)std::max(1, D->sm_count);
        double best = -1.0;
        int best_nb = nb;
        for (int c = 2; c <= 7; c++) {
            if (l0 / (size_t)c < min_piece || (uint32_t)c > nby) break;
            const uint64_t tiles = ((uint64_t)((nby + c - 1) / c) * (w / 4) + tb - 1) / tb;
            const double eff = (double)tiles / (double)(((tiles + sms - 1) / sms) * sms);
            if (eff >= best - 1e-9) { best = eff; best_nb = c; }
        }
        nb = best_nb;
    }
    if (const char* e = getenv("HKBCN_PIPE_BANDS")) nb = std::max(1, std::min(7, atoi(e))); // A/B timing
    if ((uint32_t)nb > nby) nb = (int)nby;
    // What nothing can overlap is the way in of the first piece and the way out of the last: with two streams the band size no
    // longer matters to the encoder, so the first band is cut once more (its first quarter goes up alone) and level 1 is a
    // piece of its own ahead of levels 2.., whose blocks are a sixteenth of the chain's.
    const bool fine = nb > 1 && !getenv("HKBCN_PIPE_COARSE");
    for (int b = 0; b < nb; b++) {
        const uint32_t r0 = (uint32_t)((uint64_t)nby * b / nb), r1 = (uint32_t)((uint64_t)nby * (b + 1) / nb);
        const uint32_t cut = (b == 0 && fine && r1 - r0 >= 8) ? r0 + (r1 - r0) / 4 : r0;
        if (cut > r0) pieces.push_back({(size_t)r0 * 4 * w * bpp, (size_t)(cut - r0) * 4 * w * bpp, (size_t)r0 * (w / 4) * bb, (size_t)(cut - r0) * (w / 4) * bb, 0, 1, r0, cut - r0});
        pieces.push_back({(size_t)cut * 4 * w * bpp, (size_t)(r1 - cut) * 4 * w * bpp, (size_t)cut * (w / 4) * bb, (size_t)(r1 - cut) * (w / 4) * bb, 0, 1, cut, r1 - cut});
    }
    if (mips > 1) {
        const size_t send = chain_src_bytes(codec, flags, w, h, mips), dend = hkbcn_chain_bytes(fmt, w, h, 1, mips), d0 = (size_t)(w / 4) * nby * bb;
        if (fine && mips > 2 && !gen) {
            const uint32_t w1 = mip_dim(w, 1), h1 = mip_dim(h, 1);
            const size_t s1 = (size_t)w1 * h1 * bpp, d1 = (size_t)(w1 / 4) * (h1 / 4) * bb;
            pieces.push_back({l0, s1, d0, d1, 1, 2, 0, 0});
            pieces.push_back({l0 + s1, send - l0 - s1, d0 + d1, dend - d0 - d1, 2, mips, 0, 0});
        } else pieces.push_back({l0, send - l0, d0, dend - d0, 1, mips, 0, 0});
    }
    if (pieces.size() > 12) { set_err("internal: too many pieces"); return HKBCN_ERR_BAD_ARG; }
    const size_t sbytes = chain_src_bytes(codec, flags, w, h, mips), dbytes = hkbcn_chain_bytes(fmt, w, h, 1, mips);
    size_t maxp = 0;
    for (auto& p : pieces)
        if (!(gen && p.lvl0 > 0)) maxp = std::max(maxp, p.sbytes);
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    const bool src_pinned = is_pinned(src), dst_pinned = is_pinned(dst);
    int rc;
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    for (int i = 0; i < 2; i++) {
        if (!src_pinned && (rc = grow_pinned(&D->h_in[i], &D->h_in_cap[i], maxp)) != HKBCN_OK) return rc;
        if (!D->ev_in[i]) CK(cudaEventCreateWithFlags(&D->ev_in[i], cudaEventDisableTiming));
    }
    if (!dst_pinned && (rc = grow_pinned(&D->h_out, &D->h_out_cap, dbytes)) != HKBCN_OK) return rc;
    if (!D->copy_stream) CK(cudaStreamCreateWithFlags(&D->copy_stream, cudaStreamNonBlocking));
    if (!D->stream2) CK(cudaStreamCreateWithFlags(&D->stream2, cudaStreamNonBlocking));
    // Consecutive pieces are launched on two alternating streams: a persistent encoder ends with a round in which SMs run
    // out of tiles one by one, and on one stream the next piece's kernel could not start until the last of them had finished
    // (six bands of the BC7 4096^2 level: ~1.5 ms of idle SMs in total); on the other stream its CTAs take over each SM as
    // it frees up.  Tile counters and the BC7 work list are per stream (sched_slot, bc7_aux).
    cudaStream_t es[2] = {D->stream, D->stream2};
    if (getenv("HKBCN_PIPE_ONE_STREAM")) es[1] = D->stream; // A/B timing
    for (size_t i = 0; i < pieces.size(); i++)
        if (!D->ev_piece[i]) CK(cudaEventCreateWithFlags(&D->ev_piece[i], cudaEventDisableTiming));
    uint8_t* const out = dst_pinned ? dst : (uint8_t*)D->h_out;
    for (size_t i = 0; i < pieces.size(); i++) {
        const Piece& p = pieces[i];
        const int slot = (int)(i & 1);
        cudaStream_t st = es[i & 1];
        if (gen && p.lvl0 > 0) {
            // every band is on the device once the copy stream has passed its last upload: build levels 1.. from level 0 there
            CK(cudaStreamWaitEvent(st, D->ev_in[0], 0));
            CK(cudaStreamWaitEvent(st, D->ev_in[1], 0));
            if ((rc = gen(D, fmt, D->d_src, w, h, mips, gen_cfg, flags, st, 1)) != HKBCN_OK) return rc;
        } else {
            const uint8_t* from = src + p.soff;
            if (!src_pinned) {
                if (i >= 2) CK(cudaEventSynchronize(D->ev_in[slot])); // the slot's previous upload has left it
                memcpy(D->h_in[slot], src + p.soff, p.sbytes);
                from = (const uint8_t*)D->h_in[slot];
            }
            CK(cudaMemcpyAsync((uint8_t*)D->d_src + p.soff, from, p.sbytes, cudaMemcpyHostToDevice, D->copy_stream));
            CK(cudaEventRecord(D->ev_in[slot], D->copy_stream));
            CK(cudaStreamWaitEvent(st, D->ev_in[slot], 0));
        }
        EncodeParams P;
        memset(&P, 0, sizeof(P));
        P.flags = flags;
        if (p.lvl0 == 0) {
            add_level(P, codec, flags, (uint8_t*)D->d_src + p.soff, (uint8_t*)D->d_dst + p.doff, w, p.rows * 4);
        } else {
            size_t so = p.soff, dof = p.doff;
            for (uint32_t l = p.lvl0; l < p.lvl1; l++) {
                const uint32_t lw = mip_dim(w, l), lh = mip_dim(h, l);
                add_level(P, codec, flags, (uint8_t*)D->d_src + so, (uint8_t*)D->d_dst + dof, lw, lh);
                so += (size_t)lw * lh * bpp;
                dof += (size_t)(lw / 4) * (lh / 4) * bb;
            }
        }
        if ((rc = launch_encode(D, fmt, codec, P, st)) != HKBCN_OK) return rc;
        CK(cudaMemcpyAsync(out + p.doff, (uint8_t*)D->d_dst + p.doff, p.dbytes, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(D->ev_piece[i], st));
    }
    for (size_t i = 0; i < pieces.size(); i++) {
        CK(cudaEventSynchronize(D->ev_piece[i]));
        if (!dst_pinned) memcpy(dst + pieces[i].doff, (uint8_t*)D->h_out + pieces[i].doff, pieces[i].dbytes);
    }
    return HKBCN_OK;
}

int compress_host_on(Device* D, int fmt, const void* src, void* dst, uint32_t w, uint32_t h, uint32_t mips, uint32_t flags,
                     uint32_t slices = 1)
{
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (!src || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    // Width % 4 != 0 (single surfaces only): rows keep their full length, the last Width % 4 texels of each are never read
    const bool ragged = (w & 3) != 0;
    if (ragged && (mips != 1 || slices != 1)) { set_err("chains and sliced storage need widths that are multiples of 4 (Image.cpp:412-413)"); return HKBCN_ERR_BAD_ARG; }
    const size_t sbytes = ragged ? (size_t)w * h * src_bpp(codec, flags) : chain_src_bytes(codec, flags, w, h, mips, slices);
    const size_t dbytes = hkbcn_chain_bytes(fmt, w, h, slices, mips);
    if (dbytes == 0) return HKBCN_OK;
    if (!ragged && slices == 1 && sbytes >= (16u << 20) && !getenv("HKBCN_NO_PIPELINE")) {
        CK(cudaSetDevice(D->id));
        // pageable callers always; pinned ones too unless HKBCN_PIPE_PINNED=0 (A/B timing of the blocking form)
        const char* pp = getenv("HKBCN_PIPE_PINNED");
        if (!is_pinned(src) || !(pp && pp[0] == '0')) return compress_host_pipelined(D, fmt, codec, (const uint8_t*)src, (uint8_t*)dst, w, h, mips, flags);
    }
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    int rc;
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src, sbytes, cudaMemcpyHostToDevice, D->stream));
    EncodeParams P;
    if (ragged) {
        memset(&P, 0, sizeof(P));
        P.flags = flags;
        add_level(P, codec, flags, D->d_src, D->d_dst, w & ~3u, h, w);
    } else if ((rc = build_chain(P, codec, flags, D->d_src, D->d_dst, w, h, mips, slices)) != HKBCN_OK) return rc;
    if ((rc = launch_encode(D, fmt, codec, P, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(dst, D->d_dst, dbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}


} // namespace

extern "C" {

const char* hkbcn_last_error(void) { return t_err; }

float hkbcn_srgb_decode_entry(int v)
{
    if (v < 0 || v > 255) return 0.0f;
    const float f = (float)v / 255.0f;
    const float x = f <= 0.04045f ? f / 12.92f : powf((f + 0.055f) / 1.055f, 2.4f);
    char buf[32];
    snprintf(buf, sizeof(buf), "%.6f", (double)x);
    return strtof(buf, nullptr);
}

int hkbcn_build_has(int feature)
{
    switch (feature) {
    case HKBCN_BUILD_BC1_ORDERINGS: return bc1_hq_available() ? 1 : 0;
    case HKBCN_BUILD_STBIR_SRGB: return mip_srgb_is_stbir() ? 1 : 0;
    default: return 0;
    }
}
const char* hkbcn_version(void) { return "hkbcn 0.1 (sm_100a)"; }

static void destroy_device(Device* D)
{
    cudaSetDevice(D->id);
    if (D->d_src) cudaFree(D->d_src);
    if (D->d_dst) cudaFree(D->d_dst);
    if (D->lut_srgb) cudaFree(D->lut_srgb);
    if (D->lut_lin) cudaFree(D->lut_lin);
    for (int i = 0; i < 2; i++) {
        if (D->h_in[i]) cudaFreeHost(D->h_in[i]);
        if (D->ev_in[i]) cudaEventDestroy(D->ev_in[i]);
    }
    if (D->h_out) cudaFreeHost(D->h_out);
    for (cudaEvent_t e : D->ev_piece) if (e) cudaEventDestroy(e);
    if (D->copy_stream) cudaStreamDestroy(D->copy_stream);
    if (D->stream2) cudaStreamDestroy(D->stream2);
    for (uint32_t* c : D->sched_chunks) cudaFree(c);
    for (auto& kv : D->aux_of_stream) cudaFree(kv.second.p);
    for (uint32_t* c : D->aux_retired) cudaFree(c);
    for (auto& sl : D->slots) {
        if (sl.d_src) cudaFree(sl.d_src);
        if (sl.d_dst) cudaFree(sl.d_dst);
        if (sl.h_src) cudaFreeHost(sl.h_src);
        if (sl.h_dst) cudaFreeHost(sl.h_dst);
        if (sl.st) cudaStreamDestroy(sl.st);
    }
    for (auto& kv : D->taps) { cudaFree(kv.second.idx); cudaFree(kv.second.w); }
    D->taps.clear();
    if (D->stream) cudaStreamDestroy(D->stream);
    delete D;
}

// one device's state; on failure everything built so far is released and *out stays null
static int create_device(int i, const cudaDeviceProp& prop, Device** out)
{
    Device* D = new Device();
    D->id = i;
    D->sm_count = prop.multiProcessorCount;
    struct Guard { Device* d; ~Guard() { if (d) destroy_device(d); } } guard{D};
    CK(cudaSetDevice(i));
    CK(cudaStreamCreateWithFlags(&D->stream, cudaStreamNonBlocking));
    // a12: lookup tables consumed by the encoders / BC6H decoder
    CK(decode_upload_tables());
    CK(bc1_upload_tables());
    cudaError_t te = bc7_upload_tables();
    if (te != cudaSuccess && te != cudaErrorNotSupported) { set_err("bc7 table upload: %s", cudaGetErrorString(te)); return HKBCN_ERR_CUDA; }
    CK(bc6h_upload_tables());
    // stbir's uint8 decode: v / 255, and sRGB -> linear through stbir__srgb_uchar_to_linear_float (stb_image_resize.h:607-633):
    // 256 six-decimal literals, which are what the transfer function gives in FLOAT arithmetic printed with %.6f -- recomputed
    // here the same way (tests/test_stbir_tables_cpu.py holds the result to the header's literals)
    float srgb[256], lin[256];
    for (int v = 0; v < 256; v++) {
        srgb[v] = hkbcn_srgb_decode_entry(v);
        lin[v] = (float)v / 255.0f;
    }
    CK(cudaMalloc((void**)&D->lut_srgb, sizeof(srgb)));
    CK(cudaMalloc((void**)&D->lut_lin, sizeof(lin)));
    CK(cudaMemcpy(D->lut_srgb, srgb, sizeof(srgb), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(D->lut_lin, lin, sizeof(lin), cudaMemcpyHostToDevice));
    guard.d = nullptr;
    *out = D;
    return HKBCN_OK;
}

int hkbcn_init(uint32_t device_mask)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_init.load(std::memory_order_acquire)) return HKBCN_OK;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_err("no CUDA device available (%s); this library has no CPU fallback", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
        return HKBCN_ERR_NO_DEVICE;
    }
    // the list is built aside and published together with g_init only when every selected device came up: a failure half way
    // leaves no half-built state behind, and a later call starts from scratch instead of appending the same GPUs again
    std::vector<Device*> devs;
    int rc = HKBCN_OK;
    for (int i = 0; i < n && rc == HKBCN_OK; i++) {
        if (device_mask && !(device_mask & (1u << i))) continue;
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, i) != cudaSuccess) { set_err("cudaGetDeviceProperties(%d) failed", i); rc = HKBCN_ERR_CUDA; break; }
        if (prop.major != 10) {
            set_err("device %d (%s, sm_%d%d) is not sm_100; kernels are built for sm_100a only", i, prop.name, prop.major, prop.minor);
            continue;
        }
        Device* D = nullptr;
        rc = create_device(i, prop, &D);
        if (rc == HKBCN_OK) devs.push_back(D);
    }
    if (rc == HKBCN_OK && devs.empty()) {
        if (!t_err[0]) set_err("no sm_100 device selected by mask 0x%x", device_mask);
        rc = HKBCN_ERR_NO_DEVICE;
    }
    if (rc != HKBCN_OK) {
        for (Device* D : devs) destroy_device(D);
        return rc;
    }
    CK(cudaSetDevice(devs[0]->id));
    g_devs = std::move(devs);
    g_init.store(true, std::memory_order_release);
    return HKBCN_OK;
}

void hkbcn_shutdown(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    g_init.store(false, std::memory_order_release);
    for (Device* D : g_devs) destroy_device(D);
    g_devs.clear();
}

int hkbcn_device_count(void) { return g_init.load(std::memory_order_acquire) ? (int)g_devs.size() : 0; }
uint64_t hkbcn_launch_count(void) { return g_launches.load(); }

int hkbcn_block_bytes(int fmt) { int c = codec_of(fmt); return c < 0 ? HKBCN_ERR_BAD_FORMAT : blk_bytes(c); }
int hkbcn_src_texel_bytes(int fmt, uint32_t flags) { int c = codec_of(fmt); return c < 0 ? HKBCN_ERR_BAD_FORMAT : src_bpp(c, flags); }
int hkbcn_dst_texel_bytes(int fmt, uint32_t flags) { int c = codec_of(fmt); return c < 0 ? HKBCN_ERR_BAD_FORMAT : dst_bpp(c, flags); }

// Image.cpp:188-216 for block-compressed formats: log2(max(w,h)/4) + 1 (dims < 4 count as 4)
uint32_t hkbcn_calc_num_mips(uint32_t width, uint32_t height)
{
    uint32_t m = std::max(width, height) / 4;
    uint32_t n = 1;
    while (m > 1) { m >>= 1; n++; }
    return n;
}

size_t hkbcn_chain_bytes(int fmt, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips)
{
    int c = codec_of(fmt);
    if (c < 0) return 0;
    size_t n = 0;
    for (uint32_t i = 0; i < num_mips; i++) n += (size_t)(mip_dim(width, i) / 4) * (mip_dim(height, i) / 4) * blk_bytes(c) * slices;
    return n;
}

// ---- one large surface over several GPUs: horizontal bands of whole block rows -------------------------
// Blocks are independent and a band of block rows [r0, r1) of a tightly packed level is one contiguous range of the source
// (texel rows 4*r0 .. 4*r1) and of the destination (blocks r0*nbx .. r1*nbx), so GPU d takes band d of EVERY level, encodes
// its pieces in one launch and copies each piece to its final offset of the caller's buffer ("host gather").  No exchange.
int hkbcn_plan_bands(uint32_t height, uint32_t num_mips, int n_gpus, int device, uint32_t* row0, uint32_t* row1)
{
    if (n_gpus <= 0 || device < 0 || device >= n_gpus || !row0 || !row1) { set_err("bad arguments"); return HKBCN_ERR_BAD_ARG; }
    for (uint32_t i = 0; i < num_mips; i++) {
        const uint64_t nby = mip_dim(height, i) / 4;
        row0[i] = (uint32_t)(nby * (uint64_t)device / (uint64_t)n_gpus);
        row1[i] = (uint32_t)(nby * (uint64_t)(device + 1) / (uint64_t)n_gpus);
    }
    return HKBCN_OK;
}

static int compress_bands_on(Device* D, int d, int nd, int fmt, int codec, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h,
                             uint32_t mips, uint32_t flags)
{
    uint32_t r0[HKBCN_MAX_LEVELS], r1[HKBCN_MAX_LEVELS];
    int rc = hkbcn_plan_bands(h, mips, nd, d, r0, r1);
    if (rc) return rc;
    const uint32_t bpp = src_bpp(codec, flags), bb = blk_bytes(codec);
    size_t sbytes = 0, dbytes = 0;
    for (uint32_t i = 0; i < mips; i++) {
        const size_t rows = r1[i] - r0[i];
        sbytes += rows * 4 * mip_dim(w, i) * bpp;
        dbytes += rows * (mip_dim(w, i) / 4) * bb;
    }
    if (dbytes == 0) return HKBCN_OK;
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    EncodeParams P;
    memset(&P, 0, sizeof(P));
    P.flags = flags;
    size_t so = 0, dof = 0, ls = 0, ld = 0; // offsets in the device staging / in the caller's chain
    for (uint32_t i = 0; i < mips; i++) {
        const uint32_t lw = mip_dim(w, i), lh = mip_dim(h, i);
        const size_t rows = r1[i] - r0[i];
        if (rows) {
            const size_t sb = rows * 4 * lw * bpp, db = rows * (lw / 4) * bb;
            CK(cudaMemcpyAsync((uint8_t*)D->d_src + so, src + ls + (size_t)r0[i] * 4 * lw * bpp, sb, cudaMemcpyHostToDevice, D->stream));
            add_level(P, codec, flags, (uint8_t*)D->d_src + so, (uint8_t*)D->d_dst + dof, lw, (uint32_t)rows * 4);
            so += sb;
            dof += db;
        }
        ls += (size_t)lw * lh * bpp;
        ld += (size_t)(lw / 4) * (lh / 4) * bb;
    }
    if ((rc = launch_encode(D, fmt, codec, P, D->stream)) != HKBCN_OK) return rc;
    dof = 0;
    ld = 0;
    for (uint32_t i = 0; i < mips; i++) {
        const uint32_t lw = mip_dim(w, i), lh = mip_dim(h, i);
        const size_t rows = r1[i] - r0[i];
        if (rows) {
            const size_t db = rows * (lw / 4) * bb;
            CK(cudaMemcpyAsync(dst + ld + (size_t)r0[i] * (lw / 4) * bb, (uint8_t*)D->d_dst + dof, db, cudaMemcpyDeviceToHost, D->stream));
            dof += db;
        }
        ld += (size_t)(lw / 4) * (lh / 4) * bb;
    }
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_compress_chain_multi(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t num_mips, uint32_t flags,
                               int n_gpus)
{
    Range nvtx_("hkbcn_compress_chain_multi");
    int rc = ensure_init();
    if (rc) return rc;
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if (!src || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    const int avail = (int)g_devs.size();
    // n_gpus > 0: that many of the selected GPUs; 0: all of them; < 0: -n_gpus bands dealt round-robin over the selected GPUs
    // (more bands than devices -- lets a one-GPU box run the banded path)
    if (n_gpus < -64) { set_err("at most 64 bands"); return HKBCN_ERR_BAD_ARG; }
    const int nd = n_gpus < 0 ? -n_gpus : (n_gpus == 0 || n_gpus > avail ? avail : n_gpus);
    if (nd == 1) return compress_host_on(g_devs[0], fmt, src, dst, width, height, num_mips, flags);
    std::vector<std::thread> th;
    std::vector<int> err(nd, 0);
    std::vector<std::string> msg(nd);
    for (int d = 0; d < nd; d++)
        th.emplace_back([&, d]() {
            err[d] = compress_bands_on(g_devs[(size_t)d % (size_t)avail], d, nd, fmt, codec, (const uint8_t*)src, (uint8_t*)dst, width, height, num_mips, flags);
            if (err[d]) msg[d] = t_err;
        });
    for (auto& t : th) t.join();
    for (int d = 0; d < nd; d++)
        if (err[d]) { set_err("%s", msg[d].c_str()); return err[d]; }
    return HKBCN_OK;
}

// a surface this large is worth splitting over the selected GPUs (each has its own PCIe link and its own SMs)
static bool worth_banding(uint32_t w, uint32_t h) { return g_devs.size() > 1 && (uint64_t)w * h >= (4u << 20); }

int hkbcn_compress(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t flags)
{
    Range nvtx_("hkbcn_compress");
    int rc = ensure_init();
    if (rc) return rc;
    if (width / 4 == 0 || height / 4 == 0) return HKBCN_OK; // reference: zero blocks, nothing written
    if (!(width & 3) && worth_banding(width, height)) return hkbcn_compress_chain_multi(fmt, src, dst, width, height & ~3u, 1, flags, 0);
    return compress_host_on(g_devs[0], fmt, src, dst, width, height & ~3u, 1, flags);
}

int hkbcn_compress_chain(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t num_mips, uint32_t flags)
{
    Range nvtx_("hkbcn_compress_chain");
    int rc = ensure_init();
    if (rc) return rc;
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if (worth_banding(width, height)) return hkbcn_compress_chain_multi(fmt, src, dst, width, height, num_mips, flags, 0);
    return compress_host_on(g_devs[0], fmt, src, dst, width, height, num_mips, flags);
}


// ---- mip cascade (Image.cpp:1755-1803 / :1852-1899) ----------------------------------------------
static int get_taps(Device* D, int in_size, int out_size, int filter, int edge, Device::Taps& out)
{
    const uint64_t key = ((uint64_t)in_size << 40) | ((uint64_t)out_size << 16) | ((uint64_t)filter << 8) | (uint64_t)edge;
    std::lock_guard<std::mutex> lk(D->taps_mu);
    auto it = D->taps.find(key);
    if (it != D->taps.end()) { out = it->second; return HKBCN_OK; }
    AxisTaps a;
    build_axis_taps(in_size, out_size, filter, edge, a);
    Device::Taps t{nullptr, nullptr, a.T, nullptr, nullptr, nullptr, a.margin, 1, a.uniform ? 1 : 0, a.first.empty() ? 0 : a.first[0], {0, 0, 0, 0, 0, 0, 0, 0}};
    for (int k = 0; k < a.T && k < 8; k++) t.u[k] = a.w[(size_t)k];
    // the separable tile kernel stages the inputs of 32 consecutive outputs (x) / 8 (y): at most 72 / 24 of them
    for (int i = 0; i < out_size; i += 8) {
        const int j32 = std::min(out_size, i + 32) - 1, j8 = std::min(out_size, i + 8) - 1;
        if ((i % 32) == 0 && a.first[(size_t)j32] + a.cnt[(size_t)j32] - a.first[(size_t)i] > 72) t.spanOk = 0;
        if (a.first[(size_t)j8] + a.cnt[(size_t)j8] - a.first[(size_t)i] > 24) t.spanOk = 0;
    }
    const size_t no = (size_t)out_size;
    CK(cudaMalloc((void**)&t.idx, (a.idx.size() + 2 * no + a.map.size()) * sizeof(int)));
    CK(cudaMalloc((void**)&t.w, a.w.size() * sizeof(float)));
    t.first = t.idx + a.idx.size();
    t.cnt = t.first + no;
    t.map = t.cnt + no;
    CK(cudaMemcpy(t.idx, a.idx.data(), a.idx.size() * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(t.first, a.first.data(), no * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(t.cnt, a.cnt.data(), no * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(t.map, a.map.data(), a.map.size() * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(t.w, a.w.data(), a.w.size() * sizeof(float), cudaMemcpyHostToDevice));
    D->taps[key] = t;
    out = t;
    return HKBCN_OK;
}

static void set_tile_taps(ResizeParams& P, const Device::Taps& tx, const Device::Taps& ty)
{
    P.firstx = tx.first; P.cntx = tx.cnt; P.mapx = tx.map; P.marginx = tx.margin;
    P.firsty = ty.first; P.cnty = ty.cnt; P.mapy = ty.map; P.marginy = ty.margin;
    P.tileOk = (tx.spanOk && ty.spanOk && !getenv("HKBCN_MIP_GATHER")) ? 1 : 0; // HKBCN_MIP_GATHER: the one-thread-per-texel kernels (A/B timing)
    P.uniform = (tx.uniform && ty.uniform && !getenv("HKBCN_MIP_GENERAL")) ? 1 : 0;
    P.x0 = tx.first0;
    P.y0 = ty.first0;
    for (int k = 0; k < 8; k++) { P.ux[k] = tx.u[k]; P.uy[k] = ty.u[k]; }
    for (int k = 0; k < 16; k++) { P.ux2[k] = tx.u[k / 2]; P.uy2[k] = ty.u[k / 2]; }
    P.zero2[0] = P.zero2[1] = 0.0f;
}

// slices > 1: d_chain is in ImageStorage order (all slices of level 0, then of level 1, ...) and every slice runs its own cascade
static int generate_mips_on(Device* D, int fmt, void* d_chain, uint32_t w, uint32_t h, uint32_t mips, const hkbcn_mip_config* cfg,
                            uint32_t flags, cudaStream_t st, uint32_t slices = 1)
{
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (codec == C_BC6H && (flags & HKBCN_FLAG_SRC_F16)) { set_err("mip cascade of RGBA16F sources is not supported (the importers resample RGBA32F)"); return HKBCN_ERR_UNSUPPORTED; }
    { const int rc = check_chain_dims(w, h, mips); if (rc) return rc; }
    const int edge = cfg && cfg->edge_mode ? cfg->edge_mode : 3, filter = cfg && cfg->filter ? cfg->filter : 5;
    if (edge < 1 || edge > 4 || filter < 1 || filter > 5) { set_err("bad edge mode %d / filter %d", edge, filter); return HKBCN_ERR_BAD_ARG; }
    const uint32_t bpp = src_bpp(codec, flags);
    const bool f32 = codec == C_BC6H;
    const int channels = f32 ? 4 : (int)bpp;
    // TexFormat[].bHasAlpha (Image.cpp:105-118): BC1/2/3/7 carry alpha, BC4/5/6H do not
    const bool has_alpha = (codec == C_BC1 || codec == C_BC2 || codec == C_BC3 || codec == C_BC7) && !(cfg && cfg->no_alpha);
    const bool srgb = fmt == HKBCN_BC1_UNORM_SRGB || fmt == HKBCN_BC2_UNORM_SRGB || fmt == HKBCN_BC3_UNORM_SRGB || fmt == HKBCN_BC7_UNORM_SRGB;
    uint8_t* cur = (uint8_t*)d_chain;
    for (uint32_t i = 1; i < mips; i++) {
        const uint32_t iw = mip_dim(w, i - 1), ih = mip_dim(h, i - 1), ow = mip_dim(w, i), oh = mip_dim(h, i);
        uint8_t* next = cur + (size_t)iw * ih * bpp * slices;
        Device::Taps tx, ty;
        int rc;
        if ((rc = get_taps(D, (int)iw, (int)ow, filter, edge, tx)) != HKBCN_OK) return rc;
        if ((rc = get_taps(D, (int)ih, (int)oh, filter, edge, ty)) != HKBCN_OK) return rc;
        ResizeParams P;
        P.src = cur; P.dst = next;
        P.ix = tx.idx; P.wx = tx.w; P.iy = ty.idx; P.wy = ty.w;
        P.lut = srgb ? D->lut_srgb : D->lut_lin;
        P.in_w = (int)iw; P.in_h = (int)ih; P.out_w = (int)ow; P.out_h = (int)oh; P.Tx = tx.T; P.Ty = ty.T;
        P.alpha = has_alpha ? channels - 1 : -1;
        P.weight = (has_alpha && !(cfg && cfg->alpha_premultiplied)) ? 1 : 0;
        P.srgb = srgb ? 1 : 0;
        set_tile_taps(P, tx, ty);
        for (uint32_t sl = 0; sl < slices; sl++) {
            P.src = cur + (size_t)sl * iw * ih * bpp;
            P.dst = next + (size_t)sl * ow * oh * bpp;
            cudaError_t e = launch_mip_resize(P, channels, f32, st);
            if (e != cudaSuccess) { set_err("mip resize launch failed: %s", cudaGetErrorString(e)); return HKBCN_ERR_CUDA; }
            g_launches.fetch_add(1, std::memory_order_relaxed);
        }
        cur = next;
    }
    return HKBCN_OK;
}

int hkbcn_mip_taps(int in_size, int out_size, int filter, int edge_mode, int* idx, float* weight, int capacity, int* taps_per_output)
{
    if (in_size <= 0 || out_size <= 0 || filter < 1 || filter > 5 || edge_mode < 1 || edge_mode > 4 || !taps_per_output) {
        set_err("bad arguments");
        return HKBCN_ERR_BAD_ARG;
    }
    AxisTaps a;
    build_axis_taps(in_size, out_size, filter, edge_mode, a);
    std::vector<int>& i = a.idx;
    std::vector<float>& w = a.w;
    const int T = a.T;
    *taps_per_output = T;
    if ((size_t)capacity < i.size()) { set_err("capacity %d < %zu", capacity, i.size()); return HKBCN_ERR_BAD_ARG; }
    if (idx) memcpy(idx, i.data(), i.size() * sizeof(int));
    if (weight) memcpy(weight, w.data(), w.size() * sizeof(float));
    return HKBCN_OK;
}

size_t hkbcn_chain_src_bytes(int fmt, uint32_t width, uint32_t height, uint32_t num_mips, uint32_t flags)
{
    const int codec = codec_of(fmt);
    return codec < 0 ? 0 : chain_src_bytes(codec, flags, width, height, num_mips);
}

int hkbcn_generate_mips_dev(int device, int fmt, void* d_chain, uint32_t width, uint32_t height, uint32_t num_mips,
                            const hkbcn_mip_config* cfg, uint32_t flags, void* stream)
{
    Range nvtx_("hkbcn_generate_mips_dev");
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = dev_at(device);
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    if (!d_chain) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    CK(cudaSetDevice(D->id));
    return generate_mips_on(D, fmt, d_chain, width, height, num_mips, cfg, flags, (cudaStream_t)stream);
}

int hkbcn_generate_mips(int fmt, const void* src_level0, void* dst_chain, uint32_t width, uint32_t height, uint32_t num_mips,
                        const hkbcn_mip_config* cfg, uint32_t flags)
{
    Range nvtx_("hkbcn_generate_mips");
    int rc = ensure_init();
    if (rc) return rc;
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (!src_level0 || !dst_chain || !width || !height) { set_err("null pointer or empty surface"); return HKBCN_ERR_BAD_ARG; }
    Device* D = g_devs[0];
    const size_t sbytes = chain_src_bytes(codec, flags, width, height, num_mips);
    const size_t l0 = (size_t)width * height * src_bpp(codec, flags);
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src_level0, l0, cudaMemcpyHostToDevice, D->stream));
    if ((rc = generate_mips_on(D, fmt, D->d_src, width, height, num_mips, cfg, flags, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(dst_chain, D->d_src, sbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_import_chain_dev(int device, int fmt, void* d_chain, void* d_dst, uint32_t width, uint32_t height, uint32_t num_mips,
                           const hkbcn_mip_config* cfg, uint32_t flags, void* stream)
{
    int rc = hkbcn_generate_mips_dev(device, fmt, d_chain, width, height, num_mips, cfg, flags, stream);
    if (rc) return rc;
    return hkbcn_compress_chain_dev(device, fmt, d_chain, d_dst, width, height, num_mips, flags, stream);
}

// level 0 in host memory -> device mip cascade -> encode -> compressed chain in host memory, on one device
static int import_host_on(Device* D, int fmt, const void* src_level0, void* dst, uint32_t width, uint32_t height, uint32_t num_mips,
                   const hkbcn_mip_config* cfg, uint32_t flags)
{
    int rc;
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if (!src_level0 || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    const size_t sbytes = chain_src_bytes(codec, flags, width, height, num_mips);
    const size_t l0 = (size_t)width * height * src_bpp(codec, flags);
    const size_t dbytes = hkbcn_chain_bytes(fmt, width, height, 1, num_mips);
    if (l0 >= (16u << 20) && !getenv("HKBCN_NO_PIPELINE")) {
        // bands of level 0 uploaded under the kernels of the bands before them, then the cascade and levels 1..
        if (num_mips > 1) { // argument errors of the cascade before anything is launched
            const int edge = cfg && cfg->edge_mode ? cfg->edge_mode : 3, filter = cfg && cfg->filter ? cfg->filter : 5;
            if (edge < 1 || edge > 4 || filter < 1 || filter > 5) { set_err("bad edge mode %d / filter %d", edge, filter); return HKBCN_ERR_BAD_ARG; }
            if (codec == C_BC6H && (flags & HKBCN_FLAG_SRC_F16)) { set_err("mip cascade of RGBA16F sources is not supported (the importers resample RGBA32F)"); return HKBCN_ERR_UNSUPPORTED; }
        }
        return compress_host_pipelined(D, fmt, codec, (const uint8_t*)src_level0, (uint8_t*)dst, width, height, num_mips, flags, &generate_mips_on, cfg);
    }
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src_level0, l0, cudaMemcpyHostToDevice, D->stream)); // level 0 only crosses PCIe
    if (num_mips > 1 && (rc = generate_mips_on(D, fmt, D->d_src, width, height, num_mips, cfg, flags, D->stream)) != HKBCN_OK) return rc;
    EncodeParams P;
    if ((rc = build_chain(P, codec, flags, D->d_src, D->d_dst, width, height, num_mips)) != HKBCN_OK) return rc;
    if ((rc = launch_encode(D, fmt, codec, P, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(dst, D->d_dst, dbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}


int hkbcn_import_chain(int fmt, const void* src_level0, void* dst, uint32_t width, uint32_t height, uint32_t num_mips,
                       const hkbcn_mip_config* cfg, uint32_t flags)
{
    Range nvtx_("hkbcn_import_chain");
    int rc = ensure_init();
    if (rc) return rc;
    return import_host_on(g_devs[0], fmt, src_level0, dst, width, height, num_mips, cfg, flags);
}

// ---- CreateImage(RawImage ...) with compression (Image.cpp:866-1289) and CreateRoughnessMap (:2399-2456) -------------------------
static uint32_t closest_pow2(uint32_t v)
{   // Math::ToClosestPowerOfTwo (Hork/Core/BaseMath.h:417-423): the nearer of the two enclosing powers, ties go to the smaller
    if (v == 0) return 1;
    uint32_t hi = 1;
    while (hi < v) hi <<= 1;
    const uint32_t lo = hi >> 1;
    return (hi - v) < (v - lo) ? hi : lo;
}

int hkbcn_plan_image(int raw_format, uint32_t width, uint32_t height, int mipmapped, uint32_t storage_flags, uint32_t import_flags,
                     hkbcn_image_plan* plan)
{
    if (!plan) { set_err("null plan"); return HKBCN_ERR_BAD_ARG; }
    memset(plan, 0, sizeof(*plan));
    const bool srgb = (import_flags & HKBCN_IMPORT_ASSUME_8BIT_RGB_IMAGES_ARE_SRGB) != 0;
    const bool hdr_ok = (import_flags & HKBCN_IMPORT_USE_COMPRESSION) && (import_flags & HKBCN_IMPORT_ALLOW_HDRI_COMPRESSION);
    int fmt = 0, ch = 0;
    switch (raw_format) {
    case HKBCN_RAW_R8: ch = 1; fmt = HKBCN_BC4_UNORM; break;
    case HKBCN_RAW_R8_ALPHA: ch = 2; fmt = HKBCN_BC5_UNORM; break;
    case HKBCN_RAW_RGB8: case HKBCN_RAW_BGR8:
        ch = 3; plan->add_alpha = 1; storage_flags |= HKBCN_STORAGE_NO_ALPHA; // BC1: no alpha channel is used
        fmt = srgb ? HKBCN_BC1_UNORM_SRGB : HKBCN_BC1_UNORM;
        plan->swap_rb = raw_format == HKBCN_RAW_BGR8;
        break;
    case HKBCN_RAW_RGBA8: case HKBCN_RAW_BGRA8:
        ch = 4;
        if (storage_flags & HKBCN_STORAGE_NO_ALPHA) fmt = srgb ? HKBCN_BC1_UNORM_SRGB : HKBCN_BC1_UNORM;
        else fmt = srgb ? HKBCN_BC3_UNORM_SRGB : HKBCN_BC3_UNORM;
        plan->swap_rb = raw_format == HKBCN_RAW_BGRA8;
        break;
    case HKBCN_RAW_R32F: ch = 1; break;        // no compressed analogue
    case HKBCN_RAW_R32_ALPHA_F: ch = 2; break; // no compressed analogue
    case HKBCN_RAW_RGB32F: case HKBCN_RAW_BGR32F:
        ch = 3;
        plan->swap_rb = raw_format == HKBCN_RAW_BGR32F;
        if (hdr_ok) { fmt = HKBCN_BC6H_UFLOAT; plan->add_alpha = 1; storage_flags |= HKBCN_STORAGE_NO_ALPHA; }
        break;
    case HKBCN_RAW_RGBA32F: case HKBCN_RAW_BGRA32F:
        ch = 4;
        plan->swap_rb = raw_format == HKBCN_RAW_BGRA32F;
        if (hdr_ok) fmt = HKBCN_BC6H_UFLOAT; // the alpha channel is lost
        break;
    default: set_err("bad raw image format %d", raw_format); return HKBCN_ERR_BAD_FORMAT;
    }
    if (!(import_flags & HKBCN_IMPORT_USE_COMPRESSION)) fmt = 0;
    plan->format = fmt;
    plan->storage_flags = storage_flags;
    plan->src_channels = ch;
    // Image.cpp:1064-1068 (block size 4 for every BCn format)
    plan->required_width = mipmapped ? std::max<uint32_t>(closest_pow2(width), 4) : (width + 3) & ~3u;
    plan->required_height = mipmapped ? std::max<uint32_t>(closest_pow2(height), 4) : (height + 3) & ~3u;
    return HKBCN_OK;
}

static int raw_format_info(int raw_format, int* channels, bool* f32)
{
    switch (raw_format) {
    case HKBCN_RAW_R8: *channels = 1; *f32 = false; return HKBCN_OK;
    case HKBCN_RAW_R8_ALPHA: *channels = 2; *f32 = false; return HKBCN_OK;
    case HKBCN_RAW_RGB8: case HKBCN_RAW_BGR8: *channels = 3; *f32 = false; return HKBCN_OK;
    case HKBCN_RAW_RGBA8: case HKBCN_RAW_BGRA8: *channels = 4; *f32 = false; return HKBCN_OK;
    case HKBCN_RAW_R32F: *channels = 1; *f32 = true; return HKBCN_OK;
    case HKBCN_RAW_R32_ALPHA_F: *channels = 2; *f32 = true; return HKBCN_OK;
    case HKBCN_RAW_RGB32F: case HKBCN_RAW_BGR32F: *channels = 3; *f32 = true; return HKBCN_OK;
    case HKBCN_RAW_RGBA32F: case HKBCN_RAW_BGRA32F: *channels = 4; *f32 = true; return HKBCN_OK;
    }
    set_err("bad raw image format %d", raw_format);
    return HKBCN_ERR_BAD_FORMAT;
}

// ResampleRawImage (Image.cpp:2188-2243) on device pointers of the current device
static int resample_on(Device* D, int raw_format, const void* d_src, uint32_t w, uint32_t h, const hkbcn_resample_params& rp, void* d_dst, cudaStream_t st)
{
    int ch = 0;
    bool f32 = false;
    int rc = raw_format_info(raw_format, &ch, &f32);
    if (rc) return rc;
    if (!w || !h || !rp.scaled_width || !rp.scaled_height) { set_err("ResampleRawImage: invalid size"); return HKBCN_ERR_BAD_ARG; }
    const int eh = rp.edge_h ? rp.edge_h : 3, ev = rp.edge_v ? rp.edge_v : 3, fh = rp.filter_h ? rp.filter_h : 5, fv = rp.filter_v ? rp.filter_v : 5;
    if (eh < 1 || eh > 4 || ev < 1 || ev > 4 || fh < 1 || fh > 5 || fv < 1 || fv > 5) { set_err("bad edge mode / filter"); return HKBCN_ERR_BAD_ARG; }
    const bool srgb = (rp.flags & HKBCN_RESAMPLE_COLORSPACE_SRGB) != 0, has_alpha = (rp.flags & HKBCN_RESAMPLE_HAS_ALPHA) != 0;
    if (f32 && srgb) { set_err("float images in the sRGB colour space are not supported"); return HKBCN_ERR_UNSUPPORTED; }
    Device::Taps tx, ty;
    if ((rc = get_taps(D, (int)w, (int)rp.scaled_width, fh, eh, tx)) != HKBCN_OK) return rc;
    if ((rc = get_taps(D, (int)h, (int)rp.scaled_height, fv, ev, ty)) != HKBCN_OK) return rc;
    ResizeParams P;
    memset(&P, 0, sizeof(P));
    P.src = (const uint8_t*)d_src; P.dst = (uint8_t*)d_dst;
    P.ix = tx.idx; P.wx = tx.w; P.iy = ty.idx; P.wy = ty.w;
    P.lut = srgb ? D->lut_srgb : D->lut_lin;
    P.in_w = (int)w; P.in_h = (int)h; P.out_w = (int)rp.scaled_width; P.out_h = (int)rp.scaled_height; P.Tx = tx.T; P.Ty = ty.T;
    P.alpha = has_alpha ? ch - 1 : -1;
    P.weight = (has_alpha && !(rp.flags & HKBCN_RESAMPLE_ALPHA_PREMULTIPLIED)) ? 1 : 0;
    P.srgb = srgb ? 1 : 0;
    CK(launch_resample_general(P, ch, f32, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

int hkbcn_resample_raw_image_dev(int device, int raw_format, const void* d_src, uint32_t width, uint32_t height, const hkbcn_resample_params* params,
                                 void* d_dst, void* stream)
{
    Range nvtx_("hkbcn_resample_raw_image_dev");
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = dev_at(device);
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    if (!d_src || !d_dst || !params) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    CK(cudaSetDevice(D->id));
    return resample_on(D, raw_format, d_src, width, height, *params, d_dst, (cudaStream_t)stream);
}

int hkbcn_resample_raw_image(int raw_format, const void* src, uint32_t width, uint32_t height, const hkbcn_resample_params* params, void* dst)
{
    Range nvtx_("hkbcn_resample_raw_image");
    int rc = ensure_init();
    if (rc) return rc;
    if (!src || !dst || !params) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    int ch = 0;
    bool f32 = false;
    if ((rc = raw_format_info(raw_format, &ch, &f32)) != HKBCN_OK) return rc;
    const size_t texel = (size_t)ch * (f32 ? 4 : 1);
    const size_t sbytes = (size_t)width * height * texel, dbytes = (size_t)params->scaled_width * params->scaled_height * texel;
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src, sbytes, cudaMemcpyHostToDevice, D->stream));
    if ((rc = resample_on(D, raw_format, D->d_src, width, height, *params, D->d_dst, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(dst, D->d_dst, dbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

// raw texels (host) -> level 0 in the encoder's source layout -> device cascade -> encode -> host; device 0
// (w, h): the size of the storage; (sw, sh): the size of the raw image -- resampled first when they differ (rs, raw_format)
static int create_image_on(Device* D, const hkbcn_image_plan& plan, const void* src, uint32_t w, uint32_t h, uint32_t mips, const hkbcn_mip_config* cfg,
                           void* dst, uint32_t sw = 0, uint32_t sh = 0, int raw_format = 0, const hkbcn_resample_params* rs = nullptr)
{
    const int fmt = plan.format, codec = codec_of(fmt);
    const bool f32 = codec == C_BC6H;
    const size_t texel = f32 ? 4 : 1;
    const bool pass = plan.add_alpha || plan.swap_rb; // the channel pass is needed
    const bool scale = rs != nullptr;
    if (!scale) { sw = w; sh = h; }
    const size_t raw_bytes = (size_t)w * h * plan.src_channels * texel, in_bytes = (size_t)sw * sh * plan.src_channels * texel;
    const size_t sbytes = chain_src_bytes(codec, 0, w, h, mips), dbytes = hkbcn_chain_bytes(fmt, w, h, 1, mips);
    // d_src: [ the chain | the raw image at the storage's size (channel pass) | the raw image as handed over (resample) ]
    const size_t raw_off = (sbytes + 255) & ~(size_t)255, in_off = (raw_off + (pass ? raw_bytes : 0) + 255) & ~(size_t)255;
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    int rc;
    if ((rc = grow(&D->d_src, &D->src_cap, in_off + (scale ? in_bytes : 0))) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    uint8_t* const d_raw = pass ? (uint8_t*)D->d_src + raw_off : (uint8_t*)D->d_src; // where the raw image at storage size goes
    if (scale) {
        uint8_t* d_in = (uint8_t*)D->d_src + in_off;
        CK(cudaMemcpyAsync(d_in, src, in_bytes, cudaMemcpyHostToDevice, D->stream));
        if ((rc = resample_on(D, raw_format, d_in, sw, sh, *rs, d_raw, D->stream)) != HKBCN_OK) return rc;
    } else {
        CK(cudaMemcpyAsync(d_raw, src, raw_bytes, cudaMemcpyHostToDevice, D->stream));
    }
    if (pass) {
        CK(launch_expand_channels(d_raw, D->d_src, (size_t)w * h, plan.src_channels, plan.swap_rb != 0, f32, D->stream));
        g_launches.fetch_add(1, std::memory_order_relaxed);
    }
    if (mips > 1 && (rc = generate_mips_on(D, fmt, D->d_src, w, h, mips, cfg, 0, D->stream)) != HKBCN_OK) return rc;
    EncodeParams P;
    if ((rc = build_chain(P, codec, 0, D->d_src, D->d_dst, w, h, mips)) != HKBCN_OK) return rc;
    if ((rc = launch_encode(D, fmt, codec, P, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(dst, D->d_dst, dbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_create_image(int raw_format, const void* src, uint32_t width, uint32_t height, const hkbcn_mip_config* cfg, uint32_t storage_flags,
                       uint32_t import_flags, hkbcn_texture_desc* desc, void* dst, size_t capacity)
{
    Range nvtx_("hkbcn_create_image");
    hkbcn_image_plan plan;
    int rc = hkbcn_plan_image(raw_format, width, height, cfg != nullptr, storage_flags, import_flags, &plan);
    if (rc) return rc;
    if (!src || !dst || !width || !height) { set_err("null pointer or empty image"); return HKBCN_ERR_BAD_ARG; }
    if (!plan.format) { set_err("the reference stores raw format %d uncompressed under these import flags: nothing to block-compress", raw_format); return HKBCN_ERR_UNSUPPORTED; }
    // Image.cpp:1070-1093: the image goes to the required size first, with the mip config's edge mode and filter on both axes
    const bool scale = plan.required_width != width || plan.required_height != height;
    hkbcn_resample_params rs;
    memset(&rs, 0, sizeof(rs));
    if (scale) {
        rs.edge_h = rs.edge_v = cfg && cfg->edge_mode ? cfg->edge_mode : 3;
        rs.filter_h = rs.filter_v = cfg && cfg->filter ? cfg->filter : 5;
        const bool srgb_fmt = plan.format == HKBCN_BC1_UNORM_SRGB || plan.format == HKBCN_BC3_UNORM_SRGB;
        if (srgb_fmt) rs.flags |= HKBCN_RESAMPLE_COLORSPACE_SRGB;
        if (plan.src_channels == 4 && !(plan.storage_flags & HKBCN_STORAGE_NO_ALPHA)) {
            rs.flags |= HKBCN_RESAMPLE_HAS_ALPHA;
            if (plan.storage_flags & HKBCN_STORAGE_ALPHA_PREMULTIPLIED) rs.flags |= HKBCN_RESAMPLE_ALPHA_PREMULTIPLIED;
        }
        rs.scaled_width = plan.required_width;
        rs.scaled_height = plan.required_height;
    }
    const uint32_t src_w = width, src_h = height;
    width = plan.required_width;
    height = plan.required_height;
    const uint32_t mips = cfg ? hkbcn_calc_num_mips(width, height) : 1;
    const size_t need = hkbcn_chain_bytes(plan.format, width, height, 1, mips);
    if (capacity < need) { set_err("dst holds %zu bytes, the storage needs %zu", capacity, need); return HKBCN_ERR_BAD_ARG; }
    if ((rc = ensure_init()) != HKBCN_OK) return rc;
    // ImageStorage::GenerateMipmaps (Image.cpp:767-774): alpha weighting unless NO_ALPHA, premultiplied flag passed on
    hkbcn_mip_config mc = cfg ? *cfg : hkbcn_mip_config{0, 0, 0, 0};
    mc.no_alpha = (plan.storage_flags & HKBCN_STORAGE_NO_ALPHA) ? 1 : 0;
    mc.alpha_premultiplied = (plan.storage_flags & HKBCN_STORAGE_ALPHA_PREMULTIPLIED) ? 1 : 0;
    if ((rc = create_image_on(g_devs[0], plan, src, width, height, mips, &mc, dst, src_w, src_h, raw_format, scale ? &rs : nullptr)) != HKBCN_OK) return rc;
    if (desc) {
        desc->type = HKBCN_TEXTURE_2D;
        desc->width = width; desc->height = height; desc->depth = 1;
        desc->num_mips = mips;
        desc->fmt = plan.format;
        desc->storage_flags = plan.storage_flags;
    }
    return HKBCN_OK;
}

int hkbcn_create_roughness_map(const uint8_t* roughness, uint32_t width, uint32_t height, int mipmapped, int edge_mode, int filter,
                               hkbcn_texture_desc* desc, void* dst, size_t capacity)
{
    Range nvtx_("hkbcn_create_roughness_map");
    if (!roughness || !dst || !width || !height) { set_err("null pointer or empty image"); return HKBCN_ERR_BAD_ARG; }
    const uint32_t rw = mipmapped ? std::max<uint32_t>(closest_pow2(width), 4) : (width + 3) & ~3u;
    const uint32_t rh = mipmapped ? std::max<uint32_t>(closest_pow2(height), 4) : (height + 3) & ~3u;
    if (rw != width || rh != height) {
        set_err("CreateRoughnessMap: The width and height of the roughness map must be a power of two and a multiple of blockSize if compression is enabled.");
        return HKBCN_ERR_BAD_ARG;
    }
    const uint32_t mips = mipmapped ? hkbcn_calc_num_mips(width, height) : 1;
    const size_t need = hkbcn_chain_bytes(HKBCN_BC4_UNORM, width, height, 1, mips);
    if (capacity < need) { set_err("dst holds %zu bytes, the storage needs %zu", capacity, need); return HKBCN_ERR_BAD_ARG; }
    int rc = ensure_init();
    if (rc) return rc;
    hkbcn_image_plan plan;
    memset(&plan, 0, sizeof(plan));
    plan.format = HKBCN_BC4_UNORM;
    plan.storage_flags = HKBCN_STORAGE_NO_ALPHA;
    plan.src_channels = 1;
    hkbcn_mip_config mc{edge_mode, filter, 1, 0};
    if ((rc = create_image_on(g_devs[0], plan, roughness, width, height, mips, &mc, dst)) != HKBCN_OK) return rc;
    if (desc) {
        desc->type = HKBCN_TEXTURE_2D;
        desc->width = width; desc->height = height; desc->depth = 1;
        desc->num_mips = mips;
        desc->fmt = HKBCN_BC4_UNORM;
        desc->storage_flags = HKBCN_STORAGE_NO_ALPHA;
    }
    return HKBCN_OK;
}

int hkbcn_encode_blocks(int fmt, const void* src, void* dst, size_t n_blocks, uint32_t flags)
{
    // n packed 4x4 blocks == a surface 4 texels wide and 4*n tall
    if (n_blocks == 0) return HKBCN_OK;
    if (n_blocks > 0x3FFFFFFFu) { set_err("too many blocks"); return HKBCN_ERR_BAD_ARG; }
    return hkbcn_compress(fmt, src, dst, 4, (uint32_t)n_blocks * 4, flags);
}

int hkbcn_compress_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height, uint32_t flags, void* stream)
{
    if (!(width & 3)) return hkbcn_compress_chain_dev(device, fmt, d_src, d_dst, width, height & ~3u, 1, flags, stream);
    // Width % 4 != 0: numBlocksX = Width / 4 with rowStride = Width * bpp (ImageEncoders.cpp:224-226)
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = dev_at(device);
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (width / 4 == 0 || height / 4 == 0) return HKBCN_OK;
    if ((rc = check_dev_ptrs(codec, d_src, d_dst)) != HKBCN_OK) return rc;
    CK(cudaSetDevice(D->id));
    EncodeParams P;
    memset(&P, 0, sizeof(P));
    P.flags = flags;
    add_level(P, codec, flags, d_src, d_dst, width & ~3u, height & ~3u, width);
    return launch_encode(D, fmt, codec, P, (cudaStream_t)stream);
}

int hkbcn_compress_chain_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height, uint32_t num_mips,
                             uint32_t flags, void* stream)
{
    Range nvtx_("hkbcn_compress_chain_dev");
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = dev_at(device);
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (width / 4 == 0 || height / 4 == 0) return HKBCN_OK;
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if ((rc = check_dev_ptrs(codec, d_src, d_dst)) != HKBCN_OK) return rc;
    CK(cudaSetDevice(D->id));
    EncodeParams P;
    if ((rc = build_chain(P, codec, flags, d_src, d_dst, width, height, num_mips)) != HKBCN_OK) return rc;
    return launch_encode(D, fmt, codec, P, (cudaStream_t)stream);
}

// Sliced storage (cube faces, array layers) in one launch: CubemapImporter's per-face CompressBC6h loop (Image.cpp:2033-2035)
// and every mip level of every face, src and dst both in ImageStorage order (all slices of level 0, then of level 1, ...).
int hkbcn_compress_slices_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height, uint32_t slices,
                              uint32_t num_mips, uint32_t flags, void* stream)
{
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = dev_at(device);
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (width == 0 || height == 0) return HKBCN_OK;
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if ((rc = check_dev_ptrs(codec, d_src, d_dst)) != HKBCN_OK) return rc;
    CK(cudaSetDevice(D->id));
    EncodeParams P;
    if ((rc = build_chain(P, codec, flags, d_src, d_dst, width, height, num_mips, slices)) != HKBCN_OK) return rc;
    return launch_encode(D, fmt, codec, P, (cudaStream_t)stream);
}

int hkbcn_compress_slices(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips,
                          uint32_t flags)
{
    Range nvtx_("hkbcn_compress_slices");
    int rc = ensure_init();
    if (rc) return rc;
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if (slices == 0) { set_err("slice count 0"); return HKBCN_ERR_BAD_ARG; }
    return compress_host_on(g_devs[0], fmt, src, dst, width, height, num_mips, flags, slices);
}

size_t hkbcn_slices_src_bytes(int fmt, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips, uint32_t flags)
{
    const int codec = codec_of(fmt);
    return codec < 0 ? 0 : chain_src_bytes(codec, flags, width, height, num_mips, slices);
}

int hkbcn_decompress_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height, size_t dst_row_stride,
                         uint32_t flags, void* stream)
{
    Range nvtx_("hkbcn_decompress_dev");
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = dev_at(device);
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    DecodeParams P;
    P.src = (const uint8_t*)d_src;
    P.dst = (uint8_t*)d_dst;
    P.nbx = width / 4;
    P.nblocks = (width / 4) * (height / 4);
    P.dstRowStride = dst_row_stride;
    P.codec = codec;
    P.is_signed = fmt == HKBCN_BC6H_SFLOAT;
    P.f32out = (flags & HKBCN_FLAG_DST_F32) != 0;
    if (P.nblocks == 0) return HKBCN_OK;
    // the decoders load whole blocks (8 / 16 bytes) and store texel rows; the block stream must be block-aligned
    if ((size_t)d_src % (size_t)blk_bytes(codec)) { set_err("d_src must be %d-byte aligned (one block)", blk_bytes(codec)); return HKBCN_ERR_BAD_ARG; }
    if ((size_t)d_dst % (size_t)(codec == C_BC6H ? 2 : 1) || (codec == C_BC6H && P.f32out && ((size_t)d_dst & 3))) { set_err("d_dst is not aligned to its component type"); return HKBCN_ERR_BAD_ARG; }
    CK(cudaSetDevice(D->id));
    if (!(P.sched = sched_slot(D, (cudaStream_t)stream))) return HKBCN_ERR_CUDA;
    cudaError_t e = launch_decode(P, D->sm_count, (cudaStream_t)stream);
    if (e != cudaSuccess) { set_err("decode launch failed: %s", cudaGetErrorString(e)); return HKBCN_ERR_CUDA; }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

int hkbcn_decompress(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, size_t dst_row_stride, uint32_t flags)
{
    Range nvtx_("hkbcn_decompress");
    int rc = ensure_init();
    if (rc) return rc;
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    const uint32_t nbx = width / 4, nby = height / 4;
    if (nbx == 0 || nby == 0) return HKBCN_OK;
    if (!src || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    const size_t rowBytes = (size_t)nbx * 4 * dst_bpp(codec, flags);
    if (dst_row_stride < rowBytes) { set_err("dst_row_stride %zu < row bytes %zu", dst_row_stride, rowBytes); return HKBCN_ERR_BAD_ARG; }
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    const size_t sbytes = (size_t)nbx * nby * blk_bytes(codec);
    // device image is tightly packed (stride = rowBytes rounded to 16 for vector stores)
    const size_t dstride = (rowBytes + 15) & ~(size_t)15;
    const size_t dbytes = dstride * nby * 4;
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src, sbytes, cudaMemcpyHostToDevice, D->stream));
    rc = hkbcn_decompress_dev(0, fmt, D->d_src, D->d_dst, nbx * 4, nby * 4, dstride, flags, D->stream);
    if (rc) return rc;
    CK(cudaMemcpy2DAsync(dst, dst_row_stride, D->d_dst, dstride, rowBytes, (size_t)nby * 4, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

// ---- whole storages: one launch per mip level, the level's slices as one tall surface ------------------------------------
size_t hkbcn_decoded_chain_bytes(int fmt, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips, uint32_t flags)
{
    const int codec = codec_of(fmt);
    if (codec < 0) return 0;
    size_t n = 0;
    for (uint32_t i = 0; i < num_mips; i++) n += (size_t)mip_dim(width, i) * mip_dim(height, i) * slices * dst_bpp(codec, flags);
    return n;
}

int hkbcn_decompress_slices_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height, uint32_t slices,
                                uint32_t num_mips, uint32_t flags, void* stream)
{
    int rc = ensure_init();
    if (rc) return rc;
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (width == 0 || height == 0) return HKBCN_OK;
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    if (slices == 0 || (uint64_t)height * slices > 0x40000000ull) { set_err("slice count %u out of range", slices); return HKBCN_ERR_BAD_ARG; }
    const uint8_t* s = (const uint8_t*)d_src;
    uint8_t* d = (uint8_t*)d_dst;
    for (uint32_t i = 0; i < num_mips; i++) {
        const uint32_t lw = mip_dim(width, i), lh = mip_dim(height, i) * slices;
        if ((rc = hkbcn_decompress_dev(device, fmt, s, d, lw, lh, (size_t)lw * dst_bpp(codec, flags), flags, stream)) != HKBCN_OK) return rc;
        s += (size_t)(lw / 4) * (lh / 4) * blk_bytes(codec);
        d += (size_t)lw * lh * dst_bpp(codec, flags);
    }
    return HKBCN_OK;
}

int hkbcn_decompress_slices(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t slices, uint32_t num_mips, uint32_t flags)
{
    Range nvtx_("hkbcn_decompress_slices");
    int rc = ensure_init();
    if (rc) return rc;
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (!src || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    if ((rc = check_chain_dims(width, height, num_mips)) != HKBCN_OK) return rc;
    const size_t sbytes = hkbcn_chain_bytes(fmt, width, height, slices, num_mips);
    const size_t dbytes = hkbcn_decoded_chain_bytes(fmt, width, height, slices, num_mips, flags);
    if (sbytes == 0) return HKBCN_OK;
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    if ((rc = grow(&D->d_src, &D->src_cap, sbytes)) != HKBCN_OK) return rc;
    if ((rc = grow(&D->d_dst, &D->dst_cap, dbytes)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src, sbytes, cudaMemcpyHostToDevice, D->stream));
    if ((rc = hkbcn_decompress_slices_dev(0, fmt, D->d_src, D->d_dst, width, height, slices, num_mips, flags, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(dst, D->d_dst, dbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_decompress_chain_dev(int device, int fmt, const void* d_src, void* d_dst, uint32_t width, uint32_t height, uint32_t num_mips,
                               uint32_t flags, void* stream)
{
    return hkbcn_decompress_slices_dev(device, fmt, d_src, d_dst, width, height, 1, num_mips, flags, stream);
}

int hkbcn_decompress_chain(int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t num_mips, uint32_t flags)
{
    return hkbcn_decompress_slices(fmt, src, dst, width, height, 1, num_mips, flags);
}

int hkbcn_decode_blocks(int fmt, const void* src, void* dst, size_t n_blocks, size_t dst_row_stride, uint32_t flags)
{
    if (n_blocks == 0) return HKBCN_OK;
    if (n_blocks > 0x3FFFFFFFu) { set_err("too many blocks"); return HKBCN_ERR_BAD_ARG; }
    return hkbcn_decompress(fmt, src, dst, 4, (uint32_t)n_blocks * 4, dst_row_stride, flags);
}

// Host-only planning step of hkbcn_compress_batch (no device needed): greedy longest-processing-time assignment of whole
// jobs to n_gpus bins by pixel-count-weighted cost.  device_of_job[i] = bin of job i.  Deterministic.
static void plan_batch(const hkbcn_job* jobs, size_t n_jobs, int n_gpus, int* device_of_job)
{
    auto cost = [](const hkbcn_job& j) {
        int c = codec_of(j.fmt);
        double wgt = c == C_BC7 ? 40.0 : c == C_BC6H ? 8000.0 : (c == C_BC4) ? 4.0 : (c == C_BC5) ? 8.0 : (c == C_BC3) ? 5.0 : (c == C_BC2) ? 1.2 : 1.0;
        return wgt * (double)j.width * (double)j.height;
    };
    std::vector<size_t> order(n_jobs);
    for (size_t i = 0; i < n_jobs; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return cost(jobs[a]) > cost(jobs[b]); });
    std::vector<double> load((size_t)n_gpus, 0.0);
    for (size_t idx : order) {
        size_t best = 0;
        for (size_t d = 1; d < (size_t)n_gpus; d++) if (load[d] < load[best]) best = d;
        device_of_job[idx] = (int)best;
        load[best] += cost(jobs[idx]);
    }
}

int hkbcn_plan_batch(const hkbcn_job* jobs, size_t n_jobs, int n_gpus, int* device_of_job)
{
    if (n_gpus <= 0 || (n_jobs && (!jobs || !device_of_job))) { set_err("bad arguments"); return HKBCN_ERR_BAD_ARG; }
    for (size_t i = 0; i < n_jobs; i++)
        if (codec_of(jobs[i].fmt) < 0) { set_err("job %zu: bad format %d", i, jobs[i].fmt); return HKBCN_ERR_BAD_FORMAT; }
    plan_batch(jobs, n_jobs, n_gpus, device_of_job);
    return HKBCN_OK;
}

int hkbcn_compress_batch(hkbcn_job* jobs, size_t n_jobs)
{
    Range nvtx_("hkbcn_compress_batch");
    int rc = ensure_init();
    if (rc) return rc;
    if (n_jobs == 0) return HKBCN_OK;
    if (!jobs) { set_err("null jobs"); return HKBCN_ERR_BAD_ARG; }
    const size_t nd = g_devs.size();
    std::vector<int> assign(n_jobs);
    plan_batch(jobs, n_jobs, (int)nd, assign.data());
    std::vector<std::vector<size_t>> bins(nd);
    for (size_t i = 0; i < n_jobs; i++) bins[(size_t)assign[i]].push_back(i);
    std::vector<std::thread> th;
    std::vector<int> first_err(nd, 0);
    std::vector<std::string> err_msg(nd);
    for (size_t d = 0; d < nd; d++) {
        th.emplace_back([&, d]() {
            for (size_t idx : bins[d]) {
                hkbcn_job& j = jobs[idx];
                if (j.width / 4 == 0 || j.height / 4 == 0) { j.status = HKBCN_OK; continue; }
                j.status = compress_host_on(g_devs[d], j.fmt, j.src, j.dst, j.width, j.height & ~3u, 1, j.flags);
                if (j.status && !first_err[d]) { first_err[d] = j.status; err_msg[d] = t_err; }
            }
        });
    }
    for (auto& t : th) t.join();
    for (size_t d = 0; d < nd; d++)
        if (first_err[d]) { set_err("%s", err_msg[d].c_str()); return first_err[d]; }
    return HKBCN_OK;
}

// independent surfaces decoded on the selected GPUs, whole surfaces per GPU (greedy by output bytes), results straight into
// each job's dst; per GPU the uploads, launches and downloads of consecutive jobs are queued without a host sync in between
int hkbcn_decompress_batch(hkbcn_job* jobs, size_t n_jobs)
{
    Range nvtx_("hkbcn_decompress_batch");
    int rc = ensure_init();
    if (rc) return rc;
    if (n_jobs == 0) return HKBCN_OK;
    if (!jobs) { set_err("null jobs"); return HKBCN_ERR_BAD_ARG; }
    const size_t nd = g_devs.size();
    std::vector<size_t> order(n_jobs);
    for (size_t i = 0; i < n_jobs; i++) order[i] = i;
    auto out_bytes = [](const hkbcn_job& j) {
        const int c = codec_of(j.fmt);
        return c < 0 ? (size_t)0 : (size_t)(j.width / 4) * 4 * (j.height / 4) * 4 * dst_bpp(c, j.flags);
    };
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return out_bytes(jobs[a]) > out_bytes(jobs[b]); });
    std::vector<std::vector<size_t>> bins(nd);
    std::vector<size_t> load(nd, 0);
    for (size_t idx : order) {
        size_t best = 0;
        for (size_t d = 1; d < nd; d++) if (load[d] < load[best]) best = d;
        bins[best].push_back(idx);
        load[best] += out_bytes(jobs[idx]);
    }
    std::vector<std::thread> th;
    std::vector<int> first_err(nd, 0);
    std::vector<std::string> err_msg(nd);
    for (size_t dv = 0; dv < nd; dv++) {
        th.emplace_back([&, dv]() {
            Device* D = g_devs[dv];
            std::lock_guard<std::mutex> lk(D->mu);
            auto fail = [&](hkbcn_job& j, int e) {
                j.status = e;
                if (!first_err[dv]) { first_err[dv] = e; err_msg[dv] = t_err; }
            };
            if (cudaSetDevice(D->id) != cudaSuccess) { set_err("cudaSetDevice failed"); for (size_t idx : bins[dv]) fail(jobs[idx], HKBCN_ERR_CUDA); return; }
            // device scratch for the whole bin: every job gets its own range, so nothing waits for a buffer
            size_t sb = 0, db = 0;
            for (size_t idx : bins[dv]) {
                const hkbcn_job& j = jobs[idx];
                const int c = codec_of(j.fmt);
                if (c < 0) continue;
                sb += ((size_t)(j.width / 4) * (j.height / 4) * blk_bytes(c) + 255) & ~(size_t)255;
                db += (out_bytes(j) + 255) & ~(size_t)255;
            }
            int e;
            if ((e = grow(&D->d_src, &D->src_cap, sb)) != HKBCN_OK || (e = grow(&D->d_dst, &D->dst_cap, db)) != HKBCN_OK) {
                for (size_t idx : bins[dv]) fail(jobs[idx], e);
                return;
            }
            size_t so = 0, dof = 0;
            for (size_t idx : bins[dv]) {
                hkbcn_job& j = jobs[idx];
                j.status = HKBCN_OK;
                const int c = codec_of(j.fmt);
                if (c < 0) { set_err("bad format %d", j.fmt); fail(j, HKBCN_ERR_BAD_FORMAT); continue; }
                const uint32_t nbx = j.width / 4, nby = j.height / 4;
                if (!nbx || !nby) continue;
                if (!j.src || !j.dst) { set_err("null pointer"); fail(j, HKBCN_ERR_BAD_ARG); continue; }
                const size_t sbytes = (size_t)nbx * nby * blk_bytes(c), dbytes = out_bytes(j);
                cudaError_t ce = cudaMemcpyAsync((uint8_t*)D->d_src + so, j.src, sbytes, cudaMemcpyHostToDevice, D->stream);
                if (ce == cudaSuccess) {
                    e = hkbcn_decompress_dev((int)dv, j.fmt, (uint8_t*)D->d_src + so, (uint8_t*)D->d_dst + dof, nbx * 4, nby * 4,
                                             (size_t)nbx * 4 * dst_bpp(c, j.flags), j.flags, D->stream);
                    if (e) { fail(j, e); continue; }
                    ce = cudaMemcpyAsync(j.dst, (uint8_t*)D->d_dst + dof, dbytes, cudaMemcpyDeviceToHost, D->stream);
                }
                if (ce != cudaSuccess) { set_err("copy failed: %s", cudaGetErrorString(ce)); fail(j, HKBCN_ERR_CUDA); }
                so += (sbytes + 255) & ~(size_t)255;
                dof += (dbytes + 255) & ~(size_t)255;
            }
            if (cudaStreamSynchronize(D->stream) != cudaSuccess) { set_err("decode batch failed on device %d", D->id); for (size_t idx : bins[dv]) fail(jobs[idx], HKBCN_ERR_CUDA); }
        });
    }
    for (auto& t : th) t.join();
    for (size_t d = 0; d < nd; d++)
        if (first_err[d]) { set_err("%s", err_msg[d].c_str()); return first_err[d]; }
    return HKBCN_OK;
}

// ---- normal maps and roughness (row f4) -------------------------------------------------------------------------------
int hkbcn_pack_texel_bytes(int variant)
{
    if (variant == NP_RGBA_BC1 || variant == NP_RGBA_BC3) return 4;
    if (variant >= NP_RG_BC5 && variant <= NP_PARABOLOID_BC5) return 2;
    return -1;
}

#define DEV_OR_RETURN(D, device)                                                                   \
    {                                                                                              \
        const int rc_ = ensure_init();                                                             \
        if (rc_) return rc_;                                                                       \
    }                                                                                              \
    Device* D = dev_at(device);                                                                    \
    if (!D) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }

int hkbcn_pack_normals_dev(int device, int variant, const float* d_n, void* d_dst, size_t count, void* stream)
{
    DEV_OR_RETURN(D, device)
    if (hkbcn_pack_texel_bytes(variant) < 0) { set_err("bad normal-map packing %d", variant); return HKBCN_ERR_BAD_ARG; }
    if (count == 0) return HKBCN_OK;
    CK(cudaSetDevice(D->id));
    CK(launch_pack_normals(variant, d_n, (uint8_t*)d_dst, count, (cudaStream_t)stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

int hkbcn_unpack_normals_dev(int device, int variant, const void* d_src, float* d_n, size_t count, void* stream)
{
    DEV_OR_RETURN(D, device)
    if (hkbcn_pack_texel_bytes(variant) < 0) { set_err("bad normal-map packing %d", variant); return HKBCN_ERR_BAD_ARG; }
    if (count == 0) return HKBCN_OK;
    CK(cudaSetDevice(D->id));
    CK(launch_unpack_normals(variant, (const uint8_t*)d_src, d_n, count, (cudaStream_t)stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

int hkbcn_normalize_vectors_dev(int device, float* d_v, size_t count, void* stream)
{
    DEV_OR_RETURN(D, device)
    if (count == 0) return HKBCN_OK;
    CK(cudaSetDevice(D->id));
    CK(launch_normalize_vectors(d_v, count, (cudaStream_t)stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

size_t hkbcn_bake_scratch_bytes(uint32_t width, uint32_t height)
{
    // two ping-pong images of averaged normals: level 1 and level 2 sizes
    return 12 * ((size_t)mip_dim(width, 1) * mip_dim(height, 1) + (size_t)mip_dim(width, 2) * mip_dim(height, 2));
}

static int bake_roughness_on(Device* D, const float* d_n, uint8_t* d_rough, void* d_scratch, uint32_t w, uint32_t h, uint32_t mips, int bake,
                             int filter, int edge, cudaStream_t st)
{
    if (bake != HKBCN_ROUGHNESS_BAKE_VMF && bake != HKBCN_ROUGHNESS_BAKE_TOKSVIG) { set_err("bad roughness bake %d", bake); return HKBCN_ERR_BAD_ARG; }
    if (!filter) filter = 2; // NormalMapResampleFilter = triangle (Image.cpp:2359)
    if (!edge) edge = 3;
    if (edge < 1 || edge > 4 || filter < 1 || filter > 5) { set_err("bad edge mode %d / filter %d", edge, filter); return HKBCN_ERR_BAD_ARG; }
    if (mips == 0 || mips > HKBCN_MAX_LEVELS || !w || !h) { set_err("bad chain %u x %u, %u mips", w, h, mips); return HKBCN_ERR_BAD_ARG; }
    uint8_t* a = (uint8_t*)d_scratch;
    uint8_t* b = a + 12 * (size_t)mip_dim(w, 1) * mip_dim(h, 1);
    const uint8_t* src = (const uint8_t*)d_n;
    uint8_t* rough = d_rough + (size_t)w * h;
    for (uint32_t i = 1; i < mips; i++) {
        const uint32_t iw = mip_dim(w, i - 1), ih = mip_dim(h, i - 1), ow = mip_dim(w, i), oh = mip_dim(h, i);
        uint8_t* dst = (i & 1) ? a : b;
        Device::Taps tx, ty;
        int rc;
        if ((rc = get_taps(D, (int)iw, (int)ow, filter, edge, tx)) != HKBCN_OK) return rc;
        if ((rc = get_taps(D, (int)ih, (int)oh, filter, edge, ty)) != HKBCN_OK) return rc;
        ResizeParams P;
        P.src = src; P.dst = dst;
        P.ix = tx.idx; P.wx = tx.w; P.iy = ty.idx; P.wy = ty.w;
        P.lut = D->lut_lin;
        P.in_w = (int)iw; P.in_h = (int)ih; P.out_w = (int)ow; P.out_h = (int)oh; P.Tx = tx.T; P.Ty = ty.T;
        P.alpha = -1; P.weight = 0; P.srgb = 0;
        set_tile_taps(P, tx, ty);
        CK(launch_mip_resize(P, 3, true, st));
        CK(launch_roughness_bake(bake, (const float*)dst, rough, (size_t)ow * oh, st));
        g_launches.fetch_add(2, std::memory_order_relaxed);
        src = dst;
        rough += (size_t)ow * oh;
    }
    return HKBCN_OK;
}

int hkbcn_bake_roughness_dev(int device, const float* d_n, uint8_t* d_rough, void* d_scratch, uint32_t w, uint32_t h, uint32_t mips, int bake,
                             int filter, int edge, void* stream)
{
    DEV_OR_RETURN(D, device)
    CK(cudaSetDevice(D->id));
    return bake_roughness_on(D, d_n, d_rough, d_scratch, w, h, mips, bake, filter, edge, (cudaStream_t)stream);
}

// host-pointer forms: device 0's scratch buffers, synchronous
static int with_scratch(Device* D, size_t in_bytes, size_t out_bytes)
{
    int rc;
    CK(cudaSetDevice(D->id));
    if ((rc = grow(&D->d_src, &D->src_cap, in_bytes)) != HKBCN_OK) return rc;
    return grow(&D->d_dst, &D->dst_cap, out_bytes);
}

int hkbcn_pack_normals(int variant, const float* n, void* dst, size_t count)
{
    int rc = ensure_init();
    if (rc) return rc;
    const int tb = hkbcn_pack_texel_bytes(variant);
    if (tb < 0) { set_err("bad normal-map packing %d", variant); return HKBCN_ERR_BAD_ARG; }
    if (count == 0) return HKBCN_OK;
    if (!n || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    if ((rc = with_scratch(D, count * 12, count * (size_t)tb)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, n, count * 12, cudaMemcpyHostToDevice, D->stream));
    CK(launch_pack_normals(variant, (const float*)D->d_src, (uint8_t*)D->d_dst, count, D->stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CK(cudaMemcpyAsync(dst, D->d_dst, count * (size_t)tb, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_unpack_normals(int variant, const void* src, float* n, size_t count)
{
    int rc = ensure_init();
    if (rc) return rc;
    const int tb = hkbcn_pack_texel_bytes(variant);
    if (tb < 0) { set_err("bad normal-map packing %d", variant); return HKBCN_ERR_BAD_ARG; }
    if (count == 0) return HKBCN_OK;
    if (!n || !src) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    if ((rc = with_scratch(D, count * (size_t)tb, count * 12)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, src, count * (size_t)tb, cudaMemcpyHostToDevice, D->stream));
    CK(launch_unpack_normals(variant, (const uint8_t*)D->d_src, (float*)D->d_dst, count, D->stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CK(cudaMemcpyAsync(n, D->d_dst, count * 12, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_normalize_vectors(float* v, size_t count)
{
    int rc = ensure_init();
    if (rc) return rc;
    if (count == 0) return HKBCN_OK;
    if (!v) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    if ((rc = with_scratch(D, count * 12, 0)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, v, count * 12, cudaMemcpyHostToDevice, D->stream));
    CK(launch_normalize_vectors((float*)D->d_src, count, D->stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CK(cudaMemcpyAsync(v, D->d_src, count * 12, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_bake_roughness(const float* n, uint8_t* rough, uint32_t w, uint32_t h, uint32_t mips, int bake, int filter, int edge)
{
    int rc = ensure_init();
    if (rc) return rc;
    if (!n || !rough) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    if (mips == 0 || mips > HKBCN_MAX_LEVELS || !w || !h) { set_err("bad chain %u x %u, %u mips", w, h, mips); return HKBCN_ERR_BAD_ARG; }
    size_t rbytes = 0;
    for (uint32_t i = 0; i < mips; i++) rbytes += (size_t)mip_dim(w, i) * mip_dim(h, i);
    const size_t nbytes = (size_t)w * h * 12, sbytes = hkbcn_bake_scratch_bytes(w, h);
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    // d_src: normals followed by the scratch; d_dst: the roughness chain
    if ((rc = with_scratch(D, nbytes + sbytes + 256, rbytes)) != HKBCN_OK) return rc;
    uint8_t* scratch = (uint8_t*)D->d_src + ((nbytes + 255) & ~(size_t)255);
    CK(cudaMemcpyAsync(D->d_src, n, nbytes, cudaMemcpyHostToDevice, D->stream));
    CK(cudaMemcpyAsync(D->d_dst, rough, rbytes, cudaMemcpyHostToDevice, D->stream));
    if ((rc = bake_roughness_on(D, (const float*)D->d_src, (uint8_t*)D->d_dst, scratch, w, h, mips, bake, filter, edge, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(rough, D->d_dst, rbytes, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

// ---- CreateNormalMap (Image.cpp:2262-2352) and CreateNormalAndRoughness (Image.cpp:2459-2638) as single calls -----------------
namespace {
struct DevBuf { // device scratch of one composed import call
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t n) { if (cudaMalloc(&p, n ? n : 1) != cudaSuccess) { set_err("cudaMalloc(%zu) failed: %s", n, cudaGetErrorString(cudaGetLastError())); return HKBCN_ERR_CUDA; } return HKBCN_OK; }
};
int pack_format(int pack) { return pack == HKBCN_NORMALS_RGBA_BC1 ? HKBCN_BC1_UNORM : pack == HKBCN_NORMALS_RGBA_BC3 ? HKBCN_BC3_UNORM : HKBCN_BC5_UNORM; }
uint32_t required_pow2(uint32_t v) { return std::max<uint32_t>(closest_pow2(v), 4); }

// normalised Float3 normals on the device -> packed level 0 -> cascade of the packed texels (NO_ALPHA) -> encode -> d_out
int normal_map_on(Device* D, const float* d_normals, uint32_t w, uint32_t h, int pack, uint32_t mips, int edge, int filter, void* d_out, cudaStream_t st)
{
    const int fmt = pack_format(pack), codec = codec_of(fmt);
    DevBuf chain;
    int rc = chain.alloc(chain_src_bytes(codec, 0, w, h, mips));
    if (rc) return rc;
    CK(launch_pack_normals(pack, d_normals, (uint8_t*)chain.p, (size_t)w * h, st)); // FIXME of the reference kept: packed once, not per level
    g_launches.fetch_add(1, std::memory_order_relaxed);
    hkbcn_mip_config mc{edge, filter, 1, 0};
    if (mips > 1 && (rc = generate_mips_on(D, fmt, chain.p, w, h, mips, &mc, 0, st)) != HKBCN_OK) return rc;
    EncodeParams P;
    if ((rc = build_chain(P, codec, 0, chain.p, d_out, w, h, mips)) != HKBCN_OK) return rc;
    if ((rc = launch_encode(D, fmt, codec, P, st)) != HKBCN_OK) return rc;
    CK(cudaStreamSynchronize(st)); // chain is freed on return
    return HKBCN_OK;
}
void fill_desc(hkbcn_texture_desc* d, uint32_t w, uint32_t h, uint32_t mips, int fmt)
{
    if (!d) return;
    d->type = HKBCN_TEXTURE_2D;
    d->width = w; d->height = h; d->depth = 1;
    d->num_mips = mips;
    d->fmt = fmt;
    d->storage_flags = HKBCN_STORAGE_NO_ALPHA;
}
} // namespace

int hkbcn_create_normal_map(const float* normals_xyz, uint32_t width, uint32_t height, int pack, int mipmapped, int edge_mode, int filter,
                            hkbcn_texture_desc* desc, void* dst, size_t capacity)
{
    Range nvtx_("hkbcn_create_normal_map");
    if (hkbcn_pack_texel_bytes(pack) < 0) { set_err("bad normal-map packing %d", pack); return HKBCN_ERR_BAD_ARG; }
    if (!normals_xyz || !width || !height) { set_err("null pointer or empty image"); return HKBCN_ERR_BAD_ARG; }
    const uint32_t rw = mipmapped ? required_pow2(width) : (width + 3) & ~3u, rh = mipmapped ? required_pow2(height) : (height + 3) & ~3u;
    if (rw != width || rh != height) {
        set_err("CreateNormalMap: The width and height of the normal map must be a power of two and a multiple of blockSize if compression is enabled.");
        return HKBCN_ERR_BAD_ARG;
    }
    const int fmt = pack_format(pack);
    const uint32_t mips = mipmapped ? hkbcn_calc_num_mips(width, height) : 1;
    const size_t need = hkbcn_chain_bytes(fmt, width, height, 1, mips);
    fill_desc(desc, width, height, mips, fmt);
    if (!dst) return HKBCN_OK; // size query
    if (capacity < need) { set_err("dst holds %zu bytes, the storage needs %zu", capacity, need); return HKBCN_ERR_BAD_ARG; }
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    DevBuf n, out;
    if ((rc = n.alloc((size_t)width * height * 12)) != HKBCN_OK || (rc = out.alloc(need)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(n.p, normals_xyz, (size_t)width * height * 12, cudaMemcpyHostToDevice, D->stream));
    if ((rc = normal_map_on(D, (const float*)n.p, width, height, pack, mips, edge_mode ? edge_mode : 3, filter ? filter : 5, out.p, D->stream)) != HKBCN_OK) return rc;
    CK(cudaMemcpy(dst, out.p, need, cudaMemcpyDeviceToHost));
    return HKBCN_OK;
}

int hkbcn_create_normal_and_roughness(const hkbcn_normal_roughness_settings* s, hkbcn_texture_desc* normal_desc, void* normal_dst, size_t normal_capacity,
                                      hkbcn_texture_desc* rough_desc, void* rough_dst, size_t rough_capacity)
{
    Range nvtx_("hkbcn_create_normal_and_roughness");
    if (!s || !s->normals_rgb32f || !s->roughness_r8 || !s->normal_width || !s->normal_height || !s->roughness_width || !s->roughness_height) {
        set_err("null pointer or empty image");
        return HKBCN_ERR_BAD_ARG;
    }
    if (hkbcn_pack_texel_bytes(s->pack) < 0) { set_err("bad normal-map packing %d", s->pack); return HKBCN_ERR_BAD_ARG; }
    if (s->roughness_bake != HKBCN_ROUGHNESS_BAKE_VMF && s->roughness_bake != HKBCN_ROUGHNESS_BAKE_TOKSVIG) { set_err("bad roughness bake %d", s->roughness_bake); return HKBCN_ERR_BAD_ARG; }
    const int edge = s->edge_mode ? s->edge_mode : 3, tri = 2; // NormalMapResampleFilter = RoughnessMapResampleFilter = triangle (:2463-2464)
    if (edge < 1 || edge > 4) { set_err("bad edge mode %d", edge); return HKBCN_ERR_BAD_ARG; }
    // sizes (:2475-2490, :2601-2602): the roughness map goes to powers of two, the normal map first to the roughness map's size
    // (for the bake), then to the powers of two of its OWN size for its storage
    const uint32_t rw = required_pow2(s->roughness_width), rh = required_pow2(s->roughness_height);
    const uint32_t nw = required_pow2(s->normal_width), nh = required_pow2(s->normal_height);
    // :2604-2626: the normalised image made for the bake (roughness-sized) is the one stored when it already has the normal map's
    // required size -- and also when the SOURCE has it (neither branch of the reference's if / else-if replaces it then)
    const bool use_bake = (nw == rw && nh == rh) || (s->normal_width == nw && s->normal_height == nh);
    const uint32_t ow = use_bake ? rw : nw, oh = use_bake ? rh : nh;
    const uint32_t rmips = hkbcn_calc_num_mips(rw, rh), omips = hkbcn_calc_num_mips(ow, oh);
    const int nfmt = pack_format(s->pack);
    const size_t rneed = hkbcn_chain_bytes(HKBCN_BC4_UNORM, rw, rh, 1, rmips), oneed = hkbcn_chain_bytes(nfmt, ow, oh, 1, omips);
    fill_desc(rough_desc, rw, rh, rmips, HKBCN_BC4_UNORM);
    fill_desc(normal_desc, ow, oh, omips, nfmt);
    if (!normal_dst && !rough_dst) return HKBCN_OK; // size query
    if (!normal_dst || !rough_dst) { set_err("both destinations or neither"); return HKBCN_ERR_BAD_ARG; }
    if (normal_capacity < oneed || rough_capacity < rneed) { set_err("destinations hold %zu / %zu bytes, the storages need %zu / %zu", normal_capacity, rough_capacity, oneed, rneed); return HKBCN_ERR_BAD_ARG; }
    int rc = ensure_init();
    if (rc) return rc;
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    cudaStream_t st = D->stream;
    const size_t src_n = (size_t)s->normal_width * s->normal_height, src_r = (size_t)s->roughness_width * s->roughness_height;
    DevBuf n_src, r_src, r_chain, n_bake, scratch, r_out, n_out, n_own;
    // ---- roughness level 0 at its power-of-two size
    if ((rc = r_src.alloc(src_r)) != HKBCN_OK || (rc = r_chain.alloc(chain_src_bytes(C_BC4, 0, rw, rh, rmips))) != HKBCN_OK) return rc;
    if (rw != s->roughness_width || rh != s->roughness_height) {
        CK(cudaMemcpyAsync(r_src.p, s->roughness_r8, src_r, cudaMemcpyHostToDevice, st));
        hkbcn_resample_params rp{0, edge, edge, tri, tri, rw, rh};
        if ((rc = resample_on(D, HKBCN_RAW_R8, r_src.p, s->roughness_width, s->roughness_height, rp, r_chain.p, st)) != HKBCN_OK) return rc;
    } else {
        CK(cudaMemcpyAsync(r_chain.p, s->roughness_r8, src_r, cudaMemcpyHostToDevice, st));
    }
    // ---- the normal map as loaded (green inverted for DirectX maps), then at the roughness map's size, normalised
    if ((rc = n_src.alloc(src_n * 12)) != HKBCN_OK || (rc = n_bake.alloc((size_t)rw * rh * 12)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(n_src.p, s->normals_rgb32f, src_n * 12, cudaMemcpyHostToDevice, st));
    if (s->convert_from_directx) { CK(launch_invert_green((float*)n_src.p, src_n, st)); g_launches.fetch_add(1, std::memory_order_relaxed); }
    if (s->normal_width != rw || s->normal_height != rh) {
        hkbcn_resample_params rp{0, edge, edge, tri, tri, rw, rh};
        if ((rc = resample_on(D, HKBCN_RAW_RGB32F, n_src.p, s->normal_width, s->normal_height, rp, n_bake.p, st)) != HKBCN_OK) return rc;
    } else {
        CK(cudaMemcpyAsync(n_bake.p, n_src.p, src_n * 12, cudaMemcpyDeviceToDevice, st));
    }
    CK(launch_normalize_vectors((float*)n_bake.p, (size_t)rw * rh, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    // ---- roughness chain (CreateRoughnessMap, uncompressed, mipmapped: :2517), the update of levels >= 1, BC4
    hkbcn_mip_config rmc{edge, tri, 1, 0};
    if (rmips > 1 && (rc = generate_mips_on(D, HKBCN_BC4_UNORM, r_chain.p, rw, rh, rmips, &rmc, 0, st)) != HKBCN_OK) return rc;
    if ((rc = scratch.alloc(hkbcn_bake_scratch_bytes(rw, rh))) != HKBCN_OK) return rc;
    if ((rc = bake_roughness_on(D, (const float*)n_bake.p, (uint8_t*)r_chain.p, scratch.p, rw, rh, rmips, s->roughness_bake, tri, edge, st)) != HKBCN_OK) return rc;
    if ((rc = r_out.alloc(rneed)) != HKBCN_OK) return rc;
    {
        EncodeParams P;
        if ((rc = build_chain(P, C_BC4, 0, r_chain.p, r_out.p, rw, rh, rmips)) != HKBCN_OK) return rc;
        if ((rc = launch_encode(D, HKBCN_BC4_UNORM, C_BC4, P, st)) != HKBCN_OK) return rc;
    }
    // ---- the normal map's own storage (:2601-2628)
    const float* d_normals = (const float*)n_bake.p;
    if (!use_bake) {
        if ((rc = n_own.alloc((size_t)nw * nh * 12)) != HKBCN_OK) return rc;
        hkbcn_resample_params rp{0, edge, edge, tri, tri, nw, nh};
        if ((rc = resample_on(D, HKBCN_RAW_RGB32F, n_src.p, s->normal_width, s->normal_height, rp, n_own.p, st)) != HKBCN_OK) return rc;
        CK(launch_normalize_vectors((float*)n_own.p, (size_t)nw * nh, st));
        g_launches.fetch_add(1, std::memory_order_relaxed);
        d_normals = (const float*)n_own.p;
    }
    if ((rc = n_out.alloc(oneed)) != HKBCN_OK) return rc;
    if ((rc = normal_map_on(D, d_normals, ow, oh, s->pack, omips, edge, tri, n_out.p, st)) != HKBCN_OK) return rc;
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpy(rough_dst, r_out.p, rneed, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(normal_dst, n_out.p, oneed, cudaMemcpyDeviceToHost));
    return HKBCN_OK;
}

// ---- PBRImporter's ORMX texture (Tools/PBRImporter/Main.cpp:284-735) as a batch job ------------------------------------------
// maps up -> CreateORMX gather -> cascade of the RGBA8 storage (NO_ALPHA) -> roughness channel of levels >= 1 from the
// triangle-filtered normals -> encode -> compressed chain down.  One job at a time per device, under the device's mutex.
static int import_pbr_on(Device* D, hkbcn_pbr_job& j, std::vector<uint8_t>& out, uint32_t& mips_out)
{
    const int codec = codec_of(j.fmt);
    const uint32_t w = j.width, h = j.height;
    const uint32_t mips = j.generate_mips ? hkbcn_calc_num_mips(w, h) : 1u;
    mips_out = mips;
    const size_t n = (size_t)w * h, cbytes = chain_src_bytes(codec, 0, w, h, mips), dbytes = hkbcn_chain_bytes(j.fmt, w, h, 1, mips);
    const bool bake = j.roughness_bake != 0 && j.normals && mips > 1; // no normals or no mips: nothing to anti-alias (Main.cpp:616)
    std::lock_guard<std::mutex> lk(D->mu);
    CK(cudaSetDevice(D->id));
    cudaStream_t st = D->stream;
    int rc;
    DevBuf maps[4], chain, blocks, normals, plane, scratch;
    if ((rc = chain.alloc(cbytes)) != HKBCN_OK || (rc = blocks.alloc(dbytes)) != HKBCN_OK) return rc;
    const void* d_maps[4] = {nullptr, nullptr, nullptr, nullptr};
    int ch[4];
    for (int k = 0; k < 4; k++) {
        ch[k] = j.channel[k];
        if (!j.maps[k]) continue;
        if ((rc = maps[k].alloc(n * 4)) != HKBCN_OK) return rc;
        CK(cudaMemcpyAsync(maps[k].p, j.maps[k], n * 4, cudaMemcpyHostToDevice, st));
        d_maps[k] = maps[k].p;
    }
    CK(launch_ormx_gather(d_maps, ch, chain.p, n, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    hkbcn_mip_config cfg = j.mip;
    cfg.no_alpha = 1; // ImageStorageDesc.Flags = IMAGE_STORAGE_NO_ALPHA (:601): the fourth channel is data, not coverage
    cfg.alpha_premultiplied = 0;
    // the cascade is that of an RGBA8 storage whatever the preset compresses to (:596-613)
    if (mips > 1 && (rc = generate_mips_on(D, HKBCN_BC7_UNORM, chain.p, w, h, mips, &cfg, 0, st)) != HKBCN_OK) return rc;
    if (bake) {
        size_t texels = 0;
        for (uint32_t i = 0; i < mips; i++) texels += (size_t)mip_dim(w, i) * mip_dim(h, i);
        if ((rc = normals.alloc(n * 12)) != HKBCN_OK || (rc = plane.alloc(texels)) != HKBCN_OK || (rc = scratch.alloc(hkbcn_bake_scratch_bytes(w, h))) != HKBCN_OK) return rc;
        CK(cudaMemcpyAsync(normals.p, j.normals, n * 12, cudaMemcpyHostToDevice, st));
        CK(launch_channel_move(chain.p, plane.p, texels, 1, false, st));
        if ((rc = bake_roughness_on(D, (const float*)normals.p, (uint8_t*)plane.p, scratch.p, w, h, mips, j.roughness_bake, 2 /* triangle, :640 */,
                                    cfg.edge_mode, st)) != HKBCN_OK) return rc;
        CK(launch_channel_move(chain.p, plane.p, texels, 1, true, st));
        g_launches.fetch_add(2, std::memory_order_relaxed);
    }
    EncodeParams P;
    if ((rc = build_chain(P, codec, 0, chain.p, blocks.p, w, h, mips)) != HKBCN_OK) return rc;
    if ((rc = launch_encode(D, j.fmt, codec, P, st)) != HKBCN_OK) return rc;
    out.resize(dbytes);
    void* to = j.dst ? j.dst : (void*)out.data();
    CK(cudaMemcpyAsync(to, blocks.p, dbytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (j.dst && j.out_path) memcpy(out.data(), j.dst, dbytes);
    return HKBCN_OK;
}

int hkbcn_import_pbr_batch(hkbcn_pbr_job* jobs, size_t n_jobs)
{
    Range nvtx_("hkbcn_import_pbr_batch");
    int rc = ensure_init();
    if (rc) return rc;
    if (n_jobs == 0) return HKBCN_OK;
    if (!jobs) { set_err("null jobs"); return HKBCN_ERR_BAD_ARG; }
    for (size_t i = 0; i < n_jobs; i++) {
        hkbcn_pbr_job& j = jobs[i];
        j.status = HKBCN_OK;
        if (j.fmt != HKBCN_BC1_UNORM && j.fmt != HKBCN_BC3_UNORM && j.fmt != HKBCN_BC7_UNORM) {
            set_err("job %zu: the ORMX presets compress to BC1, BC3 or BC7, not format %d", i, j.fmt);
            return HKBCN_ERR_BAD_FORMAT;
        }
        if (!j.maps[0] && !j.maps[1] && !j.maps[2] && !j.maps[3]) { set_err("job %zu: at least one source map is needed", i); return HKBCN_ERR_BAD_ARG; }
        if (!j.out_path && !j.dst) { set_err("job %zu: neither out_path nor dst", i); return HKBCN_ERR_BAD_ARG; }
        for (int k = 0; k < 4; k++)
            if (j.channel[k] > 3) { set_err("job %zu: invalid channel index %d", i, (int)j.channel[k]); return HKBCN_ERR_BAD_ARG; }
        if (j.roughness_bake != 0 && j.roughness_bake != HKBCN_ROUGHNESS_BAKE_VMF && j.roughness_bake != HKBCN_ROUGHNESS_BAKE_TOKSVIG) {
            set_err("job %zu: bad roughness bake %d", i, j.roughness_bake);
            return HKBCN_ERR_BAD_ARG;
        }
        if ((rc = check_chain_dims(j.width, j.height, j.generate_mips ? hkbcn_calc_num_mips(j.width, j.height) : 1u)) != HKBCN_OK) return rc;
    }
    const size_t nd = g_devs.size();
    std::vector<int> first_err(nd, 0);
    std::vector<std::string> err_msg(nd);
    std::vector<std::thread> th;
    for (size_t dv = 0; dv < nd && dv < n_jobs; dv++) {
        th.emplace_back([&, dv]() {
            std::vector<uint8_t> out;
            for (size_t i = dv; i < n_jobs; i += nd) {
                hkbcn_pbr_job& j = jobs[i];
                uint32_t mips = 1;
                int r = import_pbr_on(g_devs[dv], j, out, mips);
                if (r == HKBCN_OK && j.out_path) {
                    hkbcn_texture_desc desc{HKBCN_TEXTURE_2D, j.width, j.height, 1, mips, j.fmt, (uint32_t)HKBCN_STORAGE_NO_ALPHA};
                    r = hkbcn_write_texture(j.out_path, &desc, out.data(), out.size());
                }
                j.status = r;
                if (r != HKBCN_OK && !first_err[dv]) { first_err[dv] = r; err_msg[dv] = t_err; }
            }
        });
    }
    for (auto& t : th) t.join();
    for (size_t dv = 0; dv < nd; dv++)
        if (first_err[dv]) { set_err("%s", err_msg[dv].c_str()); return first_err[dv]; }
    return HKBCN_OK;
}

// ---- environment probes (row f3) --------------------------------------------------------------------------------------
size_t hkbcn_envprobe_texels(int max_lod) { return (max_lod < 1 || max_lod > 12) ? 0 : envprobe_texels(max_lod); }

static int probe_args_ok(uint32_t src_size, int max_lod)
{
    if (src_size < 1 || src_size > 8192) { set_err("bad source cube size %u", src_size); return HKBCN_ERR_BAD_ARG; }
    if (max_lod < 1 || max_lod > 12) { set_err("max_lod %d out of range 1..12 (roughness of level L is L / max_lod)", max_lod); return HKBCN_ERR_BAD_ARG; }
    return HKBCN_OK;
}

int hkbcn_envprobe_prefilter_dev(int device, const float* d_cube, uint32_t src_size, int max_lod, float* d_dst, void* stream)
{
    DEV_OR_RETURN(D, device)
    int rc = probe_args_ok(src_size, max_lod);
    if (rc) return rc;
    CK(cudaSetDevice(D->id));
    CK(launch_envprobe_prefilter(d_cube, (int)src_size, max_lod, d_dst, (cudaStream_t)stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

int hkbcn_irradiance_dev(int device, const float* d_cube, uint32_t src_size, uint32_t out_size, float* d_dst, void* stream)
{
    DEV_OR_RETURN(D, device)
    int rc = probe_args_ok(src_size, 1);
    if (rc) return rc;
    if (out_size < 1 || out_size > 1024) { set_err("bad irradiance size %u", out_size); return HKBCN_ERR_BAD_ARG; }
    CK(cudaSetDevice(D->id));
    CK(launch_irradiance(d_cube, (int)src_size, (int)out_size, d_dst, (cudaStream_t)stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return HKBCN_OK;
}

int hkbcn_envprobe_prefilter(const float* cube, uint32_t src_size, int max_lod, float* dst)
{
    int rc = ensure_init();
    if (rc) return rc;
    if ((rc = probe_args_ok(src_size, max_lod)) != HKBCN_OK) return rc;
    if (!cube || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    const size_t in = (size_t)6 * src_size * src_size * 16, out = (size_t)envprobe_texels(max_lod) * 16;
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    if ((rc = with_scratch(D, in, out)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, cube, in, cudaMemcpyHostToDevice, D->stream));
    CK(launch_envprobe_prefilter((const float*)D->d_src, (int)src_size, max_lod, (float*)D->d_dst, D->stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CK(cudaMemcpyAsync(dst, D->d_dst, out, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

int hkbcn_irradiance(const float* cube, uint32_t src_size, uint32_t out_size, float* dst)
{
    int rc = ensure_init();
    if (rc) return rc;
    if ((rc = probe_args_ok(src_size, 1)) != HKBCN_OK) return rc;
    if (out_size < 1 || out_size > 1024) { set_err("bad irradiance size %u", out_size); return HKBCN_ERR_BAD_ARG; }
    if (!cube || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    const size_t in = (size_t)6 * src_size * src_size * 16, out = (size_t)6 * out_size * out_size * 16;
    Device* D = g_devs[0];
    std::lock_guard<std::mutex> lk(D->mu);
    if ((rc = with_scratch(D, in, out)) != HKBCN_OK) return rc;
    CK(cudaMemcpyAsync(D->d_src, cube, in, cudaMemcpyHostToDevice, D->stream));
    CK(launch_irradiance((const float*)D->d_src, (int)src_size, (int)out_size, (float*)D->d_dst, D->stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CK(cudaMemcpyAsync(dst, D->d_dst, out, cudaMemcpyDeviceToHost, D->stream));
    CK(cudaStreamSynchronize(D->stream));
    return HKBCN_OK;
}

// ---- .texture files (row f2) ------------------------------------------------------------------------------------------
static std::atomic<uint32_t> g_tmp_seq{0};
static void put_u32(uint8_t* p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }
static uint32_t get_u32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

int hkbcn_texture_header(const hkbcn_texture_desc* d, uint32_t payload_bytes, void* dst30)
{
    if (!d || !dst30) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    uint8_t* p = (uint8_t*)dst30;
    // MakeResourceMagic(RESOURCE_TEXTURE, 1) written with WriteUInt32: 'H', 'k', 4, 1
    p[0] = 'H'; p[1] = 'k'; p[2] = 4; p[3] = 1;
    p[4] = (uint8_t)d->type;
    put_u32(p + 5, d->width);
    put_u32(p + 9, d->height);
    put_u32(p + 13, d->depth);
    put_u32(p + 17, d->num_mips);
    p[21] = (uint8_t)d->fmt;
    put_u32(p + 22, d->storage_flags);
    put_u32(p + 26, payload_bytes);
    return HKBCN_OK;
}

int hkbcn_write_texture(const char* path, const hkbcn_texture_desc* d, const void* payload, size_t payload_bytes)
{
    if (!path || !d || (!payload && payload_bytes)) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    if (payload_bytes > 0xFFFFFFFFull) { set_err("payload of %zu bytes does not fit the format's u32 size field", payload_bytes); return HKBCN_ERR_BAD_ARG; }
    uint8_t h[HKBCN_TEXTURE_HEADER_BYTES];
    hkbcn_texture_header(d, (uint32_t)payload_bytes, h);
    // written beside the target and renamed into place: a reader (or a re-run with HKBCN_FLAG_SKIP_EXISTING after a crash)
    // never sees a half-written texture (checkpoint / resume of the batch driver)
    const std::string tmp = std::string(path) + ".tmp." + std::to_string((long)getpid()) + "." + std::to_string(g_tmp_seq.fetch_add(1));
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { set_err("cannot open %s for writing", tmp.c_str()); return HKBCN_ERR_BAD_ARG; }
    const bool ok = fwrite(h, 1, sizeof(h), f) == sizeof(h) && (payload_bytes == 0 || fwrite(payload, 1, payload_bytes, f) == payload_bytes);
    if (fclose(f) != 0 || !ok) { remove(tmp.c_str()); set_err("short write to %s", tmp.c_str()); return HKBCN_ERR_BAD_ARG; }
    if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); set_err("cannot rename %s to %s", tmp.c_str(), path); return HKBCN_ERR_BAD_ARG; }
    return HKBCN_OK;
}

int hkbcn_read_texture(const char* path, hkbcn_texture_desc* d, void* payload, size_t capacity, size_t* payload_bytes)
{
    if (!path || !d) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    FILE* f = fopen(path, "rb");
    if (!f) { set_err("cannot open %s", path); return HKBCN_ERR_BAD_ARG; }
    uint8_t h[HKBCN_TEXTURE_HEADER_BYTES];
    if (fread(h, 1, sizeof(h), f) != sizeof(h) || h[0] != 'H' || h[1] != 'k' || h[2] != 4 || h[3] != 1) {
        fclose(f);
        set_err("%s is not a version-1 texture resource", path);
        return HKBCN_ERR_BAD_FORMAT;
    }
    d->type = h[4];
    d->width = get_u32(h + 5);
    d->height = get_u32(h + 9);
    d->depth = get_u32(h + 13);
    d->num_mips = get_u32(h + 17);
    d->fmt = h[21];
    d->storage_flags = get_u32(h + 22);
    const size_t n = get_u32(h + 26);
    if (payload_bytes) *payload_bytes = n;
    int rc = HKBCN_OK;
    if (payload) {
        const size_t want = n < capacity ? n : capacity;
        if (fread(payload, 1, want, f) != want) { set_err("%s is truncated", path); rc = HKBCN_ERR_BAD_FORMAT; }
    }
    fclose(f);
    return rc;
}

int hkbcn_plan_import_batch(const hkbcn_import_job* jobs, size_t n_jobs, int n_gpus, int* device_of_job)
{
    if (n_gpus == 0) n_gpus = (int)g_devs.size();
    if (n_gpus <= 0 || (n_jobs && (!jobs || !device_of_job))) { set_err("bad arguments (n_gpus = 0 needs hkbcn_init first)"); return HKBCN_ERR_BAD_ARG; }
    std::vector<hkbcn_job> plain(n_jobs);
    for (size_t i = 0; i < n_jobs; i++) {
        if (codec_of(jobs[i].fmt) < 0) { set_err("job %zu: bad format %d", i, jobs[i].fmt); return HKBCN_ERR_BAD_FORMAT; }
        plain[i] = hkbcn_job{jobs[i].fmt, jobs[i].width, jobs[i].height * std::max(1u, jobs[i].slices), jobs[i].flags, nullptr, nullptr, 0};
    }
    plan_batch(plain.data(), n_jobs, n_gpus, device_of_job);
    return HKBCN_OK;
}

int hkbcn_batch_stats(double* busy_ms, int capacity)
{
    std::lock_guard<std::mutex> lk(g_batch_mu);
    for (int i = 0; i < capacity && i < (int)g_batch_ms.size(); i++) busy_ms[i] = g_batch_ms[(size_t)i];
    return (int)g_batch_ms.size();
}

// re-run of an interrupted batch: the output is there, is a texture resource of the job's shape and holds its whole payload
static bool complete_texture_exists(const hkbcn_import_job& j)
{
    hkbcn_texture_desc d;
    size_t n = 0;
    if (hkbcn_read_texture(j.out_path, &d, nullptr, 0, &n) != HKBCN_OK) return false;
    const uint32_t mips = (j.generate_mips || (j.flags & HKBCN_FLAG_SRC_CHAIN)) ? hkbcn_calc_num_mips(j.width, j.height) : 1u;
    const uint32_t slices = std::max(1u, j.slices);
    if (d.width != j.width || d.height != j.height || d.fmt != j.fmt || d.num_mips != mips || d.depth != slices) return false;
    if (n != hkbcn_chain_bytes(j.fmt, j.width, j.height, slices, mips)) return false;
    FILE* f = fopen(j.out_path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fclose(f);
    return sz == (long)(HKBCN_TEXTURE_HEADER_BYTES + n);
}

int hkbcn_import_batch(hkbcn_import_job* jobs, size_t n_jobs)
{
    Range nvtx_("hkbcn_import_batch");
    int rc = ensure_init();
    if (rc) return rc;
    if (n_jobs == 0) return HKBCN_OK;
    if (!jobs) { set_err("null jobs"); return HKBCN_ERR_BAD_ARG; }
    const size_t nd = g_devs.size();
    // plan as hkbcn_compress_batch does (pixel count x per-format weight; a mip chain adds a third everywhere alike)
    std::vector<hkbcn_job> plain(n_jobs);
    for (size_t i = 0; i < n_jobs; i++) {
        plain[i] = hkbcn_job{jobs[i].fmt, jobs[i].width, jobs[i].height * std::max(1u, jobs[i].slices), jobs[i].flags, nullptr, nullptr, 0};
        jobs[i].status = HKBCN_OK;
    }
    std::vector<int> assign(n_jobs);
    plan_batch(plain.data(), n_jobs, (int)nd, assign.data());
    std::vector<std::vector<size_t>> bins(nd);
    for (size_t i = 0; i < n_jobs; i++) bins[(size_t)assign[i]].push_back(i);

    // encoded chains queue up for one writer thread, so a GPU never waits for the file system
    struct Done { size_t idx; std::vector<uint8_t> own; const void* data; size_t bytes; uint32_t mips; };
    std::mutex qmu;
    std::condition_variable qcv;
    std::deque<Done> queue;
    size_t workers_left = nd;
    std::vector<int> first_err(nd + 1, 0);
    std::vector<std::string> err_msg(nd + 1);

    std::thread writer([&]() {
        for (;;) {
            Done d;
            {
                std::unique_lock<std::mutex> lk(qmu);
                qcv.wait(lk, [&] { return !queue.empty() || workers_left == 0; });
                if (queue.empty()) return;
                d = std::move(queue.front());
                queue.pop_front();
            }
            hkbcn_import_job& j = jobs[d.idx];
            const uint32_t nsl = std::max(1u, j.slices);
            hkbcn_texture_desc desc{nsl == 1 ? HKBCN_TEXTURE_2D : nsl == 6 ? HKBCN_TEXTURE_CUBE : HKBCN_TEXTURE_2D_ARRAY, j.width, j.height, nsl, d.mips, j.fmt,
                                    (j.mip.no_alpha ? (uint32_t)HKBCN_STORAGE_NO_ALPHA : 0u) | (j.mip.alpha_premultiplied ? (uint32_t)HKBCN_STORAGE_ALPHA_PREMULTIPLIED : 0u)};
            const int wrc = hkbcn_write_texture(j.out_path, &desc, d.data, d.bytes);
            if (wrc) {
                j.status = wrc;
                if (!first_err[nd]) { first_err[nd] = wrc; err_msg[nd] = t_err; }
            }
        }
    });
    std::vector<std::thread> th;
    std::vector<double> busy(nd, 0.0);
    for (size_t dv = 0; dv < nd; dv++) {
        th.emplace_back([&, dv]() {
            const auto t_begin = std::chrono::steady_clock::now();
            struct Stamp { double& out; std::chrono::steady_clock::time_point t0; ~Stamp() { out = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); } } stamp{busy[dv], t_begin};
            Device* D = g_devs[dv];
            std::lock_guard<std::mutex> blk(D->batch_mu);
            auto fail = [&](hkbcn_import_job& j, int rc) {
                j.status = rc;
                if (!first_err[dv]) { first_err[dv] = rc; err_msg[dv] = t_err; }
            };
            struct InFlight { long idx = -1; Done d; void* out = nullptr; bool staged = false; } fl[2];
            // wait for a slot's texture, move its chain out of the staging buffer and queue the file
            auto retire = [&](int s) {
                InFlight& f = fl[s];
                if (f.idx < 0) return;
                hkbcn_import_job& j = jobs[(size_t)f.idx];
                const cudaError_t e = cudaStreamSynchronize(D->slots[s].st);
                f.idx = -1;
                if (e != cudaSuccess) { set_err("import batch: %s", cudaGetErrorString(e)); fail(j, HKBCN_ERR_CUDA); return; }
                if (f.staged) memcpy(f.out, D->slots[s].h_dst, f.d.bytes);
                if (!j.out_path) return;
                f.d.data = f.out;
                {
                    std::lock_guard<std::mutex> lk(qmu);
                    queue.push_back(std::move(f.d));
                }
                qcv.notify_one();
            };
            auto submit = [&](int s, size_t idx) -> int {
                hkbcn_import_job& j = jobs[idx];
                Device::Slot& S = D->slots[s];
                int rc;
                const int codec = codec_of(j.fmt);
                if (codec < 0) { set_err("bad format %d", j.fmt); return HKBCN_ERR_BAD_FORMAT; }
                const bool prebuilt = (j.flags & HKBCN_FLAG_SRC_CHAIN) != 0; // the caller's own chain: upload all of it, generate nothing
                const uint32_t mips = (j.generate_mips || prebuilt) ? hkbcn_calc_num_mips(j.width, j.height) : 1u;
                if ((rc = check_chain_dims(j.width, j.height, mips)) != HKBCN_OK) return rc;
                if (!j.src_level0) { set_err("null source"); return HKBCN_ERR_BAD_ARG; }
                const uint32_t slices = std::max(1u, j.slices);
                if (slices > 1 && j.width != j.height && slices == 6) { set_err("cube faces must be square"); return HKBCN_ERR_BAD_ARG; }
                const size_t sbytes = chain_src_bytes(codec, j.flags, j.width, j.height, mips, slices);
                const size_t l0 = prebuilt ? sbytes : (size_t)j.width * j.height * src_bpp(codec, j.flags) * slices;
                const size_t dbytes = hkbcn_chain_bytes(j.fmt, j.width, j.height, slices, mips);
                CK(cudaSetDevice(D->id));
                if (!S.st) CK(cudaStreamCreateWithFlags(&S.st, cudaStreamNonBlocking));
                if ((rc = grow(&S.d_src, &S.dsc, sbytes)) != HKBCN_OK) return rc;
                if ((rc = grow(&S.d_dst, &S.ddc, dbytes)) != HKBCN_OK) return rc;
                InFlight& f = fl[s];
                f.d = Done{idx, {}, nullptr, dbytes, mips};
                f.out = j.dst;
                if (!f.out) {
                    f.d.own.resize(dbytes);
                    f.out = f.d.own.data();
                }
                const void* hsrc = j.src_level0;
                const bool src_dev = (j.flags & HKBCN_FLAG_SRC_DEVICE) != 0;
                if (!src_dev && !is_pinned(hsrc)) { // pageable memory: stage through pinned so that the copy engine, not this thread, feeds the GPU
                    if ((rc = grow_pinned(&S.h_src, &S.hsc, l0)) != HKBCN_OK) return rc;
                    memcpy(S.h_src, hsrc, l0);
                    hsrc = S.h_src;
                }
                f.staged = !is_pinned(f.out);
                if (f.staged && (rc = grow_pinned(&S.h_dst, &S.hdc, dbytes)) != HKBCN_OK) return rc;
                // level 0 only crosses PCIe -- or nothing does, when the source was produced on this GPU (HKBCN_FLAG_SRC_DEVICE)
                CK(cudaMemcpyAsync(S.d_src, hsrc, l0, src_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, S.st));
                if (mips > 1 && !prebuilt && (rc = generate_mips_on(D, j.fmt, S.d_src, j.width, j.height, mips, &j.mip, j.flags, S.st, slices)) != HKBCN_OK) return rc;
                EncodeParams P;
                if ((rc = build_chain(P, codec, j.flags, S.d_src, S.d_dst, j.width, j.height, mips, slices)) != HKBCN_OK) return rc;
                if ((rc = launch_encode(D, j.fmt, codec, P, S.st)) != HKBCN_OK) return rc;
                CK(cudaMemcpyAsync(f.staged ? S.h_dst : f.out, S.d_dst, dbytes, cudaMemcpyDeviceToHost, S.st));
                f.idx = (long)idx;
                return HKBCN_OK;
            };
            int s = 0;
            for (size_t idx : bins[dv]) {
                if ((jobs[idx].flags & HKBCN_FLAG_SKIP_EXISTING) && jobs[idx].out_path && !jobs[idx].dst && complete_texture_exists(jobs[idx])) {
                    jobs[idx].status = HKBCN_SKIPPED;
                    continue;
                }
                retire(s); // the slot's previous texture (two submissions ago)
                const int rc = submit(s, idx);
                if (rc) fail(jobs[idx], rc);
                s ^= 1;
            }
            retire(s);
            retire(s ^ 1);
            {
                std::lock_guard<std::mutex> lk(qmu);
                workers_left--;
            }
            qcv.notify_one();
        });
    }
    for (auto& t : th) t.join();
    writer.join();
    {
        std::lock_guard<std::mutex> lk(g_batch_mu);
        g_batch_ms = busy;
    }
    for (size_t d = 0; d <= nd; d++)
        if (first_err[d]) { set_err("%s", err_msg[d].c_str()); return first_err[d]; }
    return HKBCN_OK;
}

void* hkbcn_alloc_pinned(size_t bytes)
{
    if (ensure_init()) return nullptr;
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { set_err("cudaHostAlloc(%zu) failed", bytes); return nullptr; }
    return p;
}
void hkbcn_free_pinned(void* p) { if (p) cudaFreeHost(p); }

} // extern "C"

// ---- hkbcn_compress_static: the host-pointer compress with no allocation of any kind ----------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under graph capture, tight serving loops, harnesses that forbid it).  hkbcn_compress stages the surface and the blocks in
// device buffers it grows with cudaMalloc and takes the persistent kernels' tile counters from sched_slot(); here all three
// live in one file-scope device array that every GPU which loads the library holds until exit:
//
//   [0, 256)                    the {next tile, finished CTAs} pair of the encode kernel -- cleared at the start of every call
//   [256, 256 + src)            the source texels as uploaded
//   [.. rounded up to 256, +dst) the encoded blocks before they are copied back
//
// In host code `g_static_scratch` is a host-side stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, asked once per device with that device current (static_scratch_of).  Every call on a device shares the
// array and finds the previous call's bytes in it; calls that overlap on one GPU would race, so they are serialised on a mutex
// per device and each call ends with a stream synchronise.
static __device__ __align__(256) unsigned char g_static_scratch[HKBCN_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu[kStaticMaxDevices];
void*      g_static_ptr[kStaticMaxDevices]; // device address of g_static_scratch on CUDA device i (the array belongs to the GPU,
                                            // not to a slot of hkbcn_init's selection, which a later init may fill differently)

int static_scratch_of(int device, Device* D, unsigned char** out)
{
    // caller holds g_static_mu[device] (device = D->id) and has made D current
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            set_err("cudaGetSymbolAddress(g_static_scratch) failed on device %d: %s", D->id, cudaGetErrorString(e));
            return HKBCN_ERR_CUDA;
        }
        g_static_ptr[device] = p;
    }
    *out = (unsigned char*)g_static_ptr[device];
    return HKBCN_OK;
}
} // namespace

extern "C" {

size_t hkbcn_static_scratch_bytes(void) { return HKBCN_STATIC_SCRATCH_BYTES; }

int hkbcn_compress_static(int device, int fmt, const void* src, void* dst, uint32_t width, uint32_t height, uint32_t flags)
{
    Range nvtx_("hkbcn_compress_static");
    // hkbcn_init allocates (streams, decode tables): it must have run already -- this call does not do it on the caller's behalf
    if (!g_init.load(std::memory_order_acquire)) { set_err("hkbcn_compress_static allocates nothing: call hkbcn_init first"); return HKBCN_ERR_NOT_INIT; }
    Device* D = dev_at(device);
    if (!D || device >= kStaticMaxDevices) { set_err("bad device index %d", device); return HKBCN_ERR_BAD_ARG; }
    const int codec = codec_of(fmt);
    if (codec < 0) { set_err("bad format %d", fmt); return HKBCN_ERR_BAD_FORMAT; }
    if (!src || !dst) { set_err("null pointer"); return HKBCN_ERR_BAD_ARG; }
    if (flags & HKBCN_FLAG_SRC_DEVICE) { set_err("hkbcn_compress_static takes host pointers"); return HKBCN_ERR_BAD_ARG; }
    const size_t sbytes = (size_t)width * height * src_bpp(codec, flags);
    const size_t dbytes = (size_t)(width / 4) * (height / 4) * blk_bytes(codec);
    if (dbytes == 0) return HKBCN_OK;
    const size_t soff = 256, doff = (soff + sbytes + 255) & ~(size_t)255;
    // BC7 keeps a work list (two counters + 4 bytes per block) behind the blocks
    const size_t loff = (doff + dbytes + 255) & ~(size_t)255, lbytes = codec == C_BC7 ? bc7_aux_bytes((uint32_t)(dbytes / 16)) : 0;
    if (loff + lbytes > HKBCN_STATIC_SCRATCH_BYTES) {
        set_err("%u x %u texels of format %d need %zu bytes of staging; the static scratch holds %zu (HKBCN_STATIC_SCRATCH_BYTES) -- use hkbcn_compress",
                width, height, fmt, loff + lbytes, (size_t)HKBCN_STATIC_SCRATCH_BYTES);
        return HKBCN_ERR_BAD_ARG;
    }
    if (D->id < 0 || D->id >= kStaticMaxDevices) { set_err("CUDA device %d is beyond the %d the static scratch is tracked for", D->id, kStaticMaxDevices); return HKBCN_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> lk(g_static_mu[D->id]);
    CK(cudaSetDevice(D->id));
    unsigned char* base = nullptr;
    int rc = static_scratch_of(D->id, D, &base);
    if (rc) return rc;
    cudaStream_t st = D->stream;
    CK(cudaMemsetAsync(base, 0, 256, st)); // the tile counters must start at zero whatever the last call left
    CK(cudaMemcpyAsync(base + soff, src, sbytes, cudaMemcpyHostToDevice, st));
    EncodeParams P;
    memset(&P, 0, sizeof(P));
    P.flags = flags;
    // Width % 4 != 0: rows keep their full length, the last Width % 4 texels of each are never read (as hkbcn_compress)
    add_level(P, codec, flags, base + soff, base + doff, width & ~3u, height, width);
    if (P.flags & HKBCN_FLAG_LEVEL_MASK) {
        const int level = (int)((P.flags & HKBCN_FLAG_LEVEL_MASK) >> 4) - 1;
        const bool ok = (codec == C_BC1 || codec == C_BC2 || codec == C_BC3) ? level <= 5 : codec == C_BC7 ? level <= 4 : false;
        if (!ok) { set_err("quality level %d of format %d is not built", level, fmt); return HKBCN_ERR_UNSUPPORTED; }
    }
    P.sched = (uint32_t*)base;
    if (codec == C_BC7) {
        P.aux = (uint32_t*)(base + loff);
        CK(cudaMemsetAsync(P.aux, 0, 16, st)); // the list's counters: whatever an earlier call staged here must go
    }
    cudaError_t e = cudaSuccess;
    switch (codec) {
    case C_BC4: e = launch_bc45_encode(1, P, D->sm_count, st); break;
    case C_BC5: e = launch_bc45_encode(2, P, D->sm_count, st); break;
    case C_BC1: case C_BC2: case C_BC3: e = launch_bc123_encode(codec, P, D->sm_count, st); break;
    case C_BC7: e = launch_bc7_encode(P, D->sm_count, st); break;
    case C_BC6H: e = launch_bc6h_encode(fmt == HKBCN_BC6H_SFLOAT, (P.flags & HKBCN_FLAG_SRC_F16) != 0, P, D->sm_count, st); break;
    }
    if (e != cudaSuccess) { set_err("kernel launch failed: %s", cudaGetErrorString(e)); return HKBCN_ERR_CUDA; }
    g_launches.fetch_add(codec == C_BC7 ? 2 : 1, std::memory_order_relaxed);
    CK(cudaMemcpyAsync(dst, base + doff, dbytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return HKBCN_OK;
}

} // extern "C"
