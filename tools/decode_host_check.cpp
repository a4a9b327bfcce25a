// TEST / DEBUG INFRASTRUCTURE -- not part of the product, never linked into libhkbcn.so.
//
// Compiles the register-level block decoders of hork-source_b200/csrc/bcn_decode_lean.cuh (and bc7_decode_lean.cuh)
// for the CPU (PRMT emulated) and compares them with the reference's decoders (oracle/_ref/libhkref.so:
// ref_decode_block = bcdec_bc1..5 / bc7decomp::unpack_bc7) over per-channel exhaustive endpoint pairs and random blocks.
// usage: decode_host_check libhkref.so [random-blocks-per-format]
#define HKBCN_HOST_EMU 1
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../hork-source_b200/csrc/bcn_decode_lean.cuh"
#ifdef WITH_BC7
#include "../hork-source_b200/csrc/bc7_decode_lean.cuh"
#include "../hork-source_b200/csrc/bc6h_decode_lean.cuh"
#endif

using namespace hkbcn;
typedef void (*ref_decode_fn)(int, const void*, void*, size_t, int);
static ref_decode_fn ref_decode;

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd()
{
    uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static long check_rgba(int fmt, const uint32_t w[4])
{
    uint32_t px[16], want[16];
    switch (fmt) {
    case 0: lean::decode_bc1(w[0], w[1], px); break;
    case 1: lean::decode_bc2(w[0], w[1], w[2], w[3], px); break;
    case 2: lean::decode_bc3(w[0], w[1], w[2], w[3], px); break;
#ifdef WITH_BC7
    case 7: if (!lean7::decode_bc7_host(w[0], w[1], w[2], w[3], px)) return 0; break; // modes 0, 2, 4: general decoder
#endif
    }
    ref_decode(fmt, w, want, 16, 0);
    return memcmp(px, want, 64) != 0;
}
static long check_alpha(const uint32_t w[2])
{
    uint32_t row[4], want[4];
    lean::alpha_rows(w[0], w[1], row);
    ref_decode(3, w, want, 4, 0);
    return memcmp(row, want, 16) != 0;
}

#ifdef WITH_BC7
struct HostSink {
    uint32_t rows[4][6];
    void row(int r, const uint32_t w[6]) { memcpy(rows[r], w, 24); }
};
static long check_bc6h(const uint32_t w[4], bool isSigned)
{
    static lean7::Tables T;
    static bool init = false;
    static lean6::FTab F;
    if (!init) { lean7::fill_tables(&T, kShape2, kAnchor2); lean6::fill_ftab(&F, kShape2); init = true; }
    HostSink sink;
    if (isSigned) lean6::decode_bc6h<true>(w[0], w[1], w[2], w[3], T, F, sink);
    else lean6::decode_bc6h<false>(w[0], w[1], w[2], w[3], T, F, sink);
    uint32_t want[4][6];
    ref_decode(isSigned ? 6 : 5, w, want, 24, 0);
    return memcmp(sink.rows, want, 96) != 0;
}
#endif

int main(int argc, char** argv)
{
    if (argc < 2) { fprintf(stderr, "usage: %s libhkref.so [n]\n", argv[0]); return 2; }
    void* h = dlopen(argv[1], RTLD_NOW);
    if (!h) { fprintf(stderr, "%s\n", dlerror()); return 2; }
    ref_decode = (ref_decode_fn)dlsym(h, "ref_decode_block");
    const long n = argc > 2 ? atol(argv[2]) : 200000;
    long bad = 0, total = 0;
    // colour endpoints: every (c0, c1) pair on a lattice that is exhaustive per channel (32x32, 64x64, 32x32)
    for (uint32_t a = 0; a < 64; a++)
        for (uint32_t b = 0; b < 64; b++)
            for (int rep = 0; rep < 4; rep++) {
                const uint32_t c0 = ((a & 31) << 11) | (a << 5) | ((a * 7 + rep) & 31), c1 = ((b & 31) << 11) | (b << 5) | ((b * 5 + 3 * rep) & 31);
                const uint32_t c0b = (((a * 3 + rep) & 31) << 11) | (((a + rep) & 63) << 5) | (a & 31), c1b = (((b + rep) & 31) << 11) | (((b * 11) & 63) << 5) | (b & 31);
                uint32_t w[4] = {(uint32_t)rnd(), (uint32_t)rnd(), c0 | (c1 << 16), (uint32_t)rnd()};
                for (int f = 1; f <= 2; f++) { bad += check_rgba(f, w); total++; }
                uint32_t v[4] = {c0 | (c1 << 16), (uint32_t)rnd(), 0, 0};
                bad += check_rgba(0, v); total++;
                v[0] = c0b | (c1b << 16);
                bad += check_rgba(0, v); total++;
                v[0] = c0 | (c0 << 16); // equal endpoints: the 3-colour branch
                bad += check_rgba(0, v); total++;
            }
    // alpha endpoints: all 65536 pairs
    for (uint32_t a0 = 0; a0 < 256; a0++)
        for (uint32_t a1 = 0; a1 < 256; a1++) {
            uint32_t w[4] = {a0 | (a1 << 8) | ((uint32_t)rnd() << 16), (uint32_t)rnd(), (uint32_t)rnd(), (uint32_t)rnd()};
            bad += check_alpha(w); total++;
            bad += check_rgba(2, w); total++;
        }
    for (long i = 0; i < n; i++) {
        uint32_t w[4] = {(uint32_t)rnd(), (uint32_t)rnd(), (uint32_t)rnd(), (uint32_t)rnd()};
        for (int f = 0; f <= 2; f++) { bad += check_rgba(f, w); total++; }
        bad += check_alpha(w); total++;
#ifdef WITH_BC7
        bad += check_rgba(7, w); total++;
        bad += check_bc6h(w, false); total++;
        bad += check_bc6h(w, true); total++;
        {   // BC6H: force each of the 14 mode codes in turn, and extreme endpoint fields now and then
            static const uint32_t codes[14] = {0, 1, 2, 6, 10, 14, 18, 22, 26, 30, 3, 7, 11, 15};
            uint32_t h[4] = {w[0], w[1], w[2], w[3]};
            const int mi = (int)(i % 14);
            h[0] = mi < 2 ? ((h[0] & ~3u) | codes[mi]) : ((h[0] & ~31u) | codes[mi]);
            if ((i & 31) == 0) { h[0] |= 0xFFFFFFE0u; h[1] = 0xFFFFFFFFu; }
            if ((i & 31) == 1) { h[0] &= 31u; h[1] = 0; }
            bad += check_bc6h(h, false); total++;
            bad += check_bc6h(h, true); total++;
        }
        // force each mode in turn so the rare high modes get the same coverage
        uint32_t m[4] = {w[0], w[1], w[2], w[3]};
        const int mode = (int)(i & 7);
        m[0] = (m[0] & ~((2u << mode) - 1u)) | (1u << mode);
        bad += check_rgba(7, m); total++;
#endif
    }
#ifdef WITH_BC7
    { uint32_t z[4] = {0, 0x12345678u, 0x9abcdef0u, 0xffffffffu}; bad += check_rgba(7, z); total++; z[0] = 0xFFFFFF00u; bad += check_rgba(7, z); total++; }
#endif
    printf("%ld blocks, %ld differ\n", total, bad);
    return bad != 0;
}
