/* TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or called from the product.
 *
 * oracle.h: plain-C restatement of the reference's BCn algorithms (serial, one block at a
 * time, in the reference's own control flow).  Each function cites the reference lines it
 * follows.  Pinned against the reference itself (oracle/_ref/libhkref.so, built from the
 * unmodified sources by oracle/Makefile) and against committed golden vectors generated from
 * it (tests/golden/, tools/gen_golden.py) -- the reference tree holds no tests or fixtures
 * of its own.
 *
 * BC7 / BC6H ENCODE are not restated: their parity bar is PSNR against the reference's own
 * encoder, and that encoder (bc7enc.cpp, bc6h_enc.h) compiles unmodified into oracle/_ref,
 * which is therefore the oracle for those two rows; this file restates their DECODERS, which
 * is what the PSNR comparison needs. */
#ifndef HK_ORACLE_H
#define HK_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
/* bit-exact encoders (restating rgbcx.cpp) */
void oracle_encode_bc4_hq(const uint8_t* px, uint32_t stride, uint8_t* dst);
void oracle_encode_bc4_fast(const uint8_t* px, uint32_t stride, uint8_t* dst);
void oracle_encode_bc1(const uint8_t* rgba, uint8_t* dst);
void oracle_encode_bc2(const uint8_t* rgba, uint8_t* dst);
void oracle_encode_bc3(const uint8_t* rgba, uint8_t* dst);
void oracle_encode_bc5(const uint8_t* rg, uint8_t* dst);
/* decoders (restating bcdec.h / bc7decomp.cpp); pitch in bytes */
void oracle_decode_bc1(const uint8_t* blk, uint8_t* dst, size_t pitch);
void oracle_decode_bc2(const uint8_t* blk, uint8_t* dst, size_t pitch);
void oracle_decode_bc3(const uint8_t* blk, uint8_t* dst, size_t pitch);
void oracle_decode_bc4(const uint8_t* blk, uint8_t* dst, size_t pitch);
void oracle_decode_bc5(const uint8_t* blk, uint8_t* dst, size_t pitch);
void oracle_decode_bc7(const uint8_t* blk, uint8_t* dst, size_t pitch);
void oracle_decode_bc6h(const uint8_t* blk, uint8_t* dst, size_t pitch, int is_signed, int f32out);
/* surfaces: fmt = 0..7 (BC1,BC2,BC3,BC4,BC5,BC6H_U,BC6H_S,BC7); returns 0, or -1 if fmt has no
 * restated encoder (BC6H, BC7) */
int oracle_compress(int fmt, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h);
int oracle_decompress(int fmt, const uint8_t* src, uint8_t* dst, uint32_t w, uint32_t h, size_t row_stride, int f32out);
#ifdef __cplusplus
}
#endif
#endif
