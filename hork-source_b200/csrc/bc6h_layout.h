// bc6h_layout.h -- header bit layouts of the 14 BC6H modes, shared by the decoder (bcn_decode.cu) and the encoder
// (bc6h_thread.cuh / bc6h_encode.cu).  Host-side parse of the token strings into uint16 tokens.
#pragma once
#include <stdint.h>

namespace hkbcn {

// Header bit layouts of the 14 BC6H modes (D3D11 BC6H format), as token strings
// "<field><lsb>[:<count>][r]": field in {rw,rx,ry,rz,gw,..,bz,d}; the bits are consumed
// from the block LSB-first right after the 2/5 mode bits; 'r' = the bits arrive MSB-first
// (bcdec__bitstream_read_bits_r, bcdec.h:235-246).  Parsed once on the host into
// g_bc6h_layout (uint16 per token: field | lsb<<4 | count<<8 | rev<<12).
static const char* const kBc6hLayoutStr[14] = {
    /* 1  00    */ "gy4,by4,bz4,rw0:10,gw0:10,bw0:10,rx0:5,gz4,gy0:4,gx0:5,bz0,gz0:4,bx0:5,bz1,by0:4,ry0:5,bz2,rz0:5,bz3,d0:5",
    /* 2  01    */ "gy5,gz4,gz5,rw0:7,bz0,bz1,by4,gw0:7,by5,bz2,gy4,bw0:7,bz3,bz5,bz4,rx0:6,gy0:4,gx0:6,gz0:4,bx0:6,by0:4,ry0:6,rz0:6,d0:5",
    /* 3  00010 */ "rw0:10,gw0:10,bw0:10,rx0:5,rw10,gy0:4,gx0:4,gw10,bz0,gz0:4,bx0:4,bw10,bz1,by0:4,ry0:5,bz2,rz0:5,bz3,d0:5",
    /* 4  00110 */ "rw0:10,gw0:10,bw0:10,rx0:4,rw10,gz4,gy0:4,gx0:5,gw10,gz0:4,bx0:4,bw10,bz1,by0:4,ry0:4,bz0,bz2,rz0:4,gy4,bz3,d0:5",
    /* 5  01010 */ "rw0:10,gw0:10,bw0:10,rx0:4,rw10,by4,gy0:4,gx0:4,gw10,bz0,gz0:4,bx0:5,bw10,by0:4,ry0:4,bz1,bz2,rz0:4,bz4,bz3,d0:5",
    /* 6  01110 */ "rw0:9,by4,gw0:9,gy4,bw0:9,bz4,rx0:5,gz4,gy0:4,gx0:5,bz0,gz0:4,bx0:5,bz1,by0:4,ry0:5,bz2,rz0:5,bz3,d0:5",
    /* 7  10010 */ "rw0:8,gz4,by4,gw0:8,bz2,gy4,bw0:8,bz3,bz4,rx0:6,gy0:4,gx0:5,bz0,gz0:4,bx0:5,bz1,by0:4,ry0:6,rz0:6,d0:5",
    /* 8  10110 */ "rw0:8,bz0,by4,gw0:8,gy5,gy4,bw0:8,gz5,bz4,rx0:5,gz4,gy0:4,gx0:6,gz0:4,bx0:5,bz1,by0:4,ry0:5,bz2,rz0:5,bz3,d0:5",
    /* 9  11010 */ "rw0:8,bz1,by4,gw0:8,by5,gy4,bw0:8,bz5,bz4,rx0:5,gz4,gy0:4,gx0:5,bz0,gz0:4,bx0:6,by0:4,ry0:5,bz2,rz0:5,bz3,d0:5",
    /* 10 11110 */ "rw0:6,gz4,bz0,bz1,by4,gw0:6,gy5,by5,bz2,gy4,bw0:6,gz5,bz3,bz5,bz4,rx0:6,gy0:4,gx0:6,gz0:4,bx0:6,by0:4,ry0:6,rz0:6,d0:5",
    /* 11 00011 */ "rw0:10,gw0:10,bw0:10,rx0:10,gx0:10,bx0:10",
    /* 12 00111 */ "rw0:10,gw0:10,bw0:10,rx0:9,rw10,gx0:9,gw10,bx0:9,bw10",
    /* 13 01011 */ "rw0:10,gw0:10,bw0:10,rx0:8,rw10:2r,gx0:8,gw10:2r,bx0:8,bw10:2r",
    /* 14 01111 */ "rw0:10,gw0:10,bw0:10,rx0:4,rw10:6r,gx0:4,gw10:6r,bx0:4,bw10:6r",
};
#define BC6H_MAX_TOKENS 28

// tab[m][k] = field | lsb << 4 | count << 8 | rev << 12; 0xFFFF terminates
inline void bc6h_parse_layout(uint16_t tab[14][BC6H_MAX_TOKENS])
{
    for (int m = 0; m < 14; m++) {
        const char* s = kBc6hLayoutStr[m];
        int n = 0;
        while (*s) {
            int ch = (*s == 'r') ? 0 : (*s == 'g') ? 1 : (*s == 'b') ? 2 : 3;
            int field;
            if (ch == 3) { field = 12; s += 1; }
            else { int w = s[1] == 'w' ? 0 : s[1] == 'x' ? 1 : s[1] == 'y' ? 2 : 3; field = ch * 4 + w; s += 2; }
            int lsb = 0, cnt = 1, rev = 0;
            while (*s >= '0' && *s <= '9') lsb = lsb * 10 + (*s++ - '0');
            if (*s == ':') { s++; cnt = 0; while (*s >= '0' && *s <= '9') cnt = cnt * 10 + (*s++ - '0'); }
            if (*s == 'r') { rev = 1; s++; }
            if (*s == ',') s++;
            tab[m][n++] = (uint16_t)(field | (lsb << 4) | (cnt << 8) | (rev << 12));
        }
        for (; n < BC6H_MAX_TOKENS; n++) tab[m][n] = 0xFFFF;
    }
}

} // namespace hkbcn
