// Host-side construction of the GF(2) constant tables the kernels use.  Pure arithmetic on constants of
// the polynomial — no message byte is ever hashed on the CPU.
#include "crc_kernels.h"

namespace b200 {

void host_build_xpow8_sq(uint32_t out[64]) {
    uint32_t v = kXpow8;
    for (int k = 0; k < 64; k++) {
        out[k] = v;
        v = gf2_mul(v, v);
    }
}

uint32_t host_xpow8(uint64_t nbytes) {
    struct Squares {
        uint32_t sq[64];
        Squares() { host_build_xpow8_sq(sq); }
    };
    static const Squares s;  // thread-safe one-time initialisation
    return gf2_xpow8(nbytes, s.sq);
}

// out[t*256 + v] = x^(8*nbytes) * (v << 8t): slicing tables for "advance a register over nbytes zero bytes"
void host_build_adv_table(uint64_t nbytes, uint32_t out[4 * 256]) {
    const uint32_t k = host_xpow8(nbytes);
    for (int t = 0; t < 4; t++) {
        // products of the 8 single-bit inputs of this byte lane, then xor-combine
        uint32_t bit[8];
        for (int b = 0; b < 8; b++) bit[b] = gf2_mul(k, 1u << (8 * t + b));
        for (uint32_t v = 0; v < 256; v++) {
            uint32_t r = 0;
            for (int b = 0; b < 8; b++)
                if (v & (1u << b)) r ^= bit[b];
            out[t * 256 + v] = r;
        }
    }
}

void host_build_col_k(uint32_t out[128]) {
    for (int c = 0; c < kRowWords; c++) out[c] = host_xpow8(4ull * (kRowWords - c));
}

void host_build_xpow128(uint32_t* out) {
    const uint32_t step = host_xpow8(16);
    out[0] = kXpow0;
    for (uint32_t j = 1; j < kXpow128Entries; j++) out[j] = gf2_mul(out[j - 1], step);
}

void host_build_byte_tab(uint32_t out[256]) {
    for (uint32_t v = 0; v < 256; v++) out[v] = crc_step_byte(0, v);
}

}  // namespace b200
