// GF(2)[x] / P(x) arithmetic for CRC-32C, shared by host and device code.
//
// Representation: the bit-reflected one Go's hash/crc32 (and every table-driven CRC-32C) uses — bit 31
// of a word is the coefficient of x^0, bit 0 is x^31.  In this form "feed one zero bit" is
// s -> (s >> 1) ^ (s & 1 ? POLY : 0), i.e. multiplication by x, so advancing a CRC register over n zero
// bytes is multiplication by x^(8n) mod P.  All combine / shard / segment logic rests on that one fact.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#endif
#if defined(__CUDA_ARCH__)
#define B200_UNROLL(n) _Pragma(#n)
#else
#define B200_UNROLL(n)
#endif

namespace b200 {

constexpr uint32_t kCastagnoli = 0x82F63B78u;  // Go: crc32.Castagnoli (reflected 0x1EDC6F41)
constexpr uint32_t kXpow0 = 0x80000000u;       // x^0
constexpr uint32_t kXpow8 = 0x00800000u;       // x^8

// a * b mod P
B200_HD uint32_t gf2_mul(uint32_t a, uint32_t b) {
    uint32_t r = 0;
B200_UNROLL(unroll 4)
    for (int i = 0; i < 32; i++) {
        r ^= b & (0u - (a >> 31));
        a <<= 1;
        b = (b >> 1) ^ (kCastagnoli & (0u - (b & 1u)));
    }
    return r;
}

// x^(8n) mod P with a caller supplied table sq[k] = x^(8 * 2^k) mod P, k < 64
B200_HD uint32_t gf2_xpow8(uint64_t nbytes, const uint32_t* sq) {
    uint32_t r = kXpow0;
    for (int k = 0; nbytes; k++, nbytes >>= 1)
        if (nbytes & 1) r = (r == kXpow0) ? sq[k] : gf2_mul(r, sq[k]);
    return r;
}

// one byte through the register, bit at a time (used for table construction and tiny heads/tails)
B200_HD uint32_t crc_step_byte(uint32_t s, uint32_t byte) {
    s ^= byte;
B200_UNROLL(unroll)
    for (int k = 0; k < 8; k++) s = (s >> 1) ^ (kCastagnoli & (0u - (s & 1u)));
    return s;
}

}  // namespace b200
