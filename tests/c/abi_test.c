/* Plain C11 caller of libb200crc.so — what cgo compiles include/b200crc.h as.  Known answers (RFC 3720 B.4 and the
 * "123456789" check value), chunked Update, the hash.Hash32 entry points, and a VerifyObjects batch on host buffers
 * against hard-coded goldens with one corrupted object.  Prints "abi_test ok" and exits 0 on success.
 *   gcc -std=c11 -Wall -Wextra -pedantic -Iinclude tests/c/abi_test.c -Lcustom-go-client-benchmark_b200 -lb200crc
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200crc.h"

#define CHECK(call)                                                                              \
    do {                                                                                         \
        int rc_ = (call);                                                                        \
        if (rc_ != 0) {                                                                          \
            fprintf(stderr, "%s -> %s: %s\n", #call, b200_strerror(rc_), b200_last_error());     \
            return 1;                                                                            \
        }                                                                                        \
    } while (0)
#define EXPECT(cond)                                          \
    do {                                                      \
        if (!(cond)) {                                        \
            fprintf(stderr, "FAILED: %s (line %d)\n", #cond, __LINE__); \
            return 1;                                         \
        }                                                     \
    } while (0)

/* splitmix64(seed, index): the payload generator of the oracle, restated (oracle/oracle.c is not linked here) */
static uint64_t splitmix64(uint64_t seed, uint64_t index) {
    uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/* bitwise CRC-32C (reflected 0x82F63B78): the definition itself, slow and obviously right */
static uint32_t crc32c_bitwise(uint32_t crc, const unsigned char* p, size_t n) {
    crc = ~crc;
    for (size_t i = 0; i < n; i++) {
        crc ^= p[i];
        for (int k = 0; k < 8; k++) crc = (crc >> 1) ^ (0x82F63B78u & (0u - (crc & 1u)));
    }
    return ~crc;
}

int main(void) {
    b200_ctx* ctx = NULL;
    int dev = 0;
    CHECK(b200_init(1, &dev, &ctx));
    uint32_t c = 1;

    /* known answers */
    CHECK(b200_crc32c_update_host(ctx, 0, 0, "", 0, &c));
    EXPECT(c == 0);
    CHECK(b200_crc32c_update_host(ctx, 0, 0, "123456789", 9, &c));
    EXPECT(c == 0xE3069283u);
    unsigned char buf32[32];
    memset(buf32, 0, 32);
    CHECK(b200_crc32c_update_host(ctx, 0, 0, buf32, 32, &c));
    EXPECT(c == 0x8A9136AAu);
    memset(buf32, 0xFF, 32);
    CHECK(b200_crc32c_update_host(ctx, 0, 0, buf32, 32, &c));
    EXPECT(c == 0x62A8AB43u);
    for (int i = 0; i < 32; i++) buf32[i] = (unsigned char)i;
    CHECK(b200_crc32c_update_host(ctx, 0, 0, buf32, 32, &c));
    EXPECT(c == 0x46DD794Eu);
    for (int i = 0; i < 32; i++) buf32[i] = (unsigned char)(31 - i);
    CHECK(b200_crc32c_update_host(ctx, 0, 0, buf32, 32, &c));
    EXPECT(c == 0x113FDB5Cu);

    /* chunked Update == one shot == the definition, on 1 MiB + 37 bytes of splitmix payload */
    const size_t n = (1u << 20) + 37;
    unsigned char* data = malloc(n + 8);
    EXPECT(data != NULL);
    for (size_t w = 0; w * 8 < n + 8; w++) {
        const uint64_t v = splitmix64(0xC0FFEE, w);
        memcpy(data + w * 8, &v, 8 > n + 8 - w * 8 ? n + 8 - w * 8 : 8);
    }
    const uint32_t want = crc32c_bitwise(0, data, n);
    CHECK(b200_crc32c_update_host(ctx, 0, 0, data, n, &c));
    EXPECT(c == want);
    uint32_t run = 0;
    for (size_t off = 0; off < n; off += 37000) {
        const size_t m = n - off < 37000 ? n - off : 37000;
        CHECK(b200_crc32c_update_host(ctx, 0, run, data + off, m, &run));
    }
    EXPECT(run == want);
    EXPECT(b200_crc32c_combine(crc32c_bitwise(0, data, 1000), crc32c_bitwise(0, data + 1000, n - 1000), n - 1000) == want);

    /* hash.Hash32 */
    b200_hasher* h = NULL;
    CHECK(b200_hasher_new(ctx, 0, &h));
    CHECK(b200_hasher_write(h, data, 12345));
    CHECK(b200_hasher_write(h, data + 12345, n - 12345));
    CHECK(b200_hasher_sum32(h, &c));
    EXPECT(c == want);
    CHECK(b200_hasher_reset(h));
    CHECK(b200_hasher_write(h, "123456789", 9));
    CHECK(b200_hasher_sum32(h, &c));
    EXPECT(c == 0xE3069283u);
    CHECK(b200_hasher_free(h));

    /* VerifyObjects on host buffers: 5 objects, one corrupted; flags exactly that one, never aborts */
    const void* bufs[5] = {"123456789", data, data + 3, "", data + 4096};
    const uint64_t len[5] = {9, n, 100000, 0, 512};
    uint32_t wantv[5] = {0xE3069283u, want, 0, 0, 0}, got[5];
    wantv[2] = crc32c_bitwise(0, data + 3, 100000);
    wantv[4] = crc32c_bitwise(0, data + 4096, 512) ^ 1u; /* wrong on purpose */
    uint8_t ok[5];
    uint64_t nbad = 99;
    CHECK(b200_crc32c_batch_host(ctx, 0, bufs, len, 5, wantv, got, ok, &nbad));
    EXPECT(nbad == 1 && ok[0] && ok[1] && ok[2] && ok[3] && !ok[4]);
    EXPECT(got[4] == (wantv[4] ^ 1u) && got[3] == 0);

    /* errors are statuses, not crashes */
    EXPECT(b200_crc32c_update_host(ctx, 0, 0, NULL, 10, &c) == B200_ERR_INVALID);
    EXPECT(b200_crc32c_update_host(NULL, 0, 0, data, 10, &c) == B200_ERR_INVALID);
    free(data);
    CHECK(b200_shutdown(ctx));
    puts("abi_test ok");
    return 0;
}
