/* Searches for 4-term multiples x^a + x^b + x^c + 1 of the CRC-32C generator
 * P(x) = x^32 + x^28 + x^27 + x^26 + x^25 + x^23 + x^22 + x^20 + x^19 + x^18 + x^14 + x^13 + x^11 + x^10 + x^9 + x^8 + x^6 + 1
 * (0x11EDC6F41).  P has even weight, so (x+1) | P and no trinomial multiple exists; 4 terms is the
 * sparsest possible.  Folding a stream modulo such a multiple costs three shifted XORs per word and
 * no table lookups (kernel K2).
 * usage: find_sparse_multiples [max_degree]
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define P_LOW 0x1EDC6F41u
int main(int argc, char **argv) {
    int A = argc > 1 ? atoi(argv[1]) : 4096;
    uint32_t *r = malloc((A + 1) * sizeof *r);
    uint32_t v = 1;
    for (int k = 0; k <= A; k++) { r[k] = v; v = (v << 1) ^ ((v >> 31) ? P_LOW : 0); }
    /* open addressing map value -> exponent */
    int H = 1; while (H < 4 * (A + 1)) H <<= 1;
    int *slot = malloc(H * sizeof *slot);
    for (int i = 0; i < H; i++) slot[i] = -1;
    for (int k = 1; k <= A; k++) {
        uint32_t h = (r[k] * 2654435761u) & (H - 1);
        while (slot[h] >= 0) h = (h + 1) & (H - 1);
        slot[h] = k;
    }
    for (int a = 3; a <= A; a++)
        for (int b = 2; b < a; b++) {
            uint32_t t = r[a] ^ r[b] ^ 1u;
            uint32_t h = (t * 2654435761u) & (H - 1);
            while (slot[h] >= 0) {
                int c = slot[h];
                if (r[c] == t && c < b) printf("a=%d b=%d c=%d  a%%32=%d b%%32=%d c%%32=%d words=%d\n", a, b, c, a % 32, b % 32, c % 32, (a + 31) / 32);
                h = (h + 1) & (H - 1);
            }
        }
    return 0;
}
