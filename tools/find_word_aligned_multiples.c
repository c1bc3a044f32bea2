/* Searches multiples of the CRC-32C generator P of the form
 *     Q(x) = y^m + sum_{j in J} y^j + x^r * sum_{i in I} y^i + 1,   y = x^32, 0 < r < 32, |J| + |I| = 4,
 * i.e. six terms whose exponents are word aligned except for ONE common bit shift r.  Folding a word stream
 * modulo such a Q needs one funnel shift and five XORs per 32-bit word and an m-word state.
 * P = (x+1) * (irreducible of degree 31); an even number of terms takes care of the factor x+1.
 * usage: find_word_aligned_multiples [max_m]
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define P_LOW 0x1EDC6F41u
static uint32_t xpow(int e) { uint32_t v = 1; for (int k = 0; k < e; k++) v = (v << 1) ^ ((v >> 31) ? P_LOW : 0); return v; }

typedef struct { uint32_t v; short a, b; } Pair;
static int cmp(const void *x, const void *y) { uint32_t a = ((const Pair*)x)->v, b = ((const Pair*)y)->v; return (a > b) - (a < b); }

int main(int argc, char **argv) {
    int M = argc > 1 ? atoi(argv[1]) : 72;
    /* pool index t: 0..M-1 aligned y^t (t>=1 used), M..2M-1 shifted x^r y^(t-M) */
    uint32_t *al = malloc((M + 1) * sizeof *al);
    for (int k = 0; k <= M; k++) al[k] = xpow(32 * k);
    uint32_t *sh = malloc(M * sizeof *sh);
    Pair *pairs = malloc(sizeof(Pair) * 4 * M * M);
    for (int r = 1; r < 32; r++) {
        for (int k = 0; k < M; k++) sh[k] = xpow(32 * k + r);
        for (int m = 8; m <= M; m++) {
            /* middle terms: aligned 1..m-1, shifted 0..m-2 */
            int n = 0;
            uint32_t pool[512]; short id[512];
            for (int j = 1; j < m; j++) { pool[n] = al[j]; id[n++] = (short)j; }
            for (int i = 0; i <= m - 2; i++) { pool[n] = sh[i]; id[n++] = (short)(1000 + i); }
            int np = 0;
            for (int a = 0; a < n; a++) for (int b = a + 1; b < n; b++) { pairs[np].v = pool[a] ^ pool[b]; pairs[np].a = id[a]; pairs[np].b = id[b]; np++; }
            qsort(pairs, np, sizeof(Pair), cmp);
            uint32_t target = al[m] ^ 1u;
            for (int p = 0; p < np; p++) {
                uint32_t want = pairs[p].v ^ target;
                int lo = 0, hi = np;
                while (lo < hi) { int mid = (lo + hi) / 2; if (pairs[mid].v < want) lo = mid + 1; else hi = mid; }
                for (int q = lo; q < np && pairs[q].v == want; q++) {
                    if (q <= p) continue;
                    short t[4] = {pairs[p].a, pairs[p].b, pairs[q].a, pairs[q].b};
                    if (t[0] == t[2] || t[0] == t[3] || t[1] == t[2] || t[1] == t[3]) continue;
                    int nsh = 0; for (int k = 0; k < 4; k++) nsh += t[k] >= 1000;
                    if (nsh == 0) continue; /* would be pure word aligned: report anyway below */
                    printf("m=%d r=%d terms:", m, r);
                    for (int k = 0; k < 4; k++) { if (t[k] >= 1000) printf(" x^%d*y^%d", r, t[k] - 1000); else printf(" y^%d", t[k]); }
                    printf("\n");
                }
            }
        }
    }
    return 0;
}
