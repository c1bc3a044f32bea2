/*
 * oracle.c — CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 * PARITY STATUS: unpinned by the reference (it has no tests); pinned here against public CRC-32C
 * known answers (RFC 3720 B.4) and numpy.histogram.  Each function cites the reference line or the
 * external module it restates.
 */
#define _GNU_SOURCE
#include "oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#if defined(__x86_64__)
#include <cpuid.h>
#include <nmmintrin.h>
#endif

#define POLY ORACLE_CASTAGNOLI_REFLECTED

/* ------------------------------------------------------------------------------------------
 * CRC-32C.  [EXT] Go hash/crc32: Update(crc, tab, p) = ^update(^crc, tab, p).
 * ---------------------------------------------------------------------------------------- */

/* The definition: reflected shift register, one bit at a time. */
uint32_t oracle_crc32c_bitwise(uint32_t crc, const void *p, size_t n) {
    const uint8_t *b = (const uint8_t *)p;
    uint32_t s = ~crc;
    for (size_t i = 0; i < n; i++) {
        s ^= b[i];
        for (int k = 0; k < 8; k++) s = (s >> 1) ^ (POLY & (0u - (s & 1u)));
    }
    return ~s;
}

static uint32_t tab8[8][256];
static pthread_once_t tab8_once = PTHREAD_ONCE_INIT;
static void tab8_build(void) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (POLY & (0u - (c & 1u)));
        tab8[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; i++)
        for (int t = 1; t < 8; t++) tab8[t][i] = (tab8[t - 1][i] >> 8) ^ tab8[0][tab8[t - 1][i] & 0xFF];
}

/* [EXT] Go's portable path for Castagnoli when SSE4.2 is absent: slicing-by-8. */
uint32_t oracle_crc32c_slice8(uint32_t crc, const void *p, size_t n) {
    pthread_once(&tab8_once, tab8_build);
    const uint8_t *b = (const uint8_t *)p;
    uint32_t s = ~crc;
    while (n >= 8) {
        uint32_t lo, hi;
        memcpy(&lo, b, 4);
        memcpy(&hi, b + 4, 4);
        lo ^= s;
        s = tab8[7][lo & 0xFF] ^ tab8[6][(lo >> 8) & 0xFF] ^ tab8[5][(lo >> 16) & 0xFF] ^ tab8[4][lo >> 24] ^
            tab8[3][hi & 0xFF] ^ tab8[2][(hi >> 8) & 0xFF] ^ tab8[1][(hi >> 16) & 0xFF] ^ tab8[0][hi >> 24];
        b += 8;
        n -= 8;
    }
    while (n--) s = (s >> 8) ^ tab8[0][(s ^ *b++) & 0xFF];
    return ~s;
}

/* GF(2)[x]/P arithmetic in the reflected representation: bit 31 is x^0, bit 0 is x^31. */
uint32_t oracle_gf2_mul(uint32_t a, uint32_t b) {
    uint32_t r = 0;
    for (int i = 31; i >= 0; i--) {           /* walk a from x^0 upwards */
        if ((a >> i) & 1u) r ^= b;
        b = (b >> 1) ^ (POLY & (0u - (b & 1u))); /* b *= x */
    }
    return r;
}

uint32_t oracle_xpow8(uint64_t nbytes) {
    uint32_t sq = 0x00800000u; /* x^8 */
    uint32_t r = 0x80000000u;  /* x^0 */
    while (nbytes) {
        if (nbytes & 1) r = oracle_gf2_mul(r, sq);
        sq = oracle_gf2_mul(sq, sq);
        nbytes >>= 1;
    }
    return r;
}

/* crc(A||B) = x^(8|B|) * crc(A) + crc(B): the all-ones init and final xor cancel between the terms. */
uint32_t oracle_crc32c_combine(uint32_t crc1, uint32_t crc2, uint64_t len2) {
    return oracle_gf2_mul(oracle_xpow8(len2), crc1) ^ crc2;
}

/* [EXT] Go's amd64 assembly for Castagnoli: SSE4.2 CRC32 instruction, three interleaved lanes for
 * long buffers, lanes merged by "advance by k zero bytes" lookup tables.  Restated in C: lane length
 * HW_LANE bytes; adv tables are built from oracle_xpow8. */
#define HW_LANE 4096u
static uint32_t hw_adv1[4][256], hw_adv2[4][256];
static pthread_once_t hw_once = PTHREAD_ONCE_INIT;
static void hw_build(void) {
    uint32_t k1 = oracle_xpow8(HW_LANE), k2 = oracle_xpow8(2 * HW_LANE);
    for (int t = 0; t < 4; t++)
        for (uint32_t v = 0; v < 256; v++) {
            hw_adv1[t][v] = oracle_gf2_mul(k1, v << (8 * t));
            hw_adv2[t][v] = oracle_gf2_mul(k2, v << (8 * t));
        }
}
static inline uint32_t adv_tab(const uint32_t t[4][256], uint32_t s) {
    return t[0][s & 0xFF] ^ t[1][(s >> 8) & 0xFF] ^ t[2][(s >> 16) & 0xFF] ^ t[3][s >> 24];
}

int oracle_crc32c_hw_available(void) {
#if defined(__x86_64__)
    unsigned a, b, c, d;
    if (!__get_cpuid(1, &a, &b, &c, &d)) return 0;
    return (c >> 20) & 1;
#else
    return 0;
#endif
}

#if defined(__x86_64__)
__attribute__((target("sse4.2"))) static uint32_t hw_update(uint32_t crc, const uint8_t *b, size_t n) {
    uint64_t s = (uint32_t)~crc;
    while (n && ((uintptr_t)b & 7)) { s = _mm_crc32_u8((uint32_t)s, *b++); n--; }
    while (n >= 3 * HW_LANE) {
        uint64_t s1 = 0, s2 = 0;
        const uint64_t *q0 = (const uint64_t *)b, *q1 = (const uint64_t *)(b + HW_LANE),
                       *q2 = (const uint64_t *)(b + 2 * HW_LANE);
        for (size_t i = 0; i < HW_LANE / 8; i++) {
            s = _mm_crc32_u64(s, q0[i]);
            s1 = _mm_crc32_u64(s1, q1[i]);
            s2 = _mm_crc32_u64(s2, q2[i]);
        }
        s = adv_tab(hw_adv2, (uint32_t)s) ^ adv_tab(hw_adv1, (uint32_t)s1) ^ (uint32_t)s2;
        b += 3 * HW_LANE;
        n -= 3 * HW_LANE;
    }
    while (n >= 8) { s = _mm_crc32_u64(s, *(const uint64_t *)b); b += 8; n -= 8; }
    while (n--) s = _mm_crc32_u8((uint32_t)s, *b++);
    return ~(uint32_t)s;
}
#endif

uint32_t oracle_crc32c_hw(uint32_t crc, const void *p, size_t n) {
#if defined(__x86_64__)
    if (oracle_crc32c_hw_available()) {
        pthread_once(&hw_once, hw_build);
        return hw_update(crc, (const uint8_t *)p, n);
    }
#endif
    return oracle_crc32c_slice8(crc, p, n);
}

/* [EXT] storage.Reader: at io.EOF, gotCRC != wantCRC -> "storage: bad CRC on read" (main.go:141-143).
 * Batched form: never aborts, flags each object. */
uint64_t oracle_verify_objects(const void *base, const uint64_t *off, const uint64_t *len, size_t nobj,
                               const uint32_t *want, uint32_t *got, uint8_t *ok) {
    uint64_t bad = 0;
    for (size_t i = 0; i < nobj; i++) {
        uint32_t c = oracle_crc32c_hw(0, (const uint8_t *)base + off[i], len[i]);
        if (got) got[i] = c;
        int good = want ? (c == want[i]) : 1;
        if (ok) ok[i] = (uint8_t)good;
        bad += !good;
    }
    return bad;
}

typedef struct {
    const void *base; const uint64_t *off, *len; const uint32_t *want; uint32_t *got; uint8_t *ok;
    size_t lo, hi; uint64_t bad;
} vjob;
static void *vjob_run(void *arg) {
    vjob *j = (vjob *)arg;
    j->bad = oracle_verify_objects(j->base, j->off + j->lo, j->len + j->lo, j->hi - j->lo,
                                   j->want ? j->want + j->lo : NULL, j->got ? j->got + j->lo : NULL,
                                   j->ok ? j->ok + j->lo : NULL);
    return NULL;
}
uint64_t oracle_verify_objects_mt(const void *base, const uint64_t *off, const uint64_t *len, size_t nobj,
                                  const uint32_t *want, uint32_t *got, uint8_t *ok, int threads) {
    if (threads < 1) threads = 1;
    if ((size_t)threads > nobj) threads = nobj ? (int)nobj : 1;
    pthread_t *th = (pthread_t *)calloc(threads, sizeof *th);
    vjob *jobs = (vjob *)calloc(threads, sizeof *jobs);
    for (int t = 0; t < threads; t++) {
        jobs[t] = (vjob){base, off, len, want, got, ok, nobj * t / threads, nobj * (t + 1) / threads, 0};
        pthread_create(&th[t], NULL, vjob_run, &jobs[t]);
    }
    uint64_t bad = 0;
    for (int t = 0; t < threads; t++) { pthread_join(th[t], NULL); bad += jobs[t].bad; }
    free(th); free(jobs);
    return bad;
}

/* ------------------------------------------------------------------------------------------
 * Synthetic inputs: integer-only so the device regenerates any slice bit-identically.
 * ---------------------------------------------------------------------------------------- */
uint64_t oracle_splitmix64(uint64_t seed, uint64_t index) {
    uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

void oracle_fill_payload(void *dst, uint64_t first_word, size_t nwords, uint64_t seed) {
    uint64_t *d = (uint64_t *)dst;
    for (size_t i = 0; i < nwords; i++) d[i] = oracle_splitmix64(seed, first_word + i); /* LE host */
}

int64_t oracle_latency_ns(uint64_t seed, uint64_t index) {
    uint64_t z = oracle_splitmix64(seed, 2 * index), z2 = oracle_splitmix64(seed, 2 * index + 1);
    uint64_t s = (z & 0xFFFF) + ((z >> 16) & 0xFFFF) + ((z >> 32) & 0xFFFF) + (z >> 48);
    uint64_t ns = 20000000ull + (s * 75000000ull) / 262141ull; /* bell-shaped bulk in [20 ms, 95 ms) */
    uint64_t q = z2 & 0xFF;
    if (q < 64) ns = (ns / 1000000ull) * 1000000ull;            /* whole milliseconds: exact duplicates */
    else if (q < 128) ns = (ns / 1000ull) * 1000ull;            /* whole microseconds */
    uint64_t sel = (z2 >> 8) % 100;
    if (sel == 0) ns = 95000000ull + (z2 >> 20) % (10000000000ull - 95000000ull); /* 1 % tail to 10 s */
    else if (sel == 1) ns = (z2 >> 20) % 20000000ull;                            /* 1 % under 20 ms   */
    return (int64_t)ns;
}

void oracle_fill_latency_ns(int64_t *dst, uint64_t first, size_t n, uint64_t seed) {
    for (size_t i = 0; i < n; i++) dst[i] = oracle_latency_ns(seed, first + i);
}

void oracle_ns_to_ms(const int64_t *ns, size_t n, int64_t *ms) {
    for (size_t i = 0; i < n; i++) ms[i] = ns[i] / 1000000; /* Duration.Milliseconds(): truncating */
}
void oracle_ns_to_us(const int64_t *ns, size_t n, int64_t *us) {
    for (size_t i = 0; i < n; i++) us[i] = ns[i] / 1000;    /* Duration.Microseconds() */
}

/* ------------------------------------------------------------------------------------------
 * Latency reduction.
 * ---------------------------------------------------------------------------------------- */

/* README.md:21-23,32.  [EXT] numpy.histogram with explicit edges: index by searchsorted(edges, v,
 * 'right') - 1, the right-most edge is inclusive, anything outside (and NaN) is not counted. */
uint64_t oracle_hist_numpy(const double *x, size_t n, const double *edges, int nedges, uint64_t *counts) {
    int nb = nedges - 1;
    uint64_t dropped = 0;
    for (int i = 0; i < nb; i++) counts[i] = 0;
    for (size_t i = 0; i < n; i++) {
        double v = x[i];
        if (!(v >= edges[0]) || !(v <= edges[nb])) { dropped++; continue; }
        int b = nb - 1;                              /* v == last edge lands in the last bin */
        for (int j = 1; j < nb; j++) if (v < edges[j]) { b = j - 1; break; }
        counts[b]++;
    }
    return dropped;
}

/* [EXT] ochttp.DefaultLatencyDistribution (go.opencensus.io v0.24.0), used at metrics_exporter.go:28 */
const double oracle_oc_default_latency_bounds[34] = {
    1, 2, 3, 4, 5, 6, 8, 10, 13, 16, 20, 25, 30, 40, 50, 65, 80, 100, 130, 160, 200, 250, 300, 400, 500, 650,
    800, 1000, 2000, 5000, 10000, 20000, 50000, 100000};

/* [EXT] view.DistributionData.addSample: linear scan, first bound strictly greater than v. */
void oracle_hist_opencensus(const double *x, size_t n, const double *bounds, int nbounds, uint64_t *counts) {
    for (int i = 0; i <= nbounds; i++) counts[i] = 0;
    for (size_t i = 0; i < n; i++) {
        int b = nbounds;
        for (int j = 0; j < nbounds; j++) if (x[i] < bounds[j]) { b = j; break; }
        counts[b]++;
    }
}

void oracle_stats_i64(const int64_t *v, size_t n, oracle_stats *o) {
    memset(o, 0, sizeof *o);
    if (!n) return;
    o->count = n;
    o->min = o->max = v[0];
    double mean = 0, ssd = 0, fsum = 0;
    uint64_t sum = 0;
    for (size_t i = 0; i < n; i++) {
        if (v[i] < o->min) o->min = v[i];
        if (v[i] > o->max) o->max = v[i];
        sum += (uint64_t)v[i];                  /* ssd_test/main.go:151-155, wrapping like Go int64 */
        double f = (double)v[i];                /* main.go:146 float64(duration.Milliseconds()) */
        double old = mean;                      /* [EXT] opencensus addSample (Welford) */
        mean = mean + (f - mean) / (double)(i + 1);
        ssd = ssd + (f - old) * (f - mean);
        fsum += f;                              /* README.md:30 sum(x) */
    }
    o->sum = sum;
    o->mean_welford = mean;
    o->sum_sq_dev = ssd;
    o->mean_seq = fsum / (double)n;
}

static int cmp_i64(const void *a, const void *b) {
    int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
    return (x > y) - (x < y);
}

/* ssd_test/main.go:147-163 */
int oracle_percentiles_i64(const int64_t *v, size_t n, oracle_percentiles *o) {
    if (!n) return -1; /* Go: integer divide by zero panic at :157 */
    int64_t *a = (int64_t *)malloc(n * sizeof *a);
    if (!a) return -2;
    memcpy(a, v, n * sizeof *a);
    qsort(a, n, sizeof *a, cmp_i64);                 /* :147-149 */
    int64_t sum = 0;
    for (size_t i = 0; i < n; i++) sum = (int64_t)((uint64_t)sum + (uint64_t)a[i]); /* :151-155 */
    o->avg = sum / (int64_t)n;                       /* :157 */
    o->p20 = a[n / 5];                               /* :158 */
    o->p50 = a[n / 2];                               /* :159 */
    o->p90 = a[(9 * n) / 10];                        /* :160 */
    o->p99 = a[(99 * n) / 100];                      /* :161 */
    o->min = a[0];                                   /* :162 */
    o->max = a[n - 1];                               /* :163 */
    free(a);
    return 0;
}

int oracle_select_kth_i64(const int64_t *v, size_t n, const uint64_t *ks, int nk, int64_t *out) {
    if (!n) return -1;
    int64_t *a = (int64_t *)malloc(n * sizeof *a);
    if (!a) return -2;
    memcpy(a, v, n * sizeof *a);
    qsort(a, n, sizeof *a, cmp_i64);
    for (int i = 0; i < nk; i++) {
        if (ks[i] >= n) { free(a); return -3; }
        out[i] = a[ks[i]];
    }
    free(a);
    return 0;
}
