/*
 * oracle.h — CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the checker or the timed CPU baseline.  The product
 * (custom-go-client-benchmark_b200/, libb200crc.so) never links, imports or calls it.
 *
 * PARITY STATUS: "parity unpinned by the reference".  The reference (Go) has no tests and no
 * golden vectors, and its CRC arithmetic lives in un-vendored modules
 *   - cloud.google.com/go/storage v1.33.0 (go.mod:6)  -> storage.Reader.Read keeps a running
 *     crc32.Update(gotCRC, castagnoliTable, p[:n]) and compares at EOF (call sites main.go:135,140,148)
 *   - Go stdlib hash/crc32 (go.mod:3, toolchain go 1.21)
 *   - go.opencensus.io v0.24.0 (go.mod:11) distribution aggregation (metrics_exporter.go:28)
 * none of which are on disk, and there is no Go toolchain in the image, so the reference cannot be
 * compiled here (no oracle/_ref).  The CRC restatement is pinned instead against the public CRC-32C
 * definition (RFC 3720 B.4 vectors, the "123456789" check value); the in-tree parts of the path
 * (README.md:15-35 histogram, benchmark-script/ssd_test/main.go:147-163 percentiles) are restated
 * line by line and the histogram is additionally pinned against numpy.histogram run in this image.
 */
#ifndef B200_ORACLE_H
#define B200_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- CRC-32C (Castagnoli), the arithmetic behind storage.Reader.Read (main.go:140) -------- */
#define ORACLE_CASTAGNOLI_REFLECTED 0x82F63B78u /* Go: crc32.Castagnoli */

/* Go hash/crc32 Update(crc, MakeTable(Castagnoli), p): pre/post conditioning is internal, so
 * Update(Update(0,a),b) == Update(0,a||b).  Three independent implementations: */
uint32_t oracle_crc32c_bitwise(uint32_t crc, const void *p, size_t n);  /* definition */
uint32_t oracle_crc32c_slice8(uint32_t crc, const void *p, size_t n);   /* Go's generic slicing-8 fallback */
uint32_t oracle_crc32c_hw(uint32_t crc, const void *p, size_t n);       /* Go's amd64 path: SSE4.2 crc32q, 3 interleaved lanes */
int oracle_crc32c_hw_available(void);

/* zlib-style combine on finalised CRCs: crc(A||B) from crc(A), crc(B), len(B). */
uint32_t oracle_crc32c_combine(uint32_t crc1, uint32_t crc2, uint64_t len2);
/* x^(8n) mod P in the bit-reflected representation (bit 31 = x^0). */
uint32_t oracle_xpow8(uint64_t nbytes);
uint32_t oracle_gf2_mul(uint32_t a, uint32_t b);

/* storage.Reader EOF check (reached from main.go:141-143): returns number of bad objects and fills
 * ok[i] = (crc(obj i) == want[i]).  Objects are base[off[i] .. off[i]+len[i]). */
uint64_t oracle_verify_objects(const void *base, const uint64_t *off, const uint64_t *len, size_t nobj,
                               const uint32_t *want, uint32_t *got, uint8_t *ok);
/* Same, on `threads` host threads (objects are independent; used as the all-cores CPU baseline). */
uint64_t oracle_verify_objects_mt(const void *base, const uint64_t *off, const uint64_t *len, size_t nobj,
                                  const uint32_t *want, uint32_t *got, uint8_t *ok, int threads);

/* ---- synthetic inputs (integer-only, counter based; identical on host and device) -------- */
uint64_t oracle_splitmix64(uint64_t seed, uint64_t index);
/* payload word i (little endian u64) = splitmix64(seed, first_word + i) */
void oracle_fill_payload(void *dst, uint64_t first_word, size_t nwords, uint64_t seed);
/* latency sample i in nanoseconds (time.Duration): 20..95 ms bulk, 1 % tail up to 10 s, duplicates */
int64_t oracle_latency_ns(uint64_t seed, uint64_t index);
void oracle_fill_latency_ns(int64_t *dst, uint64_t first, size_t n, uint64_t seed);

/* ---- latency reduction ------------------------------------------------------------------- */
/* README.md:21-23,32  plt.hist(x, bins) == numpy.histogram(x, bins): nedges-1 bins, [a,b) except the
 * last which is [a,b]; samples outside [edges[0], edges[last]] are dropped (returned count). */
uint64_t oracle_hist_numpy(const double *x, size_t n, const double *edges, int nedges, uint64_t *counts);
/* metrics_exporter.go:28 ochttp.DefaultLatencyDistribution: bucket = first i with v < bounds[i],
 * else the overflow bucket nbounds.  counts has nbounds+1 entries. */
void oracle_hist_opencensus(const double *x, size_t n, const double *bounds, int nbounds, uint64_t *counts);
extern const double oracle_oc_default_latency_bounds[34];

typedef struct {
    uint64_t count;
    int64_t min, max;
    uint64_t sum;      /* two's complement wrapping sum, as Go's int64 += (ssd_test/main.go:151-155) */
    double mean_welford;      /* OpenCensus DistributionData running mean (float64) */
    double sum_sq_dev;        /* OpenCensus SumOfSquaredDev */
    double mean_seq;          /* README.md:30  sum(x)/len(x), left to right float64 */
} oracle_stats;
/* Samples are int64 in `unit` of the caller's choosing; stats follow main.go:145-146 (float64 of an
 * integer millisecond count) when fed ms, and ssd_test/main.go:151-163 when fed microseconds. */
void oracle_stats_i64(const int64_t *v, size_t n, oracle_stats *out);

typedef struct {
    int64_t avg;  /* sum / size, Go integer division (ssd_test/main.go:157) */
    int64_t p20, p50, p90, p99, min, max; /* a[size/5], a[size/2], a[9*size/10], a[99*size/100], a[0], a[size-1] */
} oracle_percentiles;
/* ssd_test/main.go:147-163.  Sorts a private copy.  Returns -1 when n == 0 (Go panics there). */
int oracle_percentiles_i64(const int64_t *v, size_t n, oracle_percentiles *out);
/* k-th smallest (0-based) for each k in ks[], by sorting a private copy. */
int oracle_select_kth_i64(const int64_t *v, size_t n, const uint64_t *ks, int nk, int64_t *out);

/* main.go:146: duration.Milliseconds() — int64 ns / 1e6, truncating toward zero. */
void oracle_ns_to_ms(const int64_t *ns, size_t n, int64_t *ms);
/* ssd_test/main.go:79-80: totalTime.Microseconds() — ns / 1e3 truncating. */
void oracle_ns_to_us(const int64_t *ns, size_t n, int64_t *us);

#ifdef __cplusplus
}
#endif
#endif
