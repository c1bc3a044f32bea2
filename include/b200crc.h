/*
 * b200crc.h — C ABI of libb200crc.so: CRC-32C verification and latency reduction on NVIDIA B200.
 *
 * This is the drop-in boundary for the one data-parallel path of tritone/custom-go-client-benchmark:
 *   - the CRC-32C integrity check storage.Reader runs on every downloaded byte
 *     (call sites main.go:135 NewReader, main.go:140 io.CopyBuffer -> Read, main.go:148 Close; the
 *     arithmetic itself is Go stdlib hash/crc32 inside cloud.google.com/go/storage v1.33.0, go.mod:6),
 *   - the per-read latency sample reduction (main.go:145-146, metrics_exporter.go:22-34,
 *     README.md:15-35, benchmark-script/ssd_test/main.go:147-163).
 * The reference exposes no FFI of its own for this path; these entry points are what a cgo binding of
 * that path binds (see INTEGRATION.md and custom-go-client-benchmark_b200/go/crc32c).
 *
 * Conventions: plain C, no exceptions cross the boundary.  Every function returns 0 on success or a
 * negative b200_status; b200_strerror() names it and b200_last_error() gives the CUDA detail for the
 * calling thread.  Every entry point selects its device itself (goroutines migrate between OS threads),
 * and is safe to call from many threads at once; calls that target the same device serialise.
 * Host pointers passed in are never retained after the call returns (cgo pointer rule).
 * There is no CPU fallback: without a usable GPU every compute entry point fails with B200_ERR_CUDA.
 */
#ifndef B200CRC_H
#define B200CRC_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    B200_OK = 0,
    B200_ERR_INVALID = -1,   /* bad argument (null pointer, misaligned arena, k out of range, ...) */
    B200_ERR_CUDA = -2,      /* CUDA runtime / driver failure, see b200_last_error() */
    B200_ERR_NOMEM = -3,     /* host or device allocation failed */
    B200_ERR_EMPTY = -4,     /* order statistic of an empty sample set (Go panics at ssd_test/main.go:157) */
    B200_ERR_TOO_LARGE = -5, /* size beyond a documented fixed capacity */
} b200_status;

typedef struct b200_ctx b200_ctx;       /* owns streams, tables and scratch on each device it was given */
typedef struct b200_hasher b200_hasher; /* streaming state behind Go's hash.Hash32 */
typedef struct b200_batch b200_batch;   /* a prepared VerifyObjects plan for device-resident objects */

const char* b200_strerror(int status);
const char* b200_last_error(void);
int b200_device_count(int* n);
/* PCI bus id ("0000:1b:00.0") of CUDA device `dev`, e.g. to find the same GPU in NVML or its NUMA node in sysfs. */
int b200_device_pci_bus_id(int dev, char* buf, int len);

/* devs == NULL means devices 0..ndev-1.  Replaces nothing in the reference (it has no device). */
int b200_init(int ndev, const int* devs, b200_ctx** out);
int b200_shutdown(b200_ctx* ctx);
int b200_ctx_ndev(const b200_ctx* ctx);
/* CUDA-event stopwatch on the device's compute stream, for timing a region of enqueued work on the device:
 * start drains the stream and records; stop records, waits and returns the elapsed milliseconds. */
int b200_timer_start(b200_ctx* ctx, int dev);
int b200_timer_stop(b200_ctx* ctx, int dev, float* ms);

/* Library-owned memory, so that Go can view it with unsafe.Slice and C keeps no Go pointers.
 * The pinned buffer replaces ReadObject's `buf := make([]byte, 2*MB)` (main.go:125). */
int b200_alloc_pinned(b200_ctx* ctx, size_t n, void** host);
int b200_free_pinned(b200_ctx* ctx, void* host);
int b200_alloc_device(b200_ctx* ctx, int dev, size_t n, void** dptr);
int b200_free_device(b200_ctx* ctx, int dev, void* dptr);
int b200_memcpy_h2d(b200_ctx* ctx, int dev, void* dptr, const void* host, size_t n);
int b200_memcpy_d2h(b200_ctx* ctx, int dev, void* host, const void* dptr, size_t n);

/* ---- CRC-32C ------------------------------------------------------------------------------------ */

/* crc32.Update(crc, castagnoliTable, p[:n]) — the statement storage.Reader.Read executes per chunk
 * (reached from main.go:140).  Synchronous; p is host memory and is not retained. */
int b200_crc32c_update_host(b200_ctx* ctx, int dev, uint32_t crc, const void* p, size_t n, uint32_t* out);
/* Same on a device-resident buffer (16-byte alignment not required). */
int b200_crc32c_update_device(b200_ctx* ctx, int dev, uint32_t crc, const void* dptr, size_t n, uint32_t* out);

/* No counterpart in the reference.  Same result as b200_crc32c_update_device, for callers that cannot
 * allocate device memory: processes that gave all of it to their own pool at start-up, and read loops
 * that cannot take the device-wide synchronisation of cudaMalloc/cudaFree.  Needs no b200_ctx and calls no
 * allocating routine: the row kernel's per-segment scratch and the constant tables live in a fixed ~270 KB
 * array inside the library image, which every GPU that loads the library holds until exit.
 *   - n > B200_NOALLOC_MAX_BYTES is rejected with B200_ERR_TOO_LARGE (split and b200_crc32c_combine).
 *   - All calls share that one array per GPU; they are serialised inside the library and each waits for
 *     its result, so they never overlap on the GPU — overlapping users of the array would race.  The array
 *     keeps the previous call's contents; what must start at zero is cleared on every call.
 *   - The first call on a GPU also does that GPU's one-time set-up (resolves the array's device address,
 *     uploads the tables); make it before starting any stream capture.
 *   - Runs on the device's legacy default stream (a copy, the kernels, a blocking copy back) and returns once
 *     *out (host memory) holds the CRC.  Because of that no call — not only the first — may be made while any
 *     thread of the process has a global-mode stream capture open.
 *   - The calling thread's current CUDA device is restored before returning. */
#define B200_NOALLOC_MAX_BYTES (1ull << 35)
int b200_crc32c_update_device_noalloc(int dev, uint32_t crc, const void* dptr, size_t n, uint32_t* out);

/* hash.Hash32 as returned by crc32.New(crc32.MakeTable(crc32.Castagnoli)): Write never fails and
 * consumes all of p; Sum32 does not disturb the state; Reset returns to the empty message. */
int b200_hasher_new(b200_ctx* ctx, int dev, b200_hasher** out);
int b200_hasher_write(b200_hasher* h, const void* p, size_t n);
/* io.ReaderFrom support: instead of io.CopyBuffer(io.Discard, rc, buf) filling a scratch buffer that is then
 * hashed (main.go:125,140), the transport reads straight into the hasher's pinned staging.  reserve returns
 * the free part of that buffer (*avail >= 1; flushes to the GPU first when it is full), commit(n) declares
 * that n <= *avail bytes were written there.  The pointer is library-owned pinned memory, valid until the
 * next call on this hasher. */
int b200_hasher_reserve(b200_hasher* h, void** ptr, size_t* avail);
int b200_hasher_commit(b200_hasher* h, size_t n);
int b200_hasher_sum32(b200_hasher* h, uint32_t* out);
int b200_hasher_reset(b200_hasher* h);
int b200_hasher_free(b200_hasher* h);

/* VerifyObjects on objects already in device memory (BASELINE config 2).  Object i is the byte range
 * base[off[i] .. off[i]+len[i]).  off/len/want/got/ok are HOST arrays; base is a 16-byte aligned device
 * pointer.  got[i] = CRC-32C, ok[i] = (got[i] == want[i]) (1 when want == NULL), *nbad = number of
 * mismatches.  This is storage.Reader's EOF check "storage: bad CRC on read: got %d, want %d"
 * (surfacing at main.go:141-143) for a whole batch; a bad object is flagged, never aborts the batch.
 * got / ok / nbad may be NULL. */
int b200_crc32c_batch_device(b200_ctx* ctx, int dev, const void* base, const uint64_t* off, const uint64_t* len,
                             size_t nobj, const uint32_t* want, uint32_t* got, uint8_t* ok, uint64_t* nbad);
/* The same split in two so a caller can plan once and verify many arenas of one shape
 * (the read loop re-reads one object M times, main.go:127): create uploads off/len/want and builds the
 * segment plan; run launches the kernels on the device stream and waits; results copies out. */
int b200_batch_create(b200_ctx* ctx, int dev, const uint64_t* off, const uint64_t* len, size_t nobj,
                      const uint32_t* want, b200_batch** out);
int b200_batch_run(b200_batch* b, const void* base);
int b200_batch_run_async(b200_batch* b, const void* base); /* enqueue only; pair with b200_batch_wait */
int b200_batch_wait(b200_batch* b);
int b200_batch_results(b200_batch* b, uint32_t* got, uint8_t* ok, uint64_t* nbad);
/* Mean CUDA-event duration (ms) of the runs since the previous call (at most the last 256): the whole run
 * and its row kernel alone.  Waits for the stream. */
int b200_batch_kernel_ms(b200_batch* b, float* total_ms, float* rows_kernel_ms);
int b200_batch_launches(const b200_batch* b, int* launches_per_run);
int b200_batch_free(b200_batch* b);
/* VerifyObjects([]buf, []want) with HOST buffers (the Go-facing form): bufs[i] points at len[i] bytes.
 * Buffers are packed into pinned staging, copied to the device in pipelined slabs and verified there. */
int b200_crc32c_batch_host(b200_ctx* ctx, int dev, const void* const* bufs, const uint64_t* len, size_t nobj,
                           const uint32_t* want, uint32_t* got, uint8_t* ok, uint64_t* nbad);

/* One large object sharded over the context's devices (BASELINE config 3): device d holds
 * len_per_dev[d] bytes at dptr_per_dev[d], shards are in message order.  Per-device CRCs are folded on
 * the host with b200_crc32c_combine; there is no GPU<->GPU exchange. */
int b200_crc32c_sharded(b200_ctx* ctx, const void* const* dptr_per_dev, const uint64_t* len_per_dev, int ndev,
                        uint32_t* out);
/* Host-resident streaming (BASELINE config 4): n bytes of (ideally pinned) host memory are copied in
 * `chunk`-byte pieces with cudaMemcpyAsync, overlapped with the CRC of the previous piece.  PCIe-bound. */
int b200_crc32c_stream_host(b200_ctx* ctx, int dev, uint32_t crc, const void* host, size_t n, size_t chunk,
                            uint32_t* out);
/* crc(A||B) from crc(A), crc(B), len(B) — zlib crc32_combine semantics on finalised CRCs.  Pure host. */
uint32_t b200_crc32c_combine(uint32_t crc1, uint32_t crc2, uint64_t len2);

/* ---- synthetic inputs generated on the device (integer-only, identical to the host oracle) ------- */
int b200_fill_payload_device(b200_ctx* ctx, int dev, void* dptr, uint64_t first_word, size_t nwords, uint64_t seed);
int b200_fill_latency_ns_device(b200_ctx* ctx, int dev, int64_t* dptr, uint64_t first, size_t n, uint64_t seed);

/* ---- latency reduction --------------------------------------------------------------------------- */

typedef struct {
    uint64_t count;
    int64_t min, max;        /* of the scaled samples v = trunc(ns / divisor) */
    uint64_t sum;            /* low 64 bits of the sum of v: Go's wrapping int64 += (ssd_test/main.go:151-155) */
    int64_t sum_hi;          /* high 64 bits of the exact 128-bit sum */
    uint64_t sumsq_lo, sumsq_hi; /* exact 128-bit sum of v*v (mean / SumOfSquaredDev follow exactly) */
    uint64_t dropped_below, dropped_above; /* numpy mode: samples outside [edges[0], edges[last]] */
} b200_lat_stats;

#define B200_HIST_NUMPY 0      /* README.md:21-23,32: numpy.histogram, [a,b) bins, last bin closed */
#define B200_HIST_OPENCENSUS 1 /* metrics_exporter.go:28: first bound strictly greater, overflow bucket */
#define B200_HIST_MAX_EDGES 64
#define B200_RADIX_BINS 4096
#define B200_SELECT_MAX_K 4

/* samples: int64 nanoseconds on the device (time.Duration).  Each is divided by `divisor` truncating
 * toward zero (1000000 = Duration.Milliseconds(), main.go:146; 1000 = Microseconds(),
 * ssd_test/main.go:79) and compared as float64 with the edges, exactly as the reference does.  counts has
 * nedges-1 entries in numpy mode and nedges+1 in OpenCensus mode (HOST array).  stats may be NULL.
 * Exact for |v| < 2^53 (float64(v) is then exact). */
int b200_hist_fixed(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                    const double* edges, int nedges, int mode, uint64_t* counts, b200_lat_stats* stats);
/* Two histograms of the same samples with the same divisor from ONE pass (for instance the README bins and the
 * OpenCensus distribution of a run, README.md:21-32 and metrics_exporter.go:22-34, which the reference computes
 * separately): counts_a / counts_b as b200_hist_fixed would fill them for (edges_a, mode_a) and (edges_b, mode_b).
 * stats (may be NULL) holds the moments and histogram a's dropped counts; dropped_b (may be NULL) receives
 * histogram b's {below, above}.  B200_ERR_TOO_LARGE if the two edge sets have more than 66 distinct thresholds. */
int b200_hist_fixed_pair(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                         const double* edges_a, int nedges_a, int mode_a, uint64_t* counts_a,
                         const double* edges_b, int nedges_b, int mode_b, uint64_t* counts_b,
                         uint64_t* dropped_b, b200_lat_stats* stats);
/* Exact order statistics: out[i] = k[i]-th smallest (0-based) of trunc(sample/divisor), i.e. sorted[k[i]]
 * of ssd_test/main.go:158-163.  Samples are not modified (no sort: one sampled histogram pass for long
 * inputs, MSD radix select otherwise).  k and out are HOST arrays.  B200_ERR_EMPTY when n == 0
 * (the reference panics there). */
int b200_select_kth(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                    const uint64_t* k, int nk, int64_t* out);
/* b200_select_kth plus count / min / max / exact sums of trunc(sample/divisor) over ALL samples: everything
 * ssd_test/main.go:147-163 prints (the avg from the sum, sorted[k], min, max) from one call.  Long inputs get
 * the moments from the same pass over the samples that settles the ranks; short ones take one extra pass.
 * stats->dropped_* are 0. */
int b200_select_kth_stats(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                          const uint64_t* k, int nk, int64_t* out, b200_lat_stats* stats);
/* The same without the sum of squares (stats->sumsq_* = 0): exactly what ssd_test/main.go:147-163 prints — the
 * average needs the sum only — and the pass over long inputs is three instructions per sample shorter. */
int b200_select_kth_avg(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                        const uint64_t* k, int nk, int64_t* out, b200_lat_stats* stats);
/* Building blocks of the sampled select for the one-process-per-GPU form (custom-go-client-benchmark_b200/latency.py;
 * b200_select_kth runs the same steps itself on one device).
 * b200_select_sample: one sample per stratum of `stride` consecutive inputs, at a hashed position inside the
 *   stratum; *nsample = n / stride of them land in a library-owned device buffer, *d_sample.
 * b200_bracket_pass: counts[c] (HOST, nthr + 1 entries) = samples with exactly c of the strictly ascending raw
 *   thresholds <= sample; the samples of ODD c (the brackets, when the thresholds are lo0, hi0+1, lo1, hi1+1, ...)
 *   are kept, in no particular order, in a library-owned device buffer *d_kept; *nkept is their number, of which
 *   only the first `cap` are stored.  stats (may be NULL): moments of trunc(sample/divisor) over all n samples.
 * b200_bracket_pass_avg: the same without the sum of squares (stats->sumsq_* = 0), as b200_select_kth_avg.
 * Both buffers stay valid until the next b200_select_* / b200_bracket_pass call on the same device. */
int b200_select_sample(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, size_t stride,
                       const int64_t** d_sample, size_t* nsample);
/* Sampling and the order statistics of the sample in ONE call (one device round trip): out[i] bounds the ranks[i]-th
 * smallest (0-based) of the same stratified sample — from below, or from above where upper[i] != 0 (upper may be
 * NULL) — to within `slack` sample ranks (0 = the order statistic itself); bounds are all the brackets of a sampled
 * select need and cost a pass less.  nranks <= 16, ranks[i] < n / stride = *nsample. */
int b200_select_sample_bounds(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, size_t stride,
                              const uint64_t* ranks, const uint8_t* upper, int nranks, uint64_t slack, int64_t* out,
                              size_t* nsample);
int b200_bracket_pass(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                      const int64_t* thresholds, int nthr, uint64_t cap, uint64_t* counts, const int64_t** d_kept,
                      uint64_t* nkept, b200_lat_stats* stats);
int b200_bracket_pass_avg(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                          const int64_t* thresholds, int nthr, uint64_t cap, uint64_t* counts, const int64_t** d_kept,
                          uint64_t* nkept, b200_lat_stats* stats);
/* The same when the caller already knows bounds lo_ns <= every sample <= hi_ns (for instance from the min/max
 * of a previous b200_hist_fixed): saves the min/max pass of the radix select.  The bounds must be true: a
 * sample outside them is either reported as B200_ERR_INVALID (radix select) or ranked like any other
 * sample (the sampled pass for long inputs does not look at the bounds). */
int b200_select_kth_range(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, int64_t lo_ns,
                          int64_t hi_ns, const uint64_t* k, int nk, int64_t* out);
/* Building blocks of b200_select_kth for the one-process-per-GPU form, where the caller all-reduces the
 * results over ranks between calls (NCCL; see custom-go-client-benchmark_b200/latency.py):
 * b200_minmax: min and max raw sample on this device (INT64_MAX / INT64_MIN when n == 0).
 * b200_radix_count: with key' = (sample ^ 2^63) - key_base, for each i < nk count the digit
 *   (key' >> shift) & (2^(top-shift) - 1) over the samples with key' >> top == prefix[i] (every sample when
 *   top >= 64).  Requires 0 < top - shift <= 12 and nk <= B200_SELECT_MAX_K.  counts is a HOST array
 *   [nk][B200_RADIX_BINS]. */
int b200_minmax(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t* min, int64_t* max);
int b200_radix_count(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, uint64_t key_base, int shift,
                     int top, int nk, const uint64_t* prefix, uint64_t* counts);
/* b200_radix_count with a key base per histogram (key_bases[i] instead of one key_base; top < 64): the order
 * statistics of a multi-rank sampled select are each resolved inside their own bracket, relative to its lower end. */
int b200_radix_count_based(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, const uint64_t* key_bases,
                           int shift, int top, int nk, const uint64_t* prefix, uint64_t* counts);

/* One process driving several devices (the other multi-GPU form is one process per GPU with the counts
 * all-reduced over NCCL by the caller, latency.py): device devs[i] holds n[i] samples at d_samples[i]; the
 * per-device kernels run concurrently, bin / digit counts are summed on the host.  devs may repeat a device. */
int b200_hist_fixed_multi(b200_ctx* ctx, const int* devs, const int64_t* const* d_samples, const size_t* n, int ndev,
                          int64_t divisor, const double* edges, int nedges, int mode, uint64_t* counts,
                          b200_lat_stats* stats);
int b200_select_kth_multi(b200_ctx* ctx, const int* devs, const int64_t* const* d_samples, const size_t* n, int ndev,
                          int64_t divisor, const uint64_t* k, int nk, int64_t* out);

/* ---- latency text files --------------------------------------------------------------------------------- */

/* README.md:25-28 `for line in f: x.append(float(line))` over the file execute_pb.sh:4,8 writes (one Go
 * Duration.String() per line with 'm' and 's' replaced by spaces).  d_text: nbytes of text on the device.
 * d_values[i] = float(line i) for the first max_lines lines (device array); d_status[i] (device, may be NULL):
 * 0 parsed, 1 ValueError in Python (empty line, status line, "850µ " ...), 2 valid syntax but outside the
 * exactly-rounded domain (> 15 significant digits or |decimal exponent| > 22; never produced by the reference
 * pipeline).  *first_bad_line = index of the first line with status != 0 — where the reference script raises —
 * or UINT64_MAX.  Lines follow Python's universal newlines.  B200_ERR_TOO_LARGE if there are more than
 * max_lines lines (*nlines is still set). */
int b200_parse_latency_text(b200_ctx* ctx, int dev, const void* d_text, size_t nbytes, double* d_values,
                            uint8_t* d_status, size_t max_lines, uint64_t* nlines, uint64_t* first_bad_line,
                            uint64_t* nbad, uint64_t* ninexact);
/* README.md:30,32 on the parsed floats: numpy.histogram / OpenCensus buckets on float64 samples (same count
 * layout as b200_hist_fixed), NaNs counted apart, and the sum of the samples as an unevaluated double-double
 * (sum_hi + sum_lo) for mean = sum/len.  All outputs are HOST pointers; all but counts may be NULL. */
int b200_hist_fixed_f64(b200_ctx* ctx, int dev, const double* d_values, size_t n, const double* edges, int nedges,
                        int mode, uint64_t* counts, uint64_t* dropped_below, uint64_t* dropped_above,
                        uint64_t* nan_count, double* sum_hi, double* sum_lo);

#ifdef __cplusplus
}
#endif
#endif
