// Internal launch interface of the latency-reduction kernels (histogram, moments, radix select).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

constexpr int kMaxThresholds = 66;      // B200_HIST_MAX_EDGES + slack
constexpr int kLutSize = 8192;          // cells of the histogram kernel's look-up table
constexpr int kRadixBits = 12;
constexpr int kRadixBins = 1 << kRadixBits;
constexpr int kMaxSelect = 4;           // order statistics resolved per pass

// Exact integer moments of the scaled samples v = trunc(ns / divisor).
struct LatMoments {
    unsigned long long count;
    long long min_ns, max_ns;           // of the raw samples (division is monotone)
    unsigned long long sum_lo;          // 128-bit two's complement sum of v
    long long sum_hi;
    unsigned long long sumsq_lo, sumsq_hi;  // 128-bit sum of v*v
};

// Device-side parameter block for the histogram kernel, built on the host.
struct HistParams {
    long long thr[kMaxThresholds];      // thresholds in raw ns, ascending: count(n) = #{j : thr[j] <= n}
    int nthr;
    unsigned long long magic;           // division by `divisor` as multiply-high, 64-bit dividends (generic path)
    int magic_shift;
    int magic_add;
    unsigned int magic32;               // the same for 32-bit dividends.  mode32: 0 = divisor 1, 2 = q is 0,
    int shift32;                        //   1: q = (t + ((n - t) >> 1)) >> shift32, t = umulhi(magic32, n)
    int pre32;                          //   3: q = umulhi(n >> pre32, magic32) >> shift32, and q < 2^28,
    int mode32;                         //      exact for n < 2^reach_bits
    int reach_bits;                     // the kernel's fast path takes 0 <= n < 2^reach_bits (32; more in mode 3)
    long long divisor;
    // Look-up table for samples 0 <= n < 2^reach_bits, indexed by min(quotient, last cell) (by_q: needs mode32 == 3;
    // chosen when it parks nothing, i.e. the thresholds in reach are multiples of the divisor and inside the table)
    // or by min(n >> lut_shift, last cell), lut_shift >= pre32 in mode 3.  lut[c] = 1 + #{thresholds <= every sample of cell c}, the counter
    // row of the cell's samples; 0 when the cell holds a threshold or (collecting) belongs to a kept bin: parked.
    int by_q;
    int lut_shift;
    int lut_bytes;                      // table cells actually in use (a multiple of 128, <= kLutSize)
    unsigned char lut[kLutSize];
};

// Optional output of the histogram pass: the samples that fall into an ODD bin (bins are numbered from 0; the
// sampled select puts its brackets there) are appended to buf in no particular order.  *count ends up as the
// number of such samples even when it exceeds cap (the excess is not stored).
struct HistCollect {
    long long* buf;
    unsigned long long* count;          // device counter, zeroed by the launch
    unsigned long long cap;
};

// Device-resident radix select: all passes are queued without a host round trip (the host-driven select pays
// one launch + copy + synchronise per pass, which dominates on inputs of a few million samples).
constexpr int kSelMax = 8;              // order statistics resolved together
constexpr int kSelRounds = 6;           // ceil(64 / kRadixBits) passes cover any int64 range
struct SelState {
    unsigned long long key_base;        // min key; key = ns ^ 2^63
    unsigned long long prefix[kSelMax], rank[kSelMax];
    long long result[kSelMax];
    int top;                            // bits of key - key_base still unresolved
    int nk, nhist;                      // ranks; histograms the next pass fills (1 on the first pass)
    int bad;                            // a rank fell outside the counted samples
    int more[kSelMax];                  // one-kernel select: round r left a bucket above approx_max
    unsigned long long stamp[24];       // one-kernel select: %globaltimer of CTA 0 at its phase ends (B200_SELECT_TRACE)
    int nstamp;
};
struct SelRanks {
    unsigned long long rank[kSelMax];
    int nk;
};

// What a histogram pass leaves behind, in ONE device block so that a call resets it with one device-to-device
// copy (from a template held beside it) and reads it back with one copy into pinned memory: a pageable copy per
// field and direction was ~10 us each, a sixth of a seven-line call.
struct LatResults {
    unsigned long long counts[kMaxThresholds + 1];
    LatMoments mom;
    unsigned long long compact_n;       // samples kept / compacted / outside the fast path (one counter, reused)
    SelState sel;                       // the resident select's state (its answer)
};

// Input of a one-kernel select that an earlier kernel on the stream prepares (bracket_ranks_kernel after a bracket
// pass): the host launches both without knowing the counts and checks them afterwards.
struct SelDeferred {
    SelRanks ranks;                     // ranks among the kept samples
    unsigned long long n;               // kept samples (<= the buffer's capacity)
    unsigned long long miss;            // a rank fell outside the brackets, or they hold too much / nothing: no select
};

struct LatScratch {
    HistParams* d_params;
    SelDeferred* d_sel_in;
    unsigned* d_sel_bar;                // [2] the one-kernel select's grid barrier (count, generation)
    LatResults* d_results;              // the block; the three pointers below point into it
    LatResults* d_results_init;         // the template: zero counts, min = INT64_MAX, max = INT64_MIN
    LatResults* h_results;              // pinned: lat_read_results() copies the block here
    HistParams* h_params;               // pinned staging of the parameter block
    unsigned long long* d_counts;       // [kMaxThresholds + 1]
    LatMoments* d_mom;
    unsigned long long* d_radix;        // [kMaxSelect][kRadixBins]
    unsigned long long* h_radix;        // pinned: the same, read back (a pageable copy of 128 KB was ~40 us)
    unsigned long long* d_prefix;       // [2][kMaxSelect]: prefixes, then (optionally) a key base per histogram
    long long* d_compact;               // samples of the buckets still in play after the first radix pass (lazy)
    size_t compact_cap;                 // in samples
    unsigned long long* d_compact_n;    // device counter
    SelState* d_sel;                    // device-resident select: &d_results->sel
    unsigned long long* d_sel_hist;     // [kSelMax][kRadixBins]
    bool* sel_hist_clean;               // host: the last one-kernel select ended well, so d_sel_hist is all zero
    long long* d_sample;                // stratified sample of the input (sampled select, lazy)
    size_t sample_cap;
    long long* d_collect;               // samples inside the brackets (sampled select, lazy)
    size_t collect_cap;
};

// Where dynamic shared memory starts for a kernel without static shared memory.  The histogram and CRC row kernels
// lay their tables out at ABSOLUTE shared addresses and need that to be at most 1 KB (the driver's reserved block);
// cudaErrorInvalidConfiguration if it is not.  d_tmp: 8 bytes of device memory.  Asked once per device at set-up.
cudaError_t smem_base_probe(unsigned long long* d_tmp);
cudaError_t lat_scratch_init(LatScratch* s);
void lat_scratch_free(LatScratch* s);
// counts = 0, moments = start values, compact_n = 0 (one device-to-device copy); the whole block -> s.h_results
cudaError_t lat_reset_results(const LatScratch& s, cudaStream_t stream);
cudaError_t lat_read_results(const LatScratch& s, cudaStream_t stream);

// counts[c] for c in 0..nthr, moments optional (want_moments).  Results stay on the device in scratch.
// (uploads P to s.d_params first; P must have been built with collect == (collect != nullptr))
// want_moments: 0 none, 1 all, 2 without the sum of squares (collecting passes only; elsewhere the same as 1)
cudaError_t launch_hist(const long long* d_samples, size_t n, const LatScratch& s, int want_moments,
                        cudaStream_t stream, const HistParams& P, const HistCollect* collect = nullptr);
// out[i] = one sample of stratum i: x[i * stride + (hash(i) mod stride)], i < s, s * stride <= n
// *d_outside (device, may be NULL) = how many of those samples are negative or >= 2^32
// d_bounds: when given, its min_ns / max_ns receive the extremes of the sample (the bounds of a select over it that
// follows on the stream).  results: the scratch whose block holds d_outside and d_bounds; it is reset first (without
// it d_outside is cleared and d_bounds taken as it is).
cudaError_t launch_sample(const long long* d_samples, size_t stride, long long* d_out, size_t s, cudaStream_t stream,
                          unsigned long long* d_outside = nullptr, LatMoments* d_bounds = nullptr,
                          const LatScratch* results = nullptr);
// All the passes of a radix select of up to kSelMax ranks over [lo, hi] (pass have_bounds = false to take the
// bounds from a min/max pass first), queued on the stream; the answer lands in s.d_sel (result[], bad).
// The same select in one cooperative kernel (select_fused_grid() != 0 says the device can).  approx_max > 0: stop
// refining a rank once its bucket holds at most that many samples and give the bucket's lowest key (highest for the
// ranks in ceil_mask): bounds on the order statistics, for the sampled select's brackets.
int select_fused_grid();
// have_bounds: 0 the kernel finds min and max itself, 1 [lo, hi] are given, 2 they are in s.d_mom already (left by
// launch_sample with `bounds` on the same stream)
// in: ranks and n come from s.d_sel_in instead (n is then only an upper bound, for the grid)
cudaError_t launch_select_fused(const long long* d_samples, size_t n, const SelRanks& ranks, int have_bounds,
                                long long lo, long long hi, unsigned long long approx_max, unsigned ceil_mask,
                                const LatScratch& s, cudaStream_t stream, bool deferred = false);
// After a bracket pass (thresholds lo0, hi0+1, lo1, hi1+1, ...: odd bins are the brackets), on the device: the ranks
// of the wanted order statistics k[0..nk) among the kept samples -> s.d_sel_in, for the select that follows.
cudaError_t launch_bracket_ranks(const LatScratch& s, int nbrackets, const unsigned long long* k, int nk,
                                 unsigned long long cap, cudaStream_t stream);
cudaError_t launch_select_resident(const long long* d_samples, size_t n, const SelRanks& ranks, bool have_bounds,
                                   long long lo, long long hi, const LatScratch& s, cudaStream_t stream);
// One MSD radix pass over key' = (ns ^ 2^63) - key_base: for each k < nk count the digit
// (key' >> shift) & (2^(top-shift) - 1) over the samples with key' >> top == d_prefix[k] (all samples when
// top >= 64).  top - shift <= kRadixBits.
cudaError_t launch_radix_pass(const long long* d_samples, size_t n, unsigned long long key_base, int shift, int top,
                              int nk, const LatScratch& s, cudaStream_t stream,
                              bool own_bases = false);  // own_bases: histogram k's keys are taken from s.d_prefix[kMaxSelect + k]
cudaError_t launch_minmax(const long long* d_samples, size_t n, const LatScratch& s, cudaStream_t stream);
// Append to s.d_compact every sample whose key' >> top equals one of the nprefix (distinct) values in
// s.d_prefix; the count lands in *s.d_compact_n.  At most s.compact_cap samples are written.
cudaError_t launch_compact(const long long* d_samples, size_t n, unsigned long long key_base, int top, int nprefix,
                           const LatScratch& s, cudaStream_t stream);

// Latency text codec + float64 histogram (text_kernels.cu)
size_t text_tiles(size_t nbytes);          // the number of lines ends up in d_scratch[text_tiles(n)]
size_t text_scratch_words(size_t nbytes);  // 64-bit words of scratch the parse needs
cudaError_t launch_text_parse(const unsigned char* d_text, size_t n, double* d_values, unsigned char* d_status,
                              size_t max_lines, unsigned long long* d_scratch /* text_scratch_words(n) */,
                              unsigned long long* d_summary /* [3] */, cudaStream_t stream);
int hist_f64_grid();
cudaError_t launch_hist_f64(const double* d_x, size_t n, const double* edges, int nedges, int mode,
                            unsigned long long* d_counts /* kMaxThresholds+1 */,
                            double* d_sums /* 4 + 2*256*hist_f64_grid() */, cudaStream_t stream);

// Host helper: fill thresholds / LUT / magic for (edges, mode, divisor).
int host_build_hist_params(const double* edges, int nedges, int mode, long long divisor, HistParams* out);
// Just the thresholds of (edges, mode, divisor): ascending (not necessarily strictly) raw-ns values such that the
// bin of a sample is the number of thresholds <= it.  Returns their number (<= kMaxThresholds) or -1.
int host_edges_to_thresholds(const double* edges, int nedges, int mode, long long divisor, long long* thr);
// The same from raw thresholds (strictly ascending int64 ns): bin(n) = #{j : thr[j] <= n}; the divisor only
// scales the moments.
int host_build_hist_params_thr(const long long* thr, int nthr, long long divisor, bool collect, HistParams* out);

}  // namespace b200
