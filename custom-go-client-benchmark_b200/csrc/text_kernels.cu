// Latency text codec and float64 histogram for sm_100a.
//
// Reference: execute_pb.sh:4,8 pipes the benchmark's stdout through `tr 'ms' ' '` into grpc_N.txt / http_N.txt
// (one Go Duration.String() per line with every 'm' and 's' turned into a space), and README.md:25-28 reads
// that file back with
//        for line in f: x.append(float(line))
// then README.md:30,32 takes sum(x)/len(x) and plt.hist(x, bins).  parse kernels = that loop, bit-exact with
// Python's float() on the supported domain; hist_f64 = numpy.histogram / the OpenCensus bucket rule on doubles.
//
// float(line) semantics restated: surrounding ASCII whitespace is stripped; [sign] digits [. digits]
// [e [sign] digits] with single underscores between digits, or inf / infinity / nan in any case; anything else
// is a ValueError (status 1) — which is what the reference hits on its own status lines and on "850µ ".
// Exactness: decimal -> binary64 is correctly rounded when the significand has <= 15 significant digits and
// the decimal exponent is within +-22 (one exact IEEE multiply or divide).  Go's Duration.String() never
// leaves that domain; inputs beyond it get status 2 ("outside the exact domain") instead of a rounded guess.
// Line model: Python universal newlines — "\n", "\r\n" and a lone "\r" end a line; a final line without a
// terminator counts; non-ASCII bytes are invalid (Python would also strip U+0085 / U+00A0; not reproduced).
#include "lat_kernels.h"

namespace b200 {

namespace {

constexpr int kTextThreads = 256;
constexpr int kTileBytes = kTextThreads * 16;

__constant__ double c_pow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

constexpr int kHalo = 256;  // bytes staged beyond the tile: lines are short, longer ones fall back to global reads

// The bytes of one tile (plus one byte before and kHalo after) staged in shared memory; anything else is read
// from global memory.
struct TextView {
    const unsigned char* g;
    const unsigned char* sm;
    size_t sbase, send;  // [sbase, send) is staged
    __device__ __forceinline__ unsigned char operator[](size_t i) const {
        return (i >= sbase && i < send) ? sm[i - sbase] : __ldg(g + i);
    }
};

// cooperative load of tile `tile` into sm (kTileBytes + kHalo + 16 bytes)
__device__ __forceinline__ TextView stage_tile(const unsigned char* __restrict__ t, size_t n, size_t tile,
                                               unsigned char* sm) {
    const size_t tbase = tile * (size_t)kTileBytes;
    const size_t sbase = tbase ? tbase - 16 : 0;                       // keeps 16-byte alignment of the copies
    size_t send = tbase + kTileBytes + kHalo;
    if (send > n) send = n;
    const size_t len = send > sbase ? send - sbase : 0;
    if ((reinterpret_cast<uintptr_t>(t) & 15) == 0) {
        const size_t nvec = len / 16;
        const uint4* src = reinterpret_cast<const uint4*>(t + sbase);
        uint4* dst = reinterpret_cast<uint4*>(sm);
        for (size_t i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = __ldg(src + i);
        for (size_t i = nvec * 16 + threadIdx.x; i < len; i += blockDim.x) sm[i] = __ldg(t + sbase + i);
    } else {
        for (size_t i = threadIdx.x; i < len; i += blockDim.x) sm[i] = __ldg(t + sbase + i);
    }
    __syncthreads();
    return TextView{t, sm, sbase, send};
}

__device__ __forceinline__ bool is_space(unsigned char c) {
    return c == ' ' || (c >= 9 && c <= 13) || (c >= 0x1c && c <= 0x1f);
}
template <typename Text>
__device__ __forceinline__ bool line_ends_at(const Text& t, size_t n, size_t i) {
    return t[i] == '\n' || (t[i] == '\r' && (i + 1 >= n || t[i + 1] != '\n'));
}

// bit k of the result: a line starts at byte base + k (k < 16)
template <typename Text>
__device__ __forceinline__ unsigned start_mask(const Text& t, size_t n, size_t base) {
    unsigned m = 0;
    if (base >= n) return 0;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        size_t i = base + k;
        if (i < n && (i == 0 || line_ends_at(t, n, i - 1))) m |= 1u << k;
    }
    return m;
}

__global__ void __launch_bounds__(kTextThreads)
text_count_kernel(const unsigned char* __restrict__ tg, size_t n, unsigned long long* __restrict__ tile_counts) {
    __shared__ unsigned wsum[kTextThreads / 32];
    __shared__ __align__(16) unsigned char stage[kTileBytes + kHalo + 16];
    const TextView t = stage_tile(tg, n, blockIdx.x, stage);
    const size_t base = ((size_t)blockIdx.x * kTextThreads + threadIdx.x) * 16;
    unsigned c = __popc(start_mask(t, n, base));
#pragma unroll
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned s = 0;
        for (int w = 0; w < kTextThreads / 32; w++) s += wsum[w];
        tile_counts[blockIdx.x] = s;
    }
}

// exclusive scan of tile counts in place (one CTA); total goes to tile_counts[ntiles]
__global__ void __launch_bounds__(1024) text_scan_kernel(unsigned long long* __restrict__ tile_counts, size_t ntiles) {
    __shared__ unsigned long long part[1024];
    const size_t per = (ntiles + 1023) / 1024;
    const size_t lo = (size_t)threadIdx.x * per, hi = lo + per < ntiles ? lo + per : ntiles;
    unsigned long long s = 0;
    for (size_t i = lo; i < hi; i++) s += tile_counts[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; i++) {
            unsigned long long v = part[i];
            part[i] = run;
            run += v;
        }
        tile_counts[ntiles] = run;
    }
    __syncthreads();
    unsigned long long run = part[threadIdx.x];
    for (size_t i = lo; i < hi; i++) {
        unsigned long long v = tile_counts[i];
        tile_counts[i] = run;
        run += v;
    }
}

struct Parsed {
    double v;
    unsigned char status;  // 0 ok, 1 ValueError, 2 outside the exact domain
};

__device__ __forceinline__ unsigned char lower(unsigned char c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }

template <typename Text>
__device__ bool match_word(const Text& t, size_t s, size_t e, const char* w) {
    size_t i = s;
    for (; *w; w++, i++)
        if (i >= e || lower(t[i]) != (unsigned char)*w) return false;
    return i == e;
}

template <typename Text>
__device__ Parsed parse_float(const Text& t, size_t n, size_t s) {
    Parsed r{0.0, 1};
    size_t e = s;
    while (e < n && !line_ends_at(t, n, e)) e++;  // e = terminator (or n)
    while (s < e && is_space(t[s])) s++;
    while (e > s && is_space(t[e - 1])) e--;
    if (s == e) return r;
    bool neg = false;
    if (t[s] == '+' || t[s] == '-') { neg = t[s] == '-'; s++; }
    if (s == e) return r;
    if (match_word(t, s, e, "inf") || match_word(t, s, e, "infinity")) {
        r.v = neg ? -__longlong_as_double(0x7FF0000000000000ll) : __longlong_as_double(0x7FF0000000000000ll);
        r.status = 0;
        return r;
    }
    if (match_word(t, s, e, "nan")) {
        r.v = __longlong_as_double(neg ? 0xFFF8000000000000ll : 0x7FF8000000000000ll);
        r.status = 0;
        return r;
    }
    unsigned long long mant = 0;
    int sig = 0, exp10 = 0, pend_int = 0, pend_frac = 0, ndig = 0;
    bool dot = false, prev_digit = false, prev_us = false, too_long = false;
    size_t i = s;
    for (; i < e; i++) {
        unsigned char c = t[i];
        if (c >= '0' && c <= '9') {
            ndig++;
            if (c == '0') {
                if (mant != 0) { if (dot) pend_frac++; else pend_int++; }
                else if (dot) exp10--;  // leading fractional zeros of a still-zero significand
            } else {
                int shift = pend_int + pend_frac;
                if (sig + shift + 1 > 19) too_long = true;
                else {
                    for (int k = 0; k < shift; k++) mant *= 10;
                    mant = mant * 10 + (c - '0');
                    sig += shift + 1;
                }
                exp10 -= pend_frac + (dot ? 1 : 0);
                pend_int = pend_frac = 0;
            }
            prev_digit = true;
            prev_us = false;
        } else if (c == '_') {
            if (!prev_digit || i + 1 >= e || t[i + 1] < '0' || t[i + 1] > '9') return r;
            prev_digit = false;
            prev_us = true;
        } else if (c == '.') {
            if (dot || prev_us) return r;
            dot = true;
            prev_digit = false;
        } else {
            break;
        }
    }
    if (ndig == 0) return r;
    exp10 += pend_int;
    if (i < e) {
        if (t[i] != 'e' && t[i] != 'E') return r;
        i++;
        bool eneg = false;
        if (i < e && (t[i] == '+' || t[i] == '-')) { eneg = t[i] == '-'; i++; }
        int ex = 0, edig = 0;
        bool pd = false;
        for (; i < e; i++) {
            unsigned char c = t[i];
            if (c >= '0' && c <= '9') {
                if (ex < 100000) ex = ex * 10 + (c - '0');
                edig++;
                pd = true;
            } else if (c == '_') {
                if (!pd || i + 1 >= e || t[i + 1] < '0' || t[i + 1] > '9') return r;
                pd = false;
            } else {
                return r;
            }
        }
        if (edig == 0) return r;
        exp10 += eneg ? -ex : ex;
    }
    if (mant == 0 && !too_long) {
        r.v = neg ? -0.0 : 0.0;
        r.status = 0;
        return r;
    }
    if (too_long || sig > 15 || exp10 < -22 || exp10 > 22) {
        // a 15-digit significand may still absorb a larger positive exponent exactly (Clinger's extension)
        if (!too_long && sig <= 15 && exp10 > 22 && exp10 <= 22 + (15 - sig)) {
            double m = (double)mant * c_pow10[exp10 - 22];
            r.v = m * c_pow10[22];
            r.v = neg ? -r.v : r.v;
            r.status = 0;
            return r;
        }
        r.status = 2;
        return r;
    }
    double m = (double)mant;  // exact: mant < 10^15 < 2^53
    r.v = exp10 < 0 ? m / c_pow10[-exp10] : m * c_pow10[exp10];
    r.v = neg ? -r.v : r.v;
    r.status = 0;
    return r;
}

__global__ void __launch_bounds__(kTextThreads)
text_parse_kernel(const unsigned char* __restrict__ tg, size_t n, const unsigned long long* __restrict__ tile_base,
                  double* __restrict__ values, unsigned char* __restrict__ status, size_t max_lines,
                  unsigned long long* __restrict__ summary /* [0] first bad line, [1] bad count, [2] inexact count */) {
    __shared__ unsigned wsum[kTextThreads / 32];
    __shared__ __align__(16) unsigned char stage[kTileBytes + kHalo + 16];
    const TextView t = stage_tile(tg, n, blockIdx.x, stage);
    const size_t base = ((size_t)blockIdx.x * kTextThreads + threadIdx.x) * 16;
    const unsigned m = start_mask(t, n, base);
    const unsigned c = __popc(m);
    unsigned incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if ((threadIdx.x & 31) >= d) incl += o;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    unsigned wbase = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) wbase += wsum[w];
    unsigned long long rank = tile_base[blockIdx.x] + wbase + incl - c;
    unsigned mm = m;
    while (mm) {
        const int k = __ffs(mm) - 1;
        mm &= mm - 1;
        if (rank < max_lines) {
            Parsed p = parse_float(t, n, base + k);
            values[rank] = p.v;
            if (status) status[rank] = p.status;
            if (p.status) {
                atomicMin(summary, rank);
                atomicAdd(summary + (p.status == 1 ? 1 : 2), 1ull);
            }
        }
        rank++;
    }
}

// ---- float64 histogram -------------------------------------------------------------------------------------

struct F64Params {
    double edges[kMaxThresholds];
    int nedges;
    int mode;  // 0 numpy, 1 opencensus
};

__global__ void __launch_bounds__(256)
hist_f64_kernel(const double* __restrict__ x, size_t n, F64Params P, unsigned long long* __restrict__ counts,
                double* __restrict__ sums /* [0] sum hi, [1] sum lo (double-double), [2] nan count */) {
    extern __shared__ unsigned tbl[];  // [warp][bin][lane]
    __shared__ double ed[kMaxThresholds];
    const int nbins = P.nedges + 1;  // count c = #{edges <= v} in 0..nedges (numpy: last edge inclusive)
    for (int i = threadIdx.x; i < P.nedges; i += 256) ed[i] = P.edges[i];
    for (int i = threadIdx.x; i < 8 * nbins * 32; i += 256) tbl[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned* mine = tbl + (size_t)warp * nbins * 32 + lane;
    double hi = 0, lo = 0;
    unsigned long long nans = 0;
    const size_t stride = (size_t)gridDim.x * 256;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        const double v = x[i];
        if (v != v) { nans++; continue; }
        int l = 0, h = P.nedges;  // c = number of edges <= v  (upper bound)
        while (l < h) {
            int mid = (l + h) >> 1;
            if (ed[mid] <= v) l = mid + 1; else h = mid;
        }
        int c = l;
        if (P.mode == 0 && c == P.nedges && v == ed[P.nedges - 1]) c = P.nedges - 1;  // right edge is closed
        mine[c * 32]++;
        // two-sum accumulation: hi + lo tracks the exact running sum to ~106 bits
        double s = hi + v, bb = s - hi;
        lo += (hi - (s - bb)) + (v - bb);
        hi = s;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nbins; b += 256) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; w++)
            for (int l2 = 0; l2 < 32; l2++) t += tbl[((size_t)w * nbins + b) * 32 + l2];
        if (t) atomicAdd(counts + b, t);
    }
    // per-thread (hi, lo) pairs are summed on the host in a fixed order (exactness over speed): write them out
    double* out = sums + 4 + 2 * ((size_t)blockIdx.x * 256 + threadIdx.x);
    out[0] = hi;
    out[1] = lo;
    if (nans) atomicAdd(reinterpret_cast<unsigned long long*>(sums + 2), nans);
}

}  // namespace

cudaError_t launch_text_parse(const unsigned char* d_text, size_t n, double* d_values, unsigned char* d_status,
                              size_t max_lines, unsigned long long* d_tiles /* ntiles+1 */, unsigned long long* d_summary,
                              cudaStream_t stream) {
    const size_t ntiles = (n + kTileBytes - 1) / kTileBytes;
    unsigned long long init[3] = {~0ull, 0, 0};
    cudaError_t e = cudaMemcpyAsync(d_summary, init, sizeof init, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
    if (!ntiles) return cudaMemsetAsync(d_tiles, 0, sizeof(unsigned long long), stream);
    text_count_kernel<<<(unsigned)ntiles, kTextThreads, 0, stream>>>(d_text, n, d_tiles);
    text_scan_kernel<<<1, 1024, 0, stream>>>(d_tiles, ntiles);
    text_parse_kernel<<<(unsigned)ntiles, kTextThreads, 0, stream>>>(d_text, n, d_tiles, d_values, d_status, max_lines,
                                                                      d_summary);
    return cudaGetLastError();
}

size_t text_tiles(size_t nbytes) { return (nbytes + kTileBytes - 1) / kTileBytes; }

int hist_f64_grid() {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms * 2;
}

cudaError_t launch_hist_f64(const double* d_x, size_t n, const double* edges, int nedges, int mode,
                            unsigned long long* d_counts, double* d_sums, cudaStream_t stream) {
    F64Params P{};
    for (int i = 0; i < nedges; i++) P.edges[i] = edges[i];
    P.nedges = nedges;
    P.mode = mode;
    const int grid = hist_f64_grid();
    cudaError_t e = cudaMemsetAsync(d_counts, 0, (kMaxThresholds + 1) * sizeof(unsigned long long), stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_sums, 0, (4 + 2 * (size_t)grid * 256) * sizeof(double), stream);
    if (e != cudaSuccess) return e;
    const size_t smem = (size_t)8 * (nedges + 1) * 32 * sizeof(unsigned);
    e = cudaFuncSetAttribute(hist_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)((size_t)8 * (kMaxThresholds + 1) * 32 * sizeof(unsigned)));
    if (e != cudaSuccess) return e;
    hist_f64_kernel<<<grid, 256, smem, stream>>>(d_x, n, P, d_counts, d_sums);
    return cudaGetLastError();
}

}  // namespace b200
