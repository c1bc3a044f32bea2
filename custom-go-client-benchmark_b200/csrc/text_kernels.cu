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
        // at most kTileBytes + kHalo + 16 bytes = 273 vectors for kTextThreads threads: one each and a few second
        // ones, written out with 32-bit indices (as a generic size_t loop this copy was 14 % of the parse kernel)
        const unsigned nvec = (unsigned)(len / 16);
        const uint4* src = reinterpret_cast<const uint4*>(t + sbase);
        uint4* dst = reinterpret_cast<uint4*>(sm);
        static_assert((kTileBytes + kHalo + 16) / 16 <= 2 * kTextThreads, "two vectors per thread cover the staging");
        const unsigned i0 = threadIdx.x, i1 = threadIdx.x + kTextThreads;
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
        if (i0 < nvec) v0 = __ldg(src + i0);
        if (i1 < nvec) v1 = __ldg(src + i1);
        if (i0 < nvec) dst[i0] = v0;
        if (i1 < nvec) dst[i1] = v1;
        const unsigned tail = nvec * 16 + threadIdx.x;  // the last, incomplete vector of the text (< 16 bytes)
        if (tail < (unsigned)len) sm[tail] = __ldg(t + sbase + tail);
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

// ---- line starts, sixteen bytes at a time (SWAR on four little-endian words) ----------------------------------
// 0x80 in every byte of w that equals c (exact, no carries between bytes)
__device__ __forceinline__ unsigned eq_flags(unsigned w, unsigned c4) {
    const unsigned x = w ^ c4;
    return ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);
}
// Bit k of the result: a line starts at byte k of the window w[0..3] (Python's universal newlines: a line starts
// after "\n", and after a "\r" that is not followed by "\n").  prev: the byte before the window (0x0A in front
// of the very first byte of the text).  Only the first `valid` bytes count.
__device__ __forceinline__ unsigned window_starts(const unsigned w[4], unsigned prev, int valid) {
    unsigned nl[4], cr[4], anycr = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        nl[j] = eq_flags(w[j], 0x0A0A0A0Au);
        cr[j] = eq_flags(w[j], 0x0D0D0D0Du);
        anycr |= cr[j];
    }
    const unsigned pnl = prev == '\n' ? 0x80000000u : 0u, pcr = prev == '\r' ? 0x80000000u : 0u;
    unsigned m = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        // flags of the PREVIOUS byte, moved onto this byte's position
        const unsigned nlp = __funnelshift_l(j ? nl[j - 1] : pnl, nl[j], 8);
        unsigned st = nlp;
        if (anycr | pcr) st |= __funnelshift_l(j ? cr[j - 1] : pcr, cr[j], 8) & ~nl[j];
        // gather the four flag bits (7, 15, 23, 31) into a nibble: the products land on distinct bits
        m |= ((((st >> 7) & 0x01010101u) * 0x00204081u >> 21) & 0xFu) << (4 * j);
    }
    return valid >= 16 ? m : (valid <= 0 ? 0u : m & ((1u << valid) - 1u));
}
// sixteen bytes of text at `base` as four words (zero beyond the end); g must be readable at [base, n)
__device__ __forceinline__ void load_window(const unsigned char* __restrict__ g, size_t n, size_t base, unsigned w[4]) {
    if (base + 16 <= n && ((reinterpret_cast<uintptr_t>(g) + base) & 15) == 0) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(g + base));
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
        return;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
        w[j] = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const size_t i = base + 4 * j + k;
            if (i < n) w[j] |= (unsigned)__ldg(g + i) << (8 * k);
        }
    }
}

// Pass 1: where the lines start.  One 16-bit mask per 16-byte window (kept for pass 3, which then need not look at
// the bytes again to find its lines) and the number of line starts per tile.
__global__ void __launch_bounds__(kTextThreads)
text_count_kernel(const unsigned char* __restrict__ tg, size_t n, unsigned long long* __restrict__ tile_counts,
                  unsigned short* __restrict__ masks) {
    __shared__ unsigned wsum[kTextThreads / 32];
    const size_t win = (size_t)blockIdx.x * kTextThreads + threadIdx.x, base = win * 16;
    unsigned m = 0;
    if (base < n) {
        unsigned w[4];
        load_window(tg, n, base, w);
        const unsigned prev = base ? __ldg(tg + base - 1) : '\n';
        m = window_starts(w, prev, (int)(n - base < 16 ? n - base : 16));
    }
    masks[win] = (unsigned short)m;
    unsigned c = __popc(m);
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

// inclusive scan over the 1024 threads of a CTA; *excl_base receives the exclusive prefix of the caller's warp
__device__ __forceinline__ unsigned long long block_scan_1024(unsigned long long v, unsigned long long* wtot,
                                                              unsigned long long* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const unsigned long long t = wtot[lane];
        unsigned long long in2 = t;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, in2, d);
            if (lane >= d) in2 += o;
        }
        wtot[lane] = in2 - t;
        if (lane == 31) *total = in2;
    }
    __syncthreads();
    return wtot[warp] + incl - v;  // exclusive
}

// Pass 2a: exclusive scan of the tile counts inside blocks of kScanBlock tiles (in place, coalesced); the block
// totals go to aux[].  Pass 2b: one CTA scans aux[] (exclusive, in place) and leaves the number of lines in
// *total.  The rank of a tile's first line is tile_counts[tile] + aux[tile / kScanBlock].
constexpr int kScanBlock = 1024;
__global__ void __launch_bounds__(kScanBlock)
text_scan_local_kernel(unsigned long long* __restrict__ tile_counts, size_t ntiles, unsigned long long* __restrict__ aux) {
    __shared__ unsigned long long wtot[32], tot;
    const size_t i = (size_t)blockIdx.x * kScanBlock + threadIdx.x;
    const unsigned long long v = i < ntiles ? tile_counts[i] : 0;
    const unsigned long long excl = block_scan_1024(v, wtot, &tot);
    if (i < ntiles) tile_counts[i] = excl;
    if (threadIdx.x == 0) aux[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(kScanBlock)
text_scan_aux_kernel(unsigned long long* __restrict__ aux, size_t nblocks, unsigned long long* __restrict__ total) {
    __shared__ unsigned long long wtot[32], tot;
    unsigned long long carry = 0;
    for (size_t c0 = 0; c0 < nblocks; c0 += kScanBlock) {  // 1024 blocks (a 4 GB text) per round
        const size_t i = c0 + threadIdx.x;
        const unsigned long long v = i < nblocks ? aux[i] : 0;
        const unsigned long long excl = block_scan_1024(v, wtot, &tot);
        if (i < nblocks) aux[i] = carry + excl;
        carry += tot;
        __syncthreads();  // tot and wtot are rewritten next round
    }
    if (threadIdx.x == 0) *total = carry;
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


// 0x80 in every byte of w that is NOT an ASCII digit (w must be ASCII: no byte >= 0x80)
__device__ __forceinline__ unsigned nondigit_flags(unsigned w) {
    const unsigned x = w ^ 0x30303030u;  // digits -> 0..9
    return (x + 0x76767676u) & 0x80808080u;  // x >= 10 (x <= 0x7F: no carry into the next byte)
}
// the number spelled by the four ASCII digits of w, first character in the low byte
__device__ __forceinline__ unsigned four_digits(unsigned w) {
    const unsigned d = w - 0x30303030u;
    const unsigned t = d * 10u + (d >> 8);  // byte 0: d0*10 + d1, byte 2: d2*10 + d3 (each <= 99: no carries)
    return (t & 0xFFu) * 100u + ((t >> 16) & 0xFFu);
}

// The common line, without a loop over its characters: [1-4 digits] [. 0-11 digits] [spaces] "\n", all inside
// the 16 bytes w[0..3] that start at the line start (what Go's Duration.String() leaves after `tr 'ms' ' '` for
// anything from 1 ms to 9 s).  The digits are taken four at a time (SWAR), the significand is the integer
// int * 10^12 + fraction padded to twelve digits (< 2^53), and the value is that integer divided by 10^12: one
// correctly rounded IEEE division of two exact doubles, hence the same double Python's float() gives.  tail(i):
// byte i of the line, for the few bytes after the digits.  Returns false when the line is anything else (the
// caller then takes the character-by-character parser).
template <typename Tail>
__device__ __forceinline__ bool parse_fast(const unsigned w[4], Tail tail, double* out) {
    if ((w[0] | w[1] | w[2] | w[3]) & 0x80808080u) return false;  // not ASCII
    const unsigned nd0 = nondigit_flags(w[0]);
    int ni;  // digits before the first other character
    if (nd0) ni = (__ffs(nd0) >> 3) - 1;
    else if (nondigit_flags(w[1]) & 0x80u) ni = 4;
    else return false;
    if (ni < 1) return false;
    const unsigned after = __funnelshift_rc(w[0], w[1], 8 * ni) & 0xFFu;  // the character after the integer part
    const unsigned iw = __funnelshift_rc(0x30303030u, w[0], 8 * ni);       // its last four digits, '0'-padded in front
    int end = ni;  // where the digits end
    unsigned f0 = 0x30303030u, f1 = 0x30303030u, f2 = 0x30303030u;
    if (after == '.') {
        const int s = ni + 1;  // 2..5
        const bool hiw = s >= 4;
        const unsigned a0 = hiw ? w[1] : w[0], a1 = hiw ? w[2] : w[1], a2 = hiw ? w[3] : w[2], a3 = hiw ? 0u : w[3];
        const int sh = 8 * (s & 3);
        f0 = __funnelshift_r(a0, a1, sh);
        f1 = __funnelshift_r(a1, a2, sh);
        f2 = __funnelshift_r(a2, a3, sh);
        const unsigned n0 = nondigit_flags(f0), n1 = nondigit_flags(f1), n2 = nondigit_flags(f2);
        int nf;
        if (n0) nf = (__ffs(n0) >> 3) - 1;
        else if (n1) nf = 4 + (__ffs(n1) >> 3) - 1;
        else if (n2) nf = 8 + (__ffs(n2) >> 3) - 1;
        else return false;  // twelve or more fraction digits
        // bytes kept of each word: the first nf, nf - 4, nf - 8 (clamped to 0..4) — a clamping funnel shift of ones
        const unsigned k0 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32 - 8 * min(nf, 4)),
                       k1 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32 - 8 * min(max(nf - 4, 0), 4)),
                       k2 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32 - 8 * min(max(nf - 8, 0), 4));
        f0 = (f0 & k0) | (0x30303030u & ~k0);
        f1 = (f1 & k1) | (0x30303030u & ~k1);
        f2 = (f2 & k2) | (0x30303030u & ~k2);
        end = s + nf;
    }
    // after the digits: spaces, then the line feed (anything else, a carriage return included: the other parser)
    for (int i = end;; i++) {
        if (i >= 16) return false;
        const unsigned c = tail(i);
        if (c == '\n') break;
        if (c != ' ') return false;
    }
    const unsigned ip = four_digits(iw);
    if (ip >= 9007u) return false;  // the significand below must stay under 2^53
    const unsigned hi8 = ip * 10000u + four_digits(f0);                // < 10^8
    const unsigned lo8 = four_digits(f1) * 10000u + four_digits(f2);  // < 10^8
    const unsigned long long mant = (unsigned long long)hi8 * 100000000ull + lo8;
    *out = (double)mant / 1e12;
    return true;
}

__global__ void __launch_bounds__(kTextThreads)
text_parse_kernel(const unsigned char* __restrict__ tg, size_t n, const unsigned long long* __restrict__ tile_base,
                  const unsigned long long* __restrict__ aux, const unsigned short* __restrict__ masks,
                  double* __restrict__ values, unsigned char* __restrict__ status, size_t max_lines,
                  unsigned long long* __restrict__ summary /* [0] first bad line, [1] bad count, [2] inexact count */) {
    __shared__ unsigned wsum[kTextThreads / 32];
    __shared__ unsigned short starts[kTileBytes];  // offsets (from the tile's first byte) of the lines that start in it
    __shared__ __align__(16) unsigned char stage[kTileBytes + kHalo + 16];
    const TextView t = stage_tile(tg, n, blockIdx.x, stage);
    const size_t tbase = (size_t)blockIdx.x * kTileBytes;
    const unsigned char* const tile = stage + (tbase - t.sbase);  // byte 0 of the tile in shared memory
    const unsigned m = masks[(size_t)blockIdx.x * kTextThreads + threadIdx.x];  // this window's line starts (pass 1)
    const unsigned c = __popc(m);
    unsigned incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if ((threadIdx.x & 31) >= d) incl += o;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    unsigned wbase = 0, total = 0;
    for (int w = 0; w < kTextThreads / 32; w++) {
        if (w < (int)(threadIdx.x >> 5)) wbase += wsum[w];
        total += wsum[w];
    }
    // the line starts of the tile, in order: lines are then dealt to the threads evenly (a thread parsing the
    // starts of its own window has none, one or two of them and the warp waits for the lane with the most)
    unsigned at = wbase + incl - c, mm = m;
    while (mm) {
        const int k = __ffs(mm) - 1;
        mm &= mm - 1;
        starts[at++] = (unsigned short)(threadIdx.x * 16 + k);
    }
    __syncthreads();
    const unsigned long long rank0 = tile_base[blockIdx.x] + aux[blockIdx.x / kScanBlock];
    // per-line quantities in 32 bits, relative to the tile: the lines this tile may write, the staged bytes past its start
    const unsigned mine = rank0 >= max_lines ? 0u : (unsigned)min((unsigned long long)total, (unsigned long long)max_lines - rank0);
    const unsigned staged_len = (unsigned)(t.send - tbase);
    double* const vout = values + rank0;
    unsigned char* const sout = status ? status + rank0 : nullptr;
    for (unsigned l = threadIdx.x; l < mine; l += kTextThreads) {
        const unsigned off = starts[l];
        Parsed p{0.0, 1};
        bool done = false;
        if (off + 20 <= staged_len) {  // sixteen bytes from the line start, read as aligned words and shifted into place
            const unsigned* aw = reinterpret_cast<const unsigned*>(tile + (off & ~3u));
            const int sh = 8 * (off & 3);
            unsigned w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) w[j] = __funnelshift_r(aw[j], aw[j + 1], sh);
            done = parse_fast(w, [&](int i) -> unsigned { return tile[off + i]; }, &p.v);
            if (done) p.status = 0;
        }
        if (!done) p = parse_float(t, n, tbase + off);
        vout[l] = p.v;
        if (sout) sout[l] = p.status;
        if (p.status) {
            atomicMin(summary, rank0 + l);
            atomicAdd(summary + (p.status == 1 ? 1 : 2), 1ull);
        }
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

// scratch: [0, ntiles) tile counts / ranks, [ntiles] the number of lines, then the scan's block totals, then the
// windows' line-start masks (16 bits per 16 bytes of text)
size_t text_tiles(size_t nbytes) { return (nbytes + kTileBytes - 1) / kTileBytes; }
size_t text_scratch_words(size_t nbytes) {
    const size_t ntiles = text_tiles(nbytes), nblocks = (ntiles + kScanBlock - 1) / kScanBlock;
    return ntiles + 1 + nblocks + (ntiles * kTextThreads * sizeof(unsigned short) + 7) / 8 + 1;
}

cudaError_t launch_text_parse(const unsigned char* d_text, size_t n, double* d_values, unsigned char* d_status,
                              size_t max_lines, unsigned long long* d_scratch /* text_scratch_words(n) */,
                              unsigned long long* d_summary, cudaStream_t stream) {
    const size_t ntiles = text_tiles(n), nblocks = (ntiles + kScanBlock - 1) / kScanBlock;
    unsigned long long init[3] = {~0ull, 0, 0};
    cudaError_t e = cudaMemcpyAsync(d_summary, init, sizeof init, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
    if (!ntiles) return cudaMemsetAsync(d_scratch, 0, sizeof(unsigned long long), stream);
    unsigned long long* const aux = d_scratch + ntiles + 1;
    unsigned short* const masks = reinterpret_cast<unsigned short*>(aux + nblocks);
    text_count_kernel<<<(unsigned)ntiles, kTextThreads, 0, stream>>>(d_text, n, d_scratch, masks);
    text_scan_local_kernel<<<(unsigned)nblocks, kScanBlock, 0, stream>>>(d_scratch, ntiles, aux);
    text_scan_aux_kernel<<<1, kScanBlock, 0, stream>>>(aux, nblocks, d_scratch + ntiles);
    text_parse_kernel<<<(unsigned)ntiles, kTextThreads, 0, stream>>>(d_text, n, d_scratch, aux, masks, d_values, d_status,
                                                                      max_lines, d_summary);
    return cudaGetLastError();
}

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
