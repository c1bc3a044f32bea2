// Latency-sample reduction kernels for sm_100a: fixed-edge histogram + exact integer moments, and
// MSD radix counting for exact order statistics.
//
// Reference semantics (bit-exact targets):
//   sample      main.go:145-146  float64(time.Since(start).Milliseconds()): int64 ns / 1e6 truncating
//   numpy bins  README.md:21-23,32  numpy.histogram(x, [20,25,..,95]): [a,b) bins, last closed, rest dropped
//   OC buckets  metrics_exporter.go:28  first bound strictly greater than v, else overflow bucket
//   percentiles benchmark-script/ssd_test/main.go:147-163  sorted[size/2] etc.
// Both bin rules are monotone step functions of the raw nanosecond value, so the host turns the float
// edges into ascending int64 nanosecond thresholds and the kernel only counts thresholds <= n; no
// per-sample division or float compare is needed for binning.  One byte look-up in an 8 KB table (indexed by
// the sample's quotient or by its high bits) gives the counter row; the few samples whose cell holds a threshold
// are parked and resolved by a search over the thresholds (hist_kernel).
// All kernels stream 8 bytes per sample — the histogram family through per-warp rings of 2 KB tiles filled by
// bulk copies — and keep their counters private per (warp, lane) in shared memory, so the skewed latency
// distribution causes no atomic contention.  Order statistics: a stratified sample brackets the wanted ranks, one
// collecting histogram pass counts below the brackets and keeps what is inside, and a one-kernel radix select
// (sel_fused_kernel) settles the ranks among the kept samples; api.cu holds the host side of that.
#include "lat_kernels.h"

#ifdef B200_DEBUG_ASSERT
#include <cassert>
#endif

#include <algorithm>
#include <new>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

namespace {

#ifndef B200_HIST_THREADS
#define B200_HIST_THREADS 256
#endif
constexpr int kHistThreads = B200_HIST_THREADS;
constexpr int kHistWarps = kHistThreads / 32;
#ifndef B200_HIST_CTAS
#define B200_HIST_CTAS 3
#endif
constexpr int kHistCtasPerSm = B200_HIST_CTAS;  // 3 x 256 threads: 85 registers each, room for the prefetched tile
constexpr int kRadixThreads = 512;

struct U128 {
    unsigned long long lo, hi;
};
__device__ __forceinline__ void add128(U128& a, unsigned long long lo, unsigned long long hi) {
    unsigned long long s = a.lo + lo;
    a.hi += hi + (s < a.lo);
    a.lo = s;
}


// Per-thread running moments.  Fast path values fit 32 bits, so their squares fit 64.
struct SlowMoments {                                       // of generic-path samples: exact 128-bit
    long long mn = 0x7FFFFFFFFFFFFFFFll, mx = (long long)0x8000000000000000ull;
    U128 sum{0, 0}, sq{0, 0};
};
struct Moments {
    unsigned mn32 = 0xFFFFFFFFu, mx32 = 0;                 // of fast-path samples
    unsigned long long fsum = 0, fsq = 0;                  // fast path: q < 2^32, so 64 bits hold ~2^32 samples of q
    unsigned fsq_carry = 0;                                //   and q*q wraps at most once per add
};

// What the generic path needs, copied to shared memory at kernel start: it runs with one or two lanes of a warp
// active while the others wait, so its latency counts — a chain of dependent global loads (a microsecond each
// under load) per sample made the whole kernel 7x slower; from shared memory the same chain is ~300 cycles.
struct SlowParams {
    long long thr[kMaxThresholds];
    unsigned long long magic;
    int nthr, magic_shift, magic_add, pad;
};

__device__ __forceinline__ unsigned long long udiv_magic(unsigned long long a, const SlowParams& P) {
    if (P.magic == 0) return a >> P.magic_shift;  // power of two divisor
    unsigned long long t = __umul64hi(a, P.magic);
    if (P.magic_add) t = ((a - t) >> 1) + t;
    return t >> P.magic_shift;
}

// Kept samples of one warp on their way to the output buffer: staged in shared memory and flushed with ONE global
// atomic per ~64 of them (a global atomic per handful, with the warp waiting for its return every time, is what
// made the collecting pass slow).
constexpr int kKeepCap = 64;
struct KeepStage {
    unsigned* n;     // shared: samples staged by this warp
    long long* q;    // shared: [kKeepCap]
};

// A sample of a kept bin -> the warp's staging; straight to the buffer when the staging is full.  The lanes that
// arrive here together share one reservation.  Out of line (called from the drain of parked samples).
__device__ __noinline__ void keep_sample(long long v, const HistCollect* col, KeepStage keep) {
    const unsigned peers = __activemask();
    const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
    const unsigned cnt = __popc(peers), rank = __popc(peers & ((1u << lane) - 1));
    unsigned pos = 0;
    if (lane == leader) {
        pos = *keep.n;
        if (pos + cnt <= (unsigned)kKeepCap) *keep.n = pos + cnt; else pos = 0xFFFFFFFFu;
    }
    pos = __shfl_sync(peers, pos, leader);
    if (pos != 0xFFFFFFFFu) {
        keep.q[pos + rank] = v;
    } else {
        unsigned long long at = 0;
        if (lane == leader) at = atomicAdd(col->count, (unsigned long long)cnt);
        at = __shfl_sync(peers, at, leader) + rank;
        if (at < col->cap) col->buf[at] = v;
    }
}

// Generic path: any int64 sample.  Out of line, so that the fast loop keeps its registers and instruction cache.
// s == nullptr: count (and collect) only, no moments.  keep: this warp's staging (collecting kernels).
__device__ __noinline__ void slow_sample(long long v, const SlowParams* __restrict__ P, unsigned* mine,
                                         SlowMoments* s, const HistCollect* col, KeepStage keep) {
    int lo = 0, hi = P->nthr;  // c = #{j : thr[j] <= v}: thr[j] <= v for j < lo, thr[j] > v for j >= hi
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (P->thr[mid] <= v) lo = mid + 1; else hi = mid;
    }
    const int c = lo;
    mine[(c + 1) * 32]++;  // row 0 is the scratch row, bin c is row c + 1
    if (col && (c & 1)) keep_sample(v, col, keep);  // odd bins are kept
    if (s) {
        s->mn = v < s->mn ? v : s->mn;
        s->mx = v > s->mx ? v : s->mx;
        const unsigned long long a = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
        const unsigned long long q = udiv_magic(a, *P);
        add128(s->sq, q * q, __umul64hi(q, q));
        if (v < 0) add128(s->sum, 0ull - q, q ? ~0ull : 0ull);
        else add128(s->sum, q, 0);
    }
}

// every thread of the CTA: copy the generic path's parameters; the caller synchronises afterwards
__device__ __forceinline__ void load_slow_params(SlowParams* sp, const HistParams* __restrict__ Pg) {
    for (int i = threadIdx.x; i < kMaxThresholds; i += kHistThreads) sp->thr[i] = Pg->thr[i];
    if (threadIdx.x == 0) {
        sp->magic = Pg->magic;
        sp->nthr = Pg->nthr;
        sp->magic_shift = Pg->magic_shift;
        sp->magic_add = Pg->magic_add;
    }
}

// Shared memory map of the histogram kernel, in ABSOLUTE shared addresses so that the hot loop addresses the
// look-up table and the counters as register + immediate (the compiler's own shared pointers carry a window
// base it keeps recomputing under register pressure).  The driver's reserved 1 KB sits below the dynamic window;
// lat_scratch_init checks the window start once per device (smem_base_probe).
constexpr unsigned kSlowAbs = 0x400;                                    // SlowParams
constexpr unsigned kParkNAbs = kSlowAbs + 640;                          // unsigned[kHistWarps]: parked samples per warp
constexpr unsigned kParkQAbs = kParkNAbs + 64;                          // the parking area, see below
// a plain pass parks ~0.5 % of the samples (multi-second stalls): queue slots per warp
#ifndef B200_PARK_PLAIN
#define B200_PARK_PLAIN 32
#endif
__host__ __device__ constexpr int park_cap(bool) { return B200_PARK_PLAIN; }
// A collecting pass parks ~1 % (everything it keeps).  Its lanes park into slots of their OWN — kLaneLevels entries
// per lane, [level][lane], and one more level behind them that takes what does not fit — which costs four predicated
// instructions per sample and no atomic; the warp moves the slots into its dense queue only when some lane holds
// three, and works the dense queue off 32 entries at a time with every lane busy (hist_kernel).
constexpr int kLaneLevels = 4;
constexpr unsigned kLaneQBytes = (kLaneLevels + 1) * 32 * 8;            // per warp
constexpr int kDenseCap = 64;                                           // per warp: < 32 left over + one level of 32
constexpr unsigned kDenseAbs = kParkQAbs + kHistWarps * kLaneQBytes;    // long long[kHistWarps][kDenseCap]
constexpr unsigned kLaneN0Abs = kDenseAbs + kHistWarps * kDenseCap * 8; // unsigned[kHistThreads]: a lane's slots at tile start
__host__ __device__ constexpr unsigned park_bytes(bool collect) {
    return collect ? kLaneN0Abs + kHistThreads * 4 - kParkQAbs : kHistWarps * park_cap(false) * 8;
}
// unsigned char[lut_bytes], lut_bytes <= kLutSize
__host__ __device__ constexpr unsigned lut_abs(bool collect) { return kParkQAbs + park_bytes(collect); }
// then, at lut_abs + lut_bytes, unsigned[kHistWarps][rows][32] counters, then the tile ring, then (collecting) the
// warps' staging for kept samples
static_assert(sizeof(SlowParams) <= 640, "SlowParams slot");

extern __shared__ __align__(1024) unsigned char hist_smem_raw[];

// B200_DEBUG_ASSERT builds (compute-sanitizer is closed on this pool): every absolute shared-memory access of the
// histogram kernel is checked against the CTA's dynamic window [base, base + %dynamic_smem_size).
#ifdef B200_DEBUG_ASSERT
__device__ __forceinline__ void check_abs(unsigned addr, unsigned bytes) {
    unsigned size;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(size));
    const unsigned base = (unsigned)__cvta_generic_to_shared(hist_smem_raw);
    assert(addr >= base && addr + bytes <= base + size && addr % bytes == 0);
}
#else
__device__ __forceinline__ void check_abs(unsigned, unsigned) {}
#endif


// ---- tile loop: a ring of tiles per WARP in shared memory, filled by 1-D bulk copies (TMA) ---------------------
// A warp's tiles are contiguous 2 KB pieces of the sample array, so the operands of a bulk copy are uniform and
// one lane can issue it.  Bandwidth here is (bytes in flight) / (loaded HBM latency, ~1.4 us): with the next
// tile prefetched into registers a thread has 64 bytes in flight, 48 KB per SM, which is where the register
// version stopped (4.85 TB/s; ncu: warps waiting on their loads, DRAM 54 %, issue 42 %).  The ring keeps nstages
// tiles per warp in flight and costs no registers: every round the lanes copy their samples of the current stage
// into registers, lane 0 re-arms the stage's barrier and issues the copy of the tile nstages rounds on, and only
// then are the samples counted.  The rings are per warp on purpose: one ring per CTA (16 KB tiles, refilled by
// the last warp to have read a stage) keeps the CTA's warps in step, and a warp that is working its parked
// samples off then stalls the other seven (ncu: a fifth of the stall samples at the barrier wait).
// (An L2 bulk prefetch running ahead of register loads was measured too: slower than no prefetch at all.)
constexpr int kWarpTileBytes = 4 * 32 * 16;              // 2 KB: four 128-bit pairs per lane
constexpr int kTileBytes = kHistWarps * kWarpTileBytes;  // ring bytes per stage and CTA
constexpr int kMaxStages = 2;  // measured: a third stage per warp is slower (5.1 vs 5.5 TB/s), one stage much slower
constexpr int kRingHeader = 8 * kHistWarps * kMaxStages;  // the barriers: [warp][stage], 8 B each

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra WAIT_%=;\n\t}"
        ::"r"(bar), "r"(parity) : "memory");
}
// this GPU's own memory only: a plain (non-multicast) bulk copy global -> this CTA's shared memory
__device__ __forceinline__ void bulk_load(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ longlong2 lds_pair(unsigned addr) {
    check_abs(addr, 16);
    longlong2 v;
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr) : "memory");
    return v;
}

// Walks x[0, n): whole warp tiles of 4 x 32 128-bit pairs go to eight(a, b, c, d) — tile t belongs to warp
// t mod (warps in the grid) — everything else (unaligned head, remainder, odd tail) goes to one(v).  after_tile()
// runs once per tile with the whole warp converged.  ring_abs: absolute shared address (128-byte aligned) of
// kRingHeader bytes followed by nstages tile buffers per warp.
template <typename Eight, typename One, typename AfterTile>
__device__ __forceinline__ void for_each_sample(const long long* __restrict__ x, size_t n, unsigned ring_abs,
                                                int nstages, Eight eight, One one, AfterTile after_tile) {
    size_t head = ((reinterpret_cast<uintptr_t>(x) & 15) ? 1 : 0);
    if (head > n) head = n;
    const size_t tid = (size_t)blockIdx.x * kHistThreads + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * kHistThreads;
    if (tid == 0 && head) one(x[0]);
    const longlong2* x2 = reinterpret_cast<const longlong2*>(x + head);
    const size_t npairs = (n - head) / 2;
    constexpr size_t kTilePairs = 4 * 32;
    const size_t ntiles = npairs / kTilePairs;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t gwarp = (size_t)blockIdx.x * kHistWarps + warp, nwarps = (size_t)gridDim.x * kHistWarps;
    const unsigned mine = gwarp < ntiles ? (unsigned)((ntiles - 1 - gwarp) / nwarps) + 1 : 0;
    const unsigned full0 = ring_abs + 8 * kMaxStages * warp;
    const unsigned buf0 = ring_abs + kRingHeader + (unsigned)(warp * nstages) * kWarpTileBytes;
    const unsigned char* const tile0 = reinterpret_cast<const unsigned char*>(x2) + gwarp * kWarpTileBytes;
    const size_t tile_step = nwarps * kWarpTileBytes;
    if (lane == 0) {
        for (int s = 0; s < nstages; s++) mbar_init(full0 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (unsigned j = 0; j < (unsigned)nstages && j < mine; j++) {
            mbar_expect_tx(full0 + 8 * j, kWarpTileBytes);
            bulk_load(buf0 + j * kWarpTileBytes, tile0 + (size_t)j * tile_step, kWarpTileBytes, full0 + 8 * j);
        }
    }
    __syncwarp();
    unsigned s = 0, ph = 0;  // tile j sits in stage j mod nstages, in that stage's (j div nstages)-th use
    for (unsigned j = 0; j < mine; j++) {
        mbar_wait(full0 + 8 * s, ph);
        const unsigned at = buf0 + s * kWarpTileBytes + lane * 16;
        const longlong2 a = lds_pair(at), b = lds_pair(at + 512), c = lds_pair(at + 1024), d = lds_pair(at + 1536);
        __syncwarp();
        if (lane == 0 && j + nstages < mine) {
            // the operands tie the re-arming to the four loads above: it is issued once their data has arrived (the
            // other lanes' loads are the same instructions), and the copy after it
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                         ::"r"(full0 + 8 * s), "r"((unsigned)kWarpTileBytes), "l"(a.x), "l"(b.x), "l"(c.x), "l"(d.x) : "memory");
            bulk_load(buf0 + s * kWarpTileBytes, tile0 + (size_t)(j + nstages) * tile_step, kWarpTileBytes, full0 + 8 * s);
        }
        if (++s == (unsigned)nstages) { s = 0; ph ^= 1; }
        eight(a, b, c, d);
        after_tile();
    }
    for (size_t i = ntiles * kTilePairs + tid; i < npairs; i += nthreads) {
        const longlong2 a = __ldg(x2 + i);
        one(a.x); one(a.y);
    }
    if (tid == 0 && head + 2 * npairs < n) one(x[n - 1]);
}

// End of a histogram kernel: the CTA's private counters -> one 64-bit atomic per bin; the threads' moments -> the
// device totals.  tbl: [kHistWarps][rows][32] counters.
template <bool MOMENTS>
__device__ __forceinline__ void hist_epilogue(const unsigned* tbl, int rows, int nbins, size_t n,
                                              unsigned long long* __restrict__ counts, LatMoments* __restrict__ mom,
                                              long long mn, long long mx, U128 sum, U128 sq) {
    const int lane = threadIdx.x & 31;
    __syncthreads();
    for (int b = threadIdx.x; b < nbins; b += kHistThreads) {
        unsigned long long t = 0;
        for (int w = 0; w < kHistWarps; w++)
            for (int l = 0; l < 32; l++) t += tbl[((size_t)w * rows + b + 1) * 32 + ((l + b) & 31)];  // bin b is row b + 1
        if (t) atomicAdd(counts + b, t);
    }
    if (MOMENTS) {
#pragma unroll
        for (int d = 16; d; d >>= 1) {
            long long omn = __shfl_xor_sync(0xFFFFFFFFu, mn, d), omx = __shfl_xor_sync(0xFFFFFFFFu, mx, d);
            mn = omn < mn ? omn : mn;
            mx = omx > mx ? omx : mx;
            unsigned long long l1 = __shfl_xor_sync(0xFFFFFFFFu, sum.lo, d), h1 = __shfl_xor_sync(0xFFFFFFFFu, sum.hi, d);
            add128(sum, l1, h1);
            unsigned long long l2 = __shfl_xor_sync(0xFFFFFFFFu, sq.lo, d), h2 = __shfl_xor_sync(0xFFFFFFFFu, sq.hi, d);
            add128(sq, l2, h2);
        }
        if (lane == 0) {
            if (mn <= mx) {
                atomicMin(&mom->min_ns, mn);
                atomicMax(&mom->max_ns, mx);
            }
            unsigned long long old = atomicAdd(&mom->sum_lo, sum.lo);
            unsigned long long carry = (old + sum.lo) < old;
            atomicAdd(reinterpret_cast<unsigned long long*>(&mom->sum_hi), sum.hi + carry);
            old = atomicAdd(&mom->sumsq_lo, sq.lo);
            carry = (old + sq.lo) < old;
            atomicAdd(&mom->sumsq_hi, sq.hi + carry);
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&mom->count, (unsigned long long)n);
    }
}

// Samples that need the generic path — negative or >= 2^32 ns (a multi-second stall: about one in 200 of a
// realistic run, so three warp rounds in four meet one somewhere, yet only one thread in 25 does), or in a cell
// with more thresholds than the kernel compares with — are parked in their warp's queue and worked off 32 lanes
// at a time; in the common code their place is taken by a SURROGATE the fast path can digest, whose effect is
// taken out again at the end of the kernel (see the kernels).  What was measured on the way here: masks around
// every fast-path update cost ~4.5 instructions per sample; doing the generic work where it is met, one or two
// lanes active, made the kernel 7x slower; parking the whole group of eight kept the warps of a CTA out of step.
// (A full queue sends the sample down the generic path on the spot.  One queue reservation per thread and group of
// four instead of one per sample was measured: the extra live values spill, slower.)
// Whole warp, converged: hand the parked samples to slow(v), 32 at a time; before_step() runs, converged, before
// every 32.
template <typename Slow, typename BeforeStep>
__device__ __forceinline__ void drain_parked(unsigned* n, const long long* q, unsigned cap, int lane, Slow slow,
                                             BeforeStep before_step, int shift = 0) {
    __syncwarp();
    const unsigned cnt = min(*n >> shift, cap);  // (collecting kernels count the queue in bytes: shift 3)
    for (unsigned base = 0; base < cnt; base += 32) {
        before_step();
        if (base + lane < cnt) slow(q[base + lane]);
        __syncwarp();
    }
    if (lane == 0) *n = 0;
    __syncwarp();
}

// the two halves of a sample, split where the compiler cannot merge the halves' arithmetic back into 64-bit
// operations (an OR of eight high words must not become eight 64-bit ORs)
__device__ __forceinline__ void split64(long long v, unsigned& lo, unsigned& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}

// Collecting kernels park without a branch and without an atomic: about one sample in a hundred is parked (kept, or
// in a cell with a threshold), so three groups of four in four meet one somewhere in the warp; a divergent block
// there was a quarter of the pass, a predicated shared-memory atomic (which ptxas wraps in a branch) still seven
// instructions per sample.  Here: when the row is even (0: the table has no bin for the cell; 2, 4, ...: a kept bin,
// counted on the fast path but still to be collected) the sample is stored in the lane's next slot.  next: address
// of the lane's next slot (levels are 256 bytes apart); at: where the next parked sample goes = min(next, lim), lim
// the level behind the last, so a fifth sample lands there and next > lim says that something was lost.
__device__ __forceinline__ void park_lane_if_even(unsigned row, long long v, unsigned& next, unsigned& at, unsigned lim) {
    check_abs(at, 8);
#ifdef B200_DEBUG_ASSERT
    assert(at <= lim && next >= at);
#endif
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .u32 t;\n\t"
        "and.b32 t, %2, 1;\n\t"
        "setp.eq.u32 p, t, 0;\n\t"
        "@p st.shared.b64 [%1], %3;\n\t"
        "@p add.u32 %0, %0, 256;\n\t"
        "@p min.u32 %1, %0, %4;\n\t}"
        : "+r"(next), "+r"(at) : "r"(row), "l"(v), "r"(lim) : "memory");
}

// the lane's private counter at absolute shared address `addr` += 1
__device__ __forceinline__ void bump_abs(unsigned addr) {
    check_abs(addr, 4);
#ifdef B200_HIST_RED
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(1u) : "memory");
    return;
#endif
    unsigned c;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c) : "r"(addr) : "memory");
    c += 1u;
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(c) : "memory");
}

// resident CTAs per SM the registers allow (two CTAs with 128 registers were measured for the variants that form
// 64-bit squares, which spill a little at 80: slower)
constexpr int hist_ctas_by_regs(bool, bool) { return kHistCtasPerSm; }

// ---- the histogram kernel ---------------------------------------------------------------------------------
// bin(n) = #{j : thr[j] <= n} for ascending thresholds.  For 0 <= n < 2^32 the bin comes from ONE byte look-up,
// lut[index(n)], where index is either
//   BYQ   the quotient q = trunc(n / divisor), which the moments need anyway, when the quotients of 32-bit samples
//         number fewer than kLutSize (divisor >= 2^19: the README bins and OpenCensus bounds in milliseconds), or
//   else  min(n >> lut_shift, kLutSize - 1), the shift chosen on the host so that the thresholds fall inside the
//         table's reach (32 us cells for thresholds below 268 ms).
// A table cell that holds a threshold — or, when the kernel also collects (HistCollect), lies inside a bin that is
// kept — has no bin of its own: its entry is 0, the counter update lands in the scratch row 0, and the sample is
// PARKED for the generic routine (binary search over the thresholds, collection).  With edges given in the
// divisor's own unit no cell holds a threshold; the bracket pass of the select parks about as many samples as it
// keeps (~1 %).  Samples outside 0 <= n < 2^32 are parked as well and replaced, in the common code, by their clamp
// to that range: such a surrogate cannot spoil the extremes (a 0 lowers the minimum only when a negative sample
// exists, which is lower still; 2^32 - 1 raises the maximum only when a sample >= 2^32 exists; a thread that saw
// nothing but surrogates leaves its fast-path extremes out), and the quotient of 2^32 - 1 is taken out of the
// sums at the end (nhi counts them).  So the common code carries no masks and no compares: per sample 1-3
// instructions for the index, 1 look-up, 1 address, 3 for the counter, ~3 for the moments.
// DIVMODE (how the fast path divides a 32-bit sample): 0 divisor is 1, 1 multiply-high magic with the add fix-up,
// 2 divisor >= 2^32 (quotient 0), 3 pre-shift + multiply-high without fix-up and quotient < 2^28; BYQ implies 3
// with quotients < 2^13, whose sums of eight (and of a thread's whole share) fit 32 bits.
// MOMENTS: 0 none; 1 count, extremes, sum and sum of squares; 2 the same without the squares (what the seven
// ssd_test lines need: three instructions per sample less)
template <int MOMENTS, int DIVMODE, bool BYQ, bool COLLECT>
__global__ void __launch_bounds__(kHistThreads, hist_ctas_by_regs(MOMENTS != 0, BYQ))
hist_kernel(const long long* __restrict__ x, size_t n, const HistParams* __restrict__ Pg,
            unsigned long long* __restrict__ counts, LatMoments* __restrict__ mom, const HistCollect col, int nstages) {
    static_assert(!BYQ || DIVMODE == 3, "the quotient index uses the pre-shift magic");
    constexpr unsigned kLutAbs = lut_abs(COLLECT), kParkCap = park_cap(COLLECT);
    constexpr bool collecting = COLLECT;
    const unsigned smem_base = (unsigned)__cvta_generic_to_shared(hist_smem_raw);
    unsigned char* const sm0 = hist_smem_raw - smem_base;  // generic pointer to absolute shared address 0
    const unsigned lut_bytes = (unsigned)Pg->lut_bytes, kCntAbs = kLutAbs + lut_bytes;
    unsigned* const tbl = reinterpret_cast<unsigned*>(sm0 + kCntAbs);  // [warp][row][lane] private counters
    SlowParams* const sp = reinterpret_cast<SlowParams*>(sm0 + kSlowAbs);
    load_slow_params(sp, Pg);
    if (threadIdx.x < kHistWarps) reinterpret_cast<unsigned*>(sm0 + kParkNAbs)[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < (int)lut_bytes / 4; i += kHistThreads)
        reinterpret_cast<unsigned*>(sm0 + kLutAbs)[i] = reinterpret_cast<const unsigned*>(Pg->lut)[i];
    const int nbins = Pg->nthr + 1, rows = nbins + 1;  // row 0: scratch
    for (int i = threadIdx.x; i < kHistWarps * rows * 32; i += kHistThreads) tbl[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned* const mine = tbl + (size_t)warp * rows * 32 + lane;
    const unsigned mine_abs = kCntAbs + (unsigned)(warp * rows * 32 + lane) * 4u;
    const unsigned magic32 = Pg->magic32;
    // DIVMODE 3 works on u = n >> pre32 (the divisor's power-of-two part shifted out of both), so its fast path
    // reaches n < 2^reach_bits with reach_bits up to 32 + pre32 (2^37 ns = 137 s for milliseconds): the shift is the
    // funnel shift that splits the sample anyway.  Other modes: u = the low word, reach 2^32.
    const int shift32 = Pg->shift32, pre = DIVMODE == 3 ? Pg->pre32 : 0, idx_shift = Pg->lut_shift - pre;
    const int reach_bits = Pg->reach_bits, out_shift = reach_bits - 32;  // outside: (high word >> out_shift) != 0
    const unsigned u_top = reach_bits - pre >= 32 ? 0xFFFFFFFFu : (1u << (reach_bits - pre)) - 1;  // largest fast-path u

    Moments m;          // 64-bit sums with a carry count
    SlowMoments sm_slow;
    unsigned nhi = 0, nsur = 0, nproc = 0;  // surrogates for samples beyond the reach; all surrogates; fast-path samples

    // collecting kernels: the warp's staging for kept samples sits past the tile ring
    const unsigned keep_abs = ((kCntAbs + (unsigned)(kHistWarps * rows * 32) * 4u + 127u) & ~127u) + kRingHeader +
                              (unsigned)nstages * kTileBytes;
    KeepStage keep{reinterpret_cast<unsigned*>(sm0 + keep_abs) + warp,
                   reinterpret_cast<long long*>(sm0 + keep_abs + 64) + (size_t)warp * kKeepCap};
    // the warp's queue of parked samples: addressed from constants where it is used (no registers held for it)
    auto park_n = [&]() -> unsigned* { return reinterpret_cast<unsigned*>(sm0 + kParkNAbs) + warp; };
    auto park_q = [&]() -> long long* { return reinterpret_cast<long long*>(sm0 + kParkQAbs) + (size_t)warp * kParkCap; };
    // collecting: the lane's own slots (level L at lane_base + 256 L), where its next parked sample goes, the warp's
    // dense queue (counted in park_n) and the word that remembers the lane's slots at the start of the tile
    const unsigned lane_base = kParkQAbs + kLaneQBytes * warp + 8u * lane, lane_lim = lane_base + 256u * kLaneLevels;
    unsigned lane_next = lane_base, lane_at = lane_base;
    long long* const dense = reinterpret_cast<long long*>(sm0 + kDenseAbs) + (size_t)warp * kDenseCap;
    unsigned* const lane_n0 = reinterpret_cast<unsigned*>(sm0 + kLaneN0Abs) + threadIdx.x;
    if (COLLECT) *lane_n0 = lane_base;
    if (collecting && lane == 0) *keep.n = 0;
    __syncwarp();
    auto flush_keep = [&]() {  // whole warp, converged
        __syncwarp();
        const unsigned total = *keep.n;
        unsigned long long at = 0;
        if (lane == 0 && total) at = atomicAdd(col.count, (unsigned long long)total);
        at = __shfl_sync(0xFFFFFFFFu, at, 0);
        for (unsigned k = lane; k < total; k += 32)
            if (at + k < col.cap) col.buf[at + k] = keep.q[k];
        __syncwarp();
        if (lane == 0) *keep.n = 0;
        __syncwarp();
    };
    auto slow = [&](long long v) {  // parked: outside the reach -> everything; else its bin (and collection) only
        const bool outside = ((unsigned long long)v >> reach_bits) != 0;
        slow_sample(v, sp, mine, MOMENTS && outside ? &sm_slow : nullptr, collecting ? &col : nullptr, keep);
    };
    auto quotient = [&](unsigned u) -> unsigned {
        if (DIVMODE == 1) {
            const unsigned t = __umulhi(magic32, u);
            return (t + ((u - t) >> 1)) >> shift32;
        }
        if (DIVMODE == 3) return __umulhi(u, magic32) >> shift32;  // u is already n >> pre32
        return DIVMODE == 0 ? u : 0u;
    };
    const unsigned last_cell = lut_bytes - 1;
    // q: the quotient when BYQ (kept for the moments), else unused
    auto index_of = [&](unsigned u, unsigned& q) -> unsigned {
        if (BYQ) {
            q = quotient(u);
            return min(q, last_cell);
        }
        return min(u >> idx_shift, last_cell);
    };
    auto shifted = [&](unsigned lo, unsigned hi) -> unsigned { return DIVMODE == 3 ? __funnelshift_r(lo, hi, pre) : lo; };
    // (the table is read-only while the kernel counts: a plain asm, so that the eight look-ups of a round are
    // issued together instead of each waiting for the previous sample's counter update)
    auto row_of = [&](unsigned idx) -> unsigned {
        unsigned r;
        check_abs(idx + kLutAbs, 1);
        asm("ld.shared.u8 %0, [%1+%2];" : "=r"(r) : "r"(idx), "n"(kLutAbs));
        return r;
    };
    auto accumulate = [&](unsigned q) {  // running 64-bit sums; q*q wraps at most once per add
        m.fsum += q;
        if (MOMENTS == 1) {
            const unsigned long long sq = (unsigned long long)q * q + m.fsq;
            m.fsq_carry += sq < m.fsq;
            m.fsq = sq;
        }
    };
    auto one = [&](long long v) {
        if ((unsigned long long)v >> reach_bits) { slow(v); return; }
        unsigned lo, hi, q = 0;
        split64(v, lo, hi);
        const unsigned u = shifted(lo, hi), row = row_of(index_of(u, q));
        if (row == 0) slow(v);  // its bin; the moments below are the fast path's
        else if (COLLECT && !(row & 1)) keep_sample(v, &col, keep);  // a kept bin the table knows
        bump_abs(mine_abs + (row << 7));
        nproc++;
        if (MOMENTS) {
            m.mn32 = min(m.mn32, u);
            m.mx32 = max(m.mx32, u);
            accumulate(BYQ ? q : quotient(u));
        }
    };
    // Four samples at a time (two such groups per round: eight at once keep too many values alive for the 80
    // registers three CTAs per SM leave a thread).  Whether any of them must be parked is one test.
    auto four = [&](const longlong2 a, const longlong2 b) {
        constexpr int K = 4;
        const long long v[K] = {a.x, a.y, b.x, b.y};
        unsigned u[K], hi[K], q[K], idx[K], row[K], any_hi = 0;
#pragma unroll
        for (int k = 0; k < K; k++) {
            unsigned lo;
            split64(v[k], lo, hi[k]);
            u[k] = shifted(lo, hi[k]);
            any_hi |= hi[k];
        }
#pragma unroll
        for (int k = 0; k < K; k++) idx[k] = index_of(u[k], q[k]);
#pragma unroll
        for (int k = 0; k < K; k++) row[k] = row_of(idx[k]);
        if constexpr (COLLECT) {
#pragma unroll
            for (int k = 0; k < K; k++) park_lane_if_even(row[k], v[k], lane_next, lane_at, lane_lim);
            if ((any_hi >> out_shift) != 0) {  // beyond the reach: parked whatever its table entry said; surrogate
#pragma unroll
                for (int k = 0; k < K; k++)
                    if ((hi[k] >> out_shift) != 0) {
                        if (row[k] & 1) park_lane_if_even(0u, v[k], lane_next, lane_at, lane_lim);
                        const bool neg = (int)hi[k] < 0;
                        u[k] = neg ? 0u : u_top;
                        if (BYQ) q[k] = quotient(u[k]);
                        row[k] = 0;
                        nhi += !neg;
                        nsur++;
                    }
            }
        } else if ((any_hi >> out_shift) != 0 || min(min(row[0], row[1]), min(row[2], row[3])) == 0) {  // rare
#pragma unroll
            for (int k = 0; k < K; k++)
                if ((hi[k] >> out_shift) != 0 || row[k] == 0) {
                    const unsigned at = atomicAdd(park_n(), 1u);
                    if (at < kParkCap) park_q()[at] = v[k];
                    else slow(v[k]);  // queue full: on the spot
                    if (hi[k] >> out_shift) {  // the surrogate: clamp to the reach, counted in the scratch row
                        const bool neg = (int)hi[k] < 0;
                        u[k] = neg ? 0u : u_top;
                        if (BYQ) q[k] = quotient(u[k]);
                        row[k] = 0;
                        nhi += !neg;
                        nsur++;
                    }
                }
        }
#pragma unroll
        for (int k = 0; k < K; k++) bump_abs(mine_abs + (row[k] << 7));
        if (MOMENTS) {
            m.mn32 = min(min(m.mn32, u[0]), u[1]);
            m.mn32 = min(min(m.mn32, u[2]), u[3]);
            m.mx32 = max(max(m.mx32, u[0]), u[1]);
            m.mx32 = max(max(m.mx32, u[2]), u[3]);
            if (DIVMODE == 3) {  // quotients < 2^28: a group's sum fits 32 bits, its squares 64
                unsigned s = 0;
                unsigned long long sq = 0;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    const unsigned qq = BYQ ? q[k] : quotient(u[k]);
                    s += qq;
                    if (MOMENTS == 1) sq += (unsigned long long)qq * qq;
                }
                m.fsum += s;
                if (MOMENTS == 1) {
                    sq += m.fsq;
                    m.fsq_carry += sq < m.fsq;
                    m.fsq = sq;
                }
            } else {
#pragma unroll
                for (int k = 0; k < K; k++) accumulate(quotient(u[k]));
            }
        }
    };
    const unsigned ring_abs = (kCntAbs + (unsigned)(kHistWarps * rows * 32) * 4u + 127u) & ~127u;
    // a drain step stages at most 32 kept samples: flush before a step that might not fit
    auto before_step = [&] {
        if (collecting && *keep.n > (unsigned)(kKeepCap - 32)) flush_keep();
    };
    // ---- collecting kernels: working the parked samples off ---------------------------------------------------
    // work_off(count): the first `count` (<= 32) entries of the dense queue, one per lane, all lanes in step.  An
    // entry in reach gets its exact bin from a branch-free binary search over the thresholds (the same trip count in
    // every lane); if the table had no bin for its cell the fast path counted it in the scratch row, so its counter
    // is bumped now; samples of kept (odd) bins are staged with one ballot.  Entries beyond the reach — rare — take
    // the generic routine, which counts, sums and collects on its own.
    auto work_off = [&](unsigned count) {
        before_step();
        const bool in = (unsigned)lane < count;
        const long long v = in ? dense[lane] : 0;
        const bool outside = in && ((unsigned long long)v >> reach_bits) != 0;
        int c = 0;
        for (int step = 64; step; step >>= 1) {  // kMaxThresholds <= 127
            const int j = c + step;
            if (j <= sp->nthr && sp->thr[j - 1] <= v) c = j;
        }
        bool kept = false;
        if (in && !outside) {
            unsigned lo, hi, q = 0;
            split64(v, lo, hi);
            if (row_of(index_of(shifted(lo, hi), q)) == 0) mine[(c + 1) * 32]++;
            kept = c & 1;
        }
        if (outside) slow(v);
        __syncwarp();
        const unsigned km = __ballot_sync(0xFFFFFFFFu, kept);
        if (km) {
            const unsigned pos = *keep.n;  // (room for 32: before_step)
            if (kept) keep.q[pos + __popc(km & ((1u << lane) - 1))] = v;
            __syncwarp();
            if (lane == 0) *keep.n = pos + __popc(km);
            __syncwarp();
        }
    };
    // the lanes' slots -> the dense queue, level by level; every full group of 32 is worked off at once
    auto compact = [&] {  // whole warp, converged
        unsigned cnt = *park_n();
#pragma unroll 1
        for (unsigned level = lane_base; level != lane_lim; level += 256u) {
            const bool has = lane_next > level;
            const unsigned m = __ballot_sync(0xFFFFFFFFu, has);
            if (!m) break;
            check_abs(level, 8);
#ifdef B200_DEBUG_ASSERT
            assert(cnt + __popc(m) <= (unsigned)kDenseCap);
#endif
            if (has) dense[cnt + __popc(m & ((1u << lane) - 1))] = *reinterpret_cast<const long long*>(sm0 + level);
            cnt += __popc(m);
            __syncwarp();
            if (cnt >= 32) {
                work_off(32);
                cnt -= 32;
                const long long rest = (unsigned)lane < cnt ? dense[32 + lane] : 0;
                __syncwarp();
                if ((unsigned)lane < cnt) dense[lane] = rest;
                __syncwarp();
            }
        }
        lane_next = lane_at = lane_base;
        if (lane == 0) *park_n() = cnt;
        __syncwarp();
    };
    auto drain = [&] {  // end of the kernel; whole warp, converged
        if constexpr (!COLLECT) {
            drain_parked(park_n(), park_q(), kParkCap, lane, slow, before_step);
        } else {
            compact();
            const unsigned cnt = *park_n();
            if (cnt) work_off(cnt);
            __syncwarp();
            if (lane == 0) *park_n() = 0;
            __syncwarp();
        }
    };
    auto eight = [&](const longlong2 a, const longlong2 b, const longlong2 c, const longlong2 d) {
        four(a, b);
        four(c, d);
        nproc += 8;
        if constexpr (COLLECT) {
            // once per tile: does some lane hold three?  (A lane with two can take two more; a lane that parks more
            // in one tile than it had room for — a few times in 10^5 lane-tiles — forgets this tile's entries and
            // takes its parked samples of the tile one by one from the registers.)
            if (__any_sync(0xFFFFFFFFu, lane_next >= lane_base + 256u * (kLaneLevels - 1))) {
                if (lane_next > lane_lim) {
                    lane_next = *lane_n0;
                    const long long v[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
#pragma unroll
                    for (int k = 0; k < 8; k++) {
                        const bool outside = ((unsigned long long)v[k] >> reach_bits) != 0;
                        unsigned row = 1;
                        if (!outside) {
                            unsigned lo, hi, q = 0;
                            split64(v[k], lo, hi);
                            row = row_of(index_of(shifted(lo, hi), q));
                        }
                        if (outside || row == 0) slow(v[k]);
                        else if (!(row & 1)) keep_sample(v[k], &col, keep);
                    }
                }
                __syncwarp();
                compact();
            }
            *lane_n0 = lane_next;
        }
    };
    for_each_sample(x, n, ring_abs, nstages, eight, one, [&] {
        if constexpr (!COLLECT) {
            __syncwarp();
            if (*park_n() >= kParkCap / 2)  // every lane reads the same counter
                drain_parked(park_n(), park_q(), kParkCap, lane, slow, before_step);
        }
    });
    drain();
    if (collecting) flush_keep();

    long long mn = sm_slow.mn, mx = sm_slow.mx;
    if (nproc > nsur) {  // this thread saw genuine fast-path samples
        // the extremes were kept of u = n >> pre: as nanoseconds, the bounds with the same quotients
        const long long lo_ns = (long long)((unsigned long long)m.mn32 << pre);
        const long long hi_ns = (long long)(((unsigned long long)m.mx32 << pre) | ((1ull << pre) - 1));
        mn = lo_ns < mn ? lo_ns : mn;
        mx = hi_ns > mx ? hi_ns : mx;
    }
    U128 sum = sm_slow.sum, sq = sm_slow.sq;
    add128(sum, m.fsum, 0);
    add128(sq, m.fsq, m.fsq_carry);
    if (MOMENTS && nhi) {  // take the surrogates of the samples beyond the reach out of the sums again: nhi * q(u_top)
        const unsigned long long qhi = quotient(u_top), q2 = qhi * qhi;
        const unsigned long long s_lo = (unsigned long long)nhi * qhi;  // < 2^64
        const unsigned long long p_lo = (unsigned long long)nhi * q2, p_hi = __umul64hi((unsigned long long)nhi, q2);
        const unsigned long long b1 = sum.lo < s_lo;
        sum.lo -= s_lo;
        sum.hi -= b1;
        const unsigned long long b2 = sq.lo < p_lo;
        sq.lo -= p_lo;
        sq.hi -= p_hi + b2;
    }
    hist_epilogue<MOMENTS != 0>(tbl, rows, nbins, n, counts, mom, mn, mx, sum, sq);
}

constexpr int kMinmaxThreads = 256;
__global__ void __launch_bounds__(kMinmaxThreads)
minmax_kernel(const long long* __restrict__ x, size_t n, LatMoments* __restrict__ mom) {
    long long mn = 0x7FFFFFFFFFFFFFFFll, mx = (long long)0x8000000000000000ull;
    auto one = [&](long long v) {
        mn = v < mn ? v : mn;
        mx = v > mx ? v : mx;
    };
    size_t head = ((reinterpret_cast<uintptr_t>(x) & 15) ? 1 : 0);
    if (head > n) head = n;
    const size_t tid = (size_t)blockIdx.x * kMinmaxThreads + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * kMinmaxThreads;
    if (tid == 0 && head) one(x[0]);
    const longlong2* x2 = reinterpret_cast<const longlong2*>(x + head);
    const size_t npairs = (n - head) / 2;
    size_t i = tid;
    for (; i + 3 * nthreads < npairs; i += 4 * nthreads) {
        longlong2 a = __ldg(x2 + i), b = __ldg(x2 + i + nthreads), c = __ldg(x2 + i + 2 * nthreads),
                  d = __ldg(x2 + i + 3 * nthreads);
        one(a.x); one(a.y); one(b.x); one(b.y); one(c.x); one(c.y); one(d.x); one(d.y);
    }
    for (; i < npairs; i += nthreads) {
        longlong2 a = __ldg(x2 + i);
        one(a.x); one(a.y);
    }
    if (tid == 0 && head + 2 * npairs < n) one(x[n - 1]);
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        long long omn = __shfl_xor_sync(0xFFFFFFFFu, mn, d), omx = __shfl_xor_sync(0xFFFFFFFFu, mx, d);
        mn = omn < mn ? omn : mn;
        mx = omx > mx ? omx : mx;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&mom->min_ns, mn);
        atomicMax(&mom->max_ns, mx);
    }
}

// Keep only the samples of the (at most kMaxSelect distinct) buckets the wanted ranks fell into.  Hits are
// staged per CTA in shared memory and flushed in blocks, so the global counter sees one atomic per ~3000 hits
// (a warp-level append alone serialises on that counter when 10-20 % of the samples survive).
constexpr int kCompactStage = 4096;                        // samples staged per CTA
constexpr int kCompactFlush = kCompactStage - 2 * kRadixThreads;  // a round adds at most 2 per thread

__global__ void __launch_bounds__(kRadixThreads)
compact_kernel(const long long* __restrict__ x, size_t n, unsigned long long key_base, int top, int nprefix,
               const unsigned long long* __restrict__ prefix, long long* __restrict__ out,
               unsigned long long* __restrict__ out_n, size_t cap) {
    __shared__ long long stage[kCompactStage];
    __shared__ unsigned stage_n;
    __shared__ unsigned long long flush_base;
    unsigned long long pf[kMaxSelect];
#pragma unroll
    for (int k = 0; k < kMaxSelect; k++) pf[k] = k < nprefix ? prefix[k] : ~0ull;  // key' >> top < 2^63: never matches
    if (threadIdx.x == 0) stage_n = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    auto emit = [&](bool hit, long long v) {  // called by all 32 lanes together
        const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
        if (m) {
            const int leader = __ffs(m) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(&stage_n, (unsigned)__popc(m));
            base = __shfl_sync(0xFFFFFFFFu, base, leader);
            if (hit) stage[base + __popc(m & ((1u << lane) - 1))] = v;
        }
    };
    auto flush = [&]() {  // whole CTA
        __syncthreads();
        const unsigned cnt = stage_n;
        if (threadIdx.x == 0) flush_base = atomicAdd(out_n, (unsigned long long)cnt);
        __syncthreads();
        for (unsigned k = threadIdx.x; k < cnt; k += kRadixThreads)
            if (flush_base + k < cap) out[flush_base + k] = stage[k];
        __syncthreads();
        if (threadIdx.x == 0) stage_n = 0;
        __syncthreads();
    };
    auto matches = [&](long long v) {
        const unsigned long long hi = (((unsigned long long)v ^ 0x8000000000000000ull) - key_base) >> top;
        return hi == pf[0] || hi == pf[1] || hi == pf[2] || hi == pf[3];
    };
    // every thread makes the same number of trips (ballots and CTA barriers); 16-byte loads in the body
    const size_t head = (reinterpret_cast<uintptr_t>(x) & 15) && n ? 1 : 0;
    const longlong2* x2 = reinterpret_cast<const longlong2*>(x + head);
    const size_t npairs = (n - head) / 2;
    const size_t nthreads = (size_t)gridDim.x * kRadixThreads;
    const size_t tid = (size_t)blockIdx.x * kRadixThreads + threadIdx.x;
    const size_t rounds = (npairs + nthreads - 1) / nthreads;
    size_t i = tid;
    for (size_t r = 0; r < rounds; r++, i += nthreads) {
        longlong2 a = make_longlong2(0, 0);
        const bool in = i < npairs;
        if (in) a = __ldg(x2 + i);
        emit(in && matches(a.x), a.x);
        emit(in && matches(a.y), a.y);
        __syncthreads();
        if (stage_n > kCompactFlush) flush();  // CTA-uniform: everyone reads the same counter after the barrier
    }
    // the unaligned first sample and an odd last one: warp 0 of block 0, lanes 0 and 1
    if (blockIdx.x == 0 && threadIdx.x < 32) {
        long long v = 0;
        bool hit = false;
        if (lane == 0 && head) { v = x[0]; hit = matches(v); }
        if (lane == 1 && head + 2 * npairs < n) { v = x[n - 1]; hit = matches(v); }
        emit(hit, v);
    }
    flush();
}

__global__ void __launch_bounds__(kRadixThreads)
radix_kernel(const long long* __restrict__ x, size_t n, unsigned long long key_base, int shift, int top, int nk,
             const unsigned long long* __restrict__ prefix, const unsigned long long* __restrict__ bases,
             unsigned long long* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char sm[];
    unsigned* hist = reinterpret_cast<unsigned*>(sm);  // [nk][kRadixBins]
    for (int i = threadIdx.x; i < nk * kRadixBins; i += kRadixThreads) hist[i] = 0;
    __shared__ unsigned long long pf[kMaxSelect], kb[kMaxSelect];  // prefix and key base of every histogram
    if (threadIdx.x < nk) {
        pf[threadIdx.x] = top >= 64 ? 0 : prefix[threadIdx.x];
        kb[threadIdx.x] = bases ? bases[threadIdx.x] : key_base;
    }
    __syncthreads();
    const unsigned mask = (top - shift >= kRadixBits) ? (kRadixBins - 1) : ((1u << (top - shift)) - 1);
    const bool all = top >= 64;
    const size_t nthreads = (size_t)gridDim.x * kRadixThreads;
    auto one = [&](long long v) {
        const unsigned long long raw = (unsigned long long)v ^ 0x8000000000000000ull;
#pragma unroll
        for (int k = 0; k < kMaxSelect; k++) {
            if (k >= nk) break;
            const unsigned long long key = raw - kb[k];
            if (all || (key >> top) == pf[k]) atomicAdd(hist + k * kRadixBins + ((unsigned)(key >> shift) & mask), 1u);
        }
    };
    size_t head = ((reinterpret_cast<uintptr_t>(x) & 15) ? 1 : 0);
    if (head > n) head = n;
    const size_t tid = (size_t)blockIdx.x * kRadixThreads + threadIdx.x;
    if (tid == 0 && head) one(x[0]);
    const longlong2* x2 = reinterpret_cast<const longlong2*>(x + head);
    const size_t npairs = (n - head) / 2;
    size_t i = tid;
    for (; i + nthreads < npairs; i += 2 * nthreads) {
        longlong2 a = __ldg(x2 + i), b = __ldg(x2 + i + nthreads);
        one(a.x); one(a.y); one(b.x); one(b.y);
    }
    for (; i < npairs; i += nthreads) {
        longlong2 a = __ldg(x2 + i);
        one(a.x); one(a.y);
    }
    if (tid == 0 && head + 2 * npairs < n) one(x[n - 1]);
    __syncthreads();
    for (int i2 = threadIdx.x; i2 < nk * kRadixBins; i2 += kRadixThreads) {
        unsigned c = hist[i2];
        if (c) atomicAdd(out + i2, (unsigned long long)c);
    }
}

// ---- device-resident radix select (see SelState) ------------------------------------------------------------

__global__ void sel_init_kernel(SelState* st, const LatMoments* __restrict__ mom, SelRanks ranks, int have_bounds,
                                long long lo, long long hi, unsigned long long n) {
    if (threadIdx.x || blockIdx.x) return;
    if (!have_bounds) {
        lo = mom->min_ns;
        hi = mom->max_ns;
    }
    const unsigned long long kmin = (unsigned long long)lo ^ 0x8000000000000000ull,
                             kmax = (unsigned long long)hi ^ 0x8000000000000000ull;
    const unsigned long long range = kmax - kmin;
    st->key_base = kmin;
    st->top = lo > hi ? 0 : 64 - __clzll((long long)range);
    st->nk = ranks.nk;
    st->nhist = 1;
    st->bad = lo > hi;
    for (int k = 0; k < kSelMax; k++) {
        st->prefix[k] = 0;
        st->rank[k] = k < ranks.nk ? ranks.rank[k] : 0;
        st->result[k] = lo;  // all samples equal: nothing to resolve
        if (k < ranks.nk && ranks.rank[k] >= n) st->bad = 1;
    }
}

__global__ void __launch_bounds__(kRadixThreads)
sel_pass_kernel(const long long* __restrict__ x, size_t n, const SelState* __restrict__ st,
                unsigned long long* __restrict__ out) {
    const int top = st->top;
    if (top == 0 || st->bad) return;
    extern __shared__ __align__(16) unsigned char sm[];
    unsigned* hist = reinterpret_cast<unsigned*>(sm);  // [nhist][kRadixBins]
    const int nhist = st->nhist, nk = st->nk;
    for (int i = threadIdx.x; i < nhist * kRadixBins; i += kRadixThreads) hist[i] = 0;
    __shared__ unsigned long long pf[kSelMax];
    if (threadIdx.x < kSelMax) pf[threadIdx.x] = threadIdx.x < nk ? st->prefix[threadIdx.x] : ~0ull;
    __syncthreads();
    const int shift = top > kRadixBits ? top - kRadixBits : 0;
    const unsigned mask = (1u << (top - shift)) - 1;
    const unsigned long long key_base = st->key_base;
    auto one = [&](long long v) {
        const unsigned long long key = ((unsigned long long)v ^ 0x8000000000000000ull) - key_base;
        const unsigned dgt = (unsigned)(key >> shift) & mask;
        const unsigned long long hi = top >= 64 ? 0 : key >> top;
        if (nhist == 1) {  // first pass (every sample has hi == 0 == every prefix: one histogram serves all), or nk == 1
            if (hi == pf[0]) atomicAdd(hist + dgt, 1u);
        } else {
#pragma unroll
            for (int k = 0; k < kSelMax; k++)
                if (hi == pf[k]) atomicAdd(hist + k * kRadixBins + dgt, 1u);
        }
    };
    size_t head = ((reinterpret_cast<uintptr_t>(x) & 15) ? 1 : 0);
    if (head > n) head = n;
    const size_t tid = (size_t)blockIdx.x * kRadixThreads + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * kRadixThreads;
    if (tid == 0 && head) one(x[0]);
    const longlong2* x2 = reinterpret_cast<const longlong2*>(x + head);
    const size_t npairs = (n - head) / 2;
    size_t i = tid;
    for (; i + nthreads < npairs; i += 2 * nthreads) {
        const longlong2 a = __ldg(x2 + i), b = __ldg(x2 + i + nthreads);
        one(a.x); one(a.y); one(b.x); one(b.y);
    }
    for (; i < npairs; i += nthreads) {
        const longlong2 a = __ldg(x2 + i);
        one(a.x); one(a.y);
    }
    if (tid == 0 && head + 2 * npairs < n) one(x[n - 1]);
    __syncthreads();
    for (int i2 = threadIdx.x; i2 < nhist * kRadixBins; i2 += kRadixThreads) {
        const unsigned c = hist[i2];
        if (c) atomicAdd(out + i2, (unsigned long long)c);
    }
}

// One CTA: for every rank find the digit whose bucket holds it, descend, and clear the histograms.
__global__ void __launch_bounds__(1024)
sel_scan_kernel(SelState* st, unsigned long long* __restrict__ hist) {
    const int top = st->top;
    if (top == 0 || st->bad) return;
    const int nk = st->nk, nhist = st->nhist;
    const int shift = top > kRadixBits ? top - kRadixBits : 0, width = top - shift;
    __shared__ unsigned long long warp_tot[kSelMax][32];
    __shared__ unsigned long long new_prefix[kSelMax], new_rank[kSelMax], rank[kSelMax];
    __shared__ int found[kSelMax];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t < kSelMax) {
        found[t] = 0;
        rank[t] = st->rank[t];
    }
    // Four ranks at a time, their loads issued together (one rank after the other is a chain of dependent global
    // loads: 18 us per launch); a thread owns bins [4t, 4t+4).  Everything is indexed statically to stay in registers.
    constexpr int kBatch = 4;
    for (int kb = 0; kb < nk; kb += kBatch) {
        unsigned long long c[kBatch][4], sum[kBatch], incl[kBatch];
#pragma unroll
        for (int k = 0; k < kBatch; k++)
#pragma unroll
            for (int j = 0; j < 4; j++)
                c[k][j] = kb + k < nk ? hist[(size_t)(nhist == 1 ? 0 : kb + k) * kRadixBins + 4 * t + j] : 0;
#pragma unroll
        for (int k = 0; k < kBatch; k++) {
            sum[k] = c[k][0] + c[k][1] + c[k][2] + c[k][3];
            incl[k] = sum[k];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, incl[k], d);
                if (lane >= d) incl[k] += o;
            }
            if (lane == 31) warp_tot[kb + k][warp] = incl[k];
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kBatch; k++) {
            if (kb + k < nk) {
                unsigned long long before = incl[k] - sum[k];
                for (int w = 0; w < warp; w++) before += warp_tot[kb + k][w];
                const unsigned long long r = rank[kb + k];
                if (r >= before && r < before + sum[k]) {  // exactly one thread
                    unsigned long long cum = before;
                    int j = 0;
#pragma unroll
                    for (int jj = 0; jj < 3; jj++)
                        if (j == jj && r >= cum + c[k][jj]) { cum += c[k][jj]; j++; }
                    new_prefix[kb + k] = (st->prefix[kb + k] << width) | (unsigned long long)(4 * t + j);
                    new_rank[kb + k] = r - cum;
                    found[kb + k] = 1;
                }
            }
        }
    }
    __syncthreads();
    for (int i = t; i < nhist * kRadixBins; i += 1024) hist[i] = 0;
    if (t == 0) {
        int bad = 0;
        for (int k = 0; k < nk; k++) {
            if (!found[k]) { bad = 1; continue; }
            st->prefix[k] = new_prefix[k];
            st->rank[k] = new_rank[k];
            if (shift == 0) st->result[k] = (long long)((new_prefix[k] + st->key_base) ^ 0x8000000000000000ull);
        }
        st->top = shift;
        st->nhist = nk;
        st->bad = bad;
    }
}

// ---- the same select in ONE kernel -----------------------------------------------------------------------------
// The launches above cost more than their work on inputs of a few million samples (the sample of a sampled select,
// the samples it keeps): min/max + init + six passes + six scans, ~12 us each with their gaps, for data that sits
// in L2.  Here one cooperative grid (a CTA per SM, all resident) does the lot with grid-wide barriers: every round
// the CTAs count their share into shared memory and add it to the global histograms; after the barrier CTA h scans
// histogram h (one histogram for all ranks on the first round, one per rank afterwards), writes the ranks' new
// prefixes into the state and clears the histogram; after a second barrier everyone reads the prefixes.
// approx_max > 0: a rank stops being refined once its bucket holds at most approx_max samples, and the result is
// the bucket's lowest key — or its highest for the ranks in ceil_mask — i.e. a bound on the order statistic good to
// approx_max ranks, which is all the brackets of the sampled select need (one round less).
constexpr int kFusedThreads = 1024, kScanBins = kRadixBins / kFusedThreads;

// Barrier over all CTAs of a grid that is resident as a whole (the launch is cooperative, which is what guarantees
// that).  bar[0] counts arrivals and is cleared by the last to arrive, bar[1] is the generation it then advances:
// both are left ready for the next barrier and the next launch (cleared once, at set-up).  The fences on either side
// make what the CTAs wrote before the barrier visible to all of them after it.
__device__ __forceinline__ void grid_barrier(unsigned* bar) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned gen = *reinterpret_cast<volatile unsigned*>(bar + 1);
        if (atomicAdd(bar, 1u) == gridDim.x - 1) {
            *reinterpret_cast<volatile unsigned*>(bar) = 0;
            __threadfence();
            atomicAdd(bar + 1, 1u);
        } else {
            while (*reinterpret_cast<volatile unsigned*>(bar + 1) == gen) {
            }
        }
        __threadfence();
    }
    __syncthreads();
}
// bounds (approx_max > 0) need not be finer than 2^kApproxBits key units (4 us for nanoseconds): identical samples
// — a latency distribution has atoms — would otherwise keep a bucket above approx_max down to the last bit
constexpr int kApproxBits = 12;
__global__ void __launch_bounds__(kFusedThreads)
sel_fused_kernel(const long long* __restrict__ x, size_t n_arg, SelState* st, unsigned long long* __restrict__ hist,
                 LatMoments* mom, SelRanks ranks_arg, int have_bounds, long long lo, long long hi,
                 unsigned long long approx_max, unsigned ceil_mask, const SelDeferred* __restrict__ in, unsigned* bar) {
    // ranks and length either from the launch or from what the kernel before this one left (the same for every CTA)
    size_t n = n_arg;
    SelRanks ranks = ranks_arg;
    if (in) {
        if (in->miss) {
            if (blockIdx.x == 0 && threadIdx.x == 0) {
                st->bad = 1;
                st->top = 0;
                st->nstamp = 0;
            }
            return;
        }
        n = (size_t)in->n;
        ranks = in->ranks;
    }
    static_assert(kRadixBins == kScanBins * kFusedThreads, "a thread scans kScanBins bins");
    extern __shared__ __align__(16) unsigned char sm[];
    unsigned* hs = reinterpret_cast<unsigned*>(sm);  // [nhist][kRadixBins]
    __shared__ unsigned long long pf[kSelMax];
    __shared__ unsigned long long wtot[kFusedThreads / 32];
    __shared__ unsigned flt[kRadixBins / 4];  // a byte per low-12-bit value of a prefix: the ranks it may belong to
    static_assert(kSelMax == 8, "a byte of rank bits; copies of the first histogram");
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    int nstamp = 0;
    auto stamp = [&] {  // CTA 0's clock at the end of a phase: where the time goes, for the trace
        if (blockIdx.x == 0 && t == 0 && nstamp < 24) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            st->stamp[nstamp++] = now;
        }
    };
    stamp();
    size_t head = ((reinterpret_cast<uintptr_t>(x) & 15) ? 1 : 0);
    if (head > n) head = n;
    const size_t tid = (size_t)blockIdx.x * kFusedThreads + t, nthreads = (size_t)gridDim.x * kFusedThreads;
    const longlong2* x2 = reinterpret_cast<const longlong2*>(x + head);
    const size_t npairs = (n - head) / 2;
    auto walk = [&](auto one) {
        if (tid == 0 && head) one(x[0]);
        size_t i = tid;
        // (the data sits in L2 at best: what a pass costs is load latency, so four loads in flight per thread)
        for (; i + 3 * nthreads < npairs; i += 4 * nthreads) {
            const longlong2 a = __ldg(x2 + i), b = __ldg(x2 + i + nthreads), c = __ldg(x2 + i + 2 * nthreads),
                            d = __ldg(x2 + i + 3 * nthreads);
            one(a.x); one(a.y); one(b.x); one(b.y); one(c.x); one(c.y); one(d.x); one(d.y);
        }
        for (; i < npairs; i += nthreads) {
            const longlong2 a = __ldg(x2 + i);
            one(a.x); one(a.y);
        }
        if (tid == 0 && head + 2 * npairs < n) one(x[n - 1]);
    };
    if (!have_bounds) {  // mom: reset by the host before the launch
        long long mn = 0x7FFFFFFFFFFFFFFFll, mx = (long long)0x8000000000000000ull;
        walk([&](long long v) {
            mn = v < mn ? v : mn;
            mx = v > mx ? v : mx;
        });
#pragma unroll
        for (int d = 16; d; d >>= 1) {
            const long long omn = __shfl_xor_sync(0xFFFFFFFFu, mn, d), omx = __shfl_xor_sync(0xFFFFFFFFu, mx, d);
            mn = omn < mn ? omn : mn;
            mx = omx > mx ? omx : mx;
        }
        if (lane == 0 && mn <= mx) {
            atomicMin(&mom->min_ns, mn);
            atomicMax(&mom->max_ns, mx);
        }
        grid_barrier(bar);
        stamp();
    }
    if (have_bounds != 1) {  // 2: an earlier kernel on the stream (the sampling) left them there
        lo = __ldcg(&mom->min_ns);
        hi = __ldcg(&mom->max_ns);
    }
    const unsigned long long kmin = (unsigned long long)lo ^ 0x8000000000000000ull,
                             kmax = (unsigned long long)hi ^ 0x8000000000000000ull, range = kmax - kmin;
    int top = lo > hi ? 0 : 64 - __clzll((long long)range);
    const int nk = ranks.nk;
    int bad = lo > hi;
    for (int k = 0; k < nk; k++) bad |= ranks.rank[k] >= n;
    if (blockIdx.x == 0 && t < kSelMax) {  // read by others only after the first barrier of a round
        st->prefix[t] = 0;
        st->rank[t] = t < nk ? ranks.rank[t] : 0;
        st->more[t] = 0;
        if (t == 0) st->bad = bad;
    }
    int nhist = 1, round = 0;
    while (!bad && top > 0) {
        const int shift = top > kRadixBits ? top - kRadixBits : 0, width = top - shift;
        const unsigned mask = (1u << width) - 1;
        // Round 0 has ONE histogram, and latencies crowd into a few of its bins (shared-memory atomics on one
        // address serialise): kSelMax copies of it, a copy per four warps.  Later rounds have a histogram per rank
        // that few samples reach: a byte per value of the prefix's low 12 bits says which ranks' prefixes end in
        // them, so a sample that matches none — nearly all — costs one look-up instead of a compare per rank.
        const int ncopies = round == 0 ? kSelMax : nhist;
        for (int i = t; i < ncopies * kRadixBins; i += kFusedThreads) hs[i] = 0;
        for (int i = t; i < kRadixBins / 4; i += kFusedThreads) flt[i] = 0;
        if (t < kSelMax) pf[t] = t >= nk ? ~0ull : round == 0 ? 0ull : __ldcg(&st->prefix[t]);
        __syncthreads();
        if (round > 0 && t < nk) {
            const unsigned slot = (unsigned)pf[t] & (kRadixBins - 1);
            atomicOr(flt + (slot >> 2), (1u << t) << (8 * (slot & 3)));
        }
        __syncthreads();
        if (round == 0) {
            unsigned* const mine = hs + (warp & (kSelMax - 1)) * kRadixBins;  // (a copy per four warps)
            walk([&](long long v) {
                const unsigned long long key = ((unsigned long long)v ^ 0x8000000000000000ull) - kmin;
                const unsigned dgt = (unsigned)(key >> shift) & mask;
                const unsigned long long up = top >= 64 ? 0 : key >> top;
                if (up == 0) atomicAdd(mine + dgt, 1u);  // (a sample outside given bounds has bits above `top`)
            });
        } else {
            const unsigned char* const flt8 = reinterpret_cast<const unsigned char*>(flt);
            walk([&](long long v) {
                const unsigned long long key = ((unsigned long long)v ^ 0x8000000000000000ull) - kmin;
                const unsigned long long up = key >> top;  // (top <= 52 after a round)
                unsigned m = flt8[(unsigned)up & (kRadixBins - 1)];
                while (m) {
                    const int k = __ffs(m) - 1;
                    m &= m - 1;
                    if (up == pf[k]) atomicAdd(hs + k * kRadixBins + ((unsigned)(key >> shift) & mask), 1u);
                }
            });
        }
        __syncthreads();
        if (round == 0) {
            for (int i = t; i < kRadixBins; i += kFusedThreads) {
                unsigned c = 0;
#pragma unroll
                for (int r = 0; r < kSelMax; r++) c += hs[r * kRadixBins + i];
                if (c) atomicAdd(hist + i, (unsigned long long)c);
            }
        } else {
            for (int i = t; i < nhist * kRadixBins; i += kFusedThreads) {
                const unsigned c = hs[i];
                if (c) atomicAdd(hist + i, (unsigned long long)c);
            }
        }
        stamp();
        grid_barrier(bar);
        stamp();
        for (int h = blockIdx.x; h < nhist; h += gridDim.x) {  // CTA h: histogram h (a small grid: several each)
            unsigned long long* const H = hist + (size_t)h * kRadixBins + kScanBins * t;
            unsigned long long c[kScanBins], sum = 0;
#pragma unroll
            for (int j = 0; j < kScanBins; j += 2) {
                const ulonglong2 p = __ldcg(reinterpret_cast<const ulonglong2*>(H + j));
                c[j] = p.x;
                c[j + 1] = p.y;
            }
#pragma unroll
            for (int j = 0; j < kScanBins; j++) sum += c[j];
            unsigned long long incl = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                if (lane >= d) incl += o;
            }
            // after the first round histogram h belongs to rank h alone; its rank and prefix are read by every thread
            // BEFORE the barriers below, because the one thread that finds the bucket overwrites them
            const unsigned long long r_own = round == 0 ? 0ull : __ldcg(&st->rank[h]),
                                     p_own = round == 0 ? 0ull : __ldcg(&st->prefix[h]);
            __syncthreads();  // (wtot: the previous histogram's readers are done)
            if (lane == 31) wtot[warp] = incl;
            __syncthreads();
            unsigned long long before = incl - sum, total = 0;
            for (int w = 0; w < kFusedThreads / 32; w++) {
                if (w < warp) before += wtot[w];
                total += wtot[w];
            }
            for (int k = 0; k < nk; k++) {
                if (nhist != 1 && k != h) continue;
                const unsigned long long r = round == 0 ? ranks.rank[k] : r_own;
                if (r >= total) {
                    if (t == 0) st->bad = 1;
                } else if (r >= before && r < before + sum) {  // exactly one thread
                    unsigned long long cum = before;
                    int j = 0;
#pragma unroll
                    for (int jj = 0; jj < kScanBins - 1; jj++)
                        if (j == jj && r >= cum + c[jj]) { cum += c[jj]; j++; }
                    unsigned long long cj = c[0];
#pragma unroll
                    for (int jj = 1; jj < kScanBins; jj++) cj = j == jj ? c[jj] : cj;
                    st->prefix[k] = (p_own << width) | (unsigned long long)(kScanBins * t + j);
                    st->rank[k] = r - cum;
                    if (cj > approx_max) st->more[round] = 1;
                }
            }
#pragma unroll
            for (int j = 0; j < kScanBins; j += 2) *reinterpret_cast<ulonglong2*>(H + j) = make_ulonglong2(0, 0);
        }
        stamp();
        grid_barrier(bar);
        stamp();
        bad = __ldcg(&st->bad);
        const int more = __ldcg(&st->more[round]);
        top = shift;
        nhist = nk;
        round++;
        if (approx_max && (!more || top <= kApproxBits)) break;
    }
    if (blockIdx.x == 0 && t < nk) {
        unsigned long long key = round == 0 ? 0ull : __ldcg(&st->prefix[t]);
        if (top > 0) {  // (only when approx_max stopped the rounds early: the bucket's ends)
            key <<= top;
            if ((ceil_mask >> t) & 1) key |= (1ull << top) - 1;
        }
        key = key < range ? key : range;
        st->result[t] = (long long)((key + kmin) ^ 0x8000000000000000ull);
    }
    if (blockIdx.x == 0 && t == 0) {
        st->top = 0;
        st->nk = nk;
        st->key_base = kmin;
        st->nstamp = nstamp;
    }
}

// dynamic window: everything up to the end of the counters (absolute address, see the map above), then the tile ring
size_t hist_smem(int nbins, int lut_bytes, int nstages, bool collect) {
    const size_t end = lut_abs(collect) + (size_t)lut_bytes + (size_t)kHistWarps * (nbins + 1) * 32 * sizeof(unsigned);
    return ((end + 127) & ~(size_t)127) + kRingHeader + (size_t)nstages * kTileBytes +
           (collect ? 64 + (size_t)kHistWarps * kKeepCap * 8 : 0);
}
// CTAs per SM and ring stages per CTA that fit the SM's 227 KB (plus 1 KB reserved per CTA): as many CTAs as the
// registers allow with two stages each, then as many more stages as still fit
void hist_shape(size_t smem_without_ring, int by_regs, int* per_sm, int* nstages) {
    const size_t budget = 227 * 1024;
    int c = by_regs, forced = 0;
    if (const char* e = getenv("B200_HIST_SHAPE")) {  // tuning override: "<ctas per SM>,<stages>"
        int cc = 0, ss = 0;
        if (sscanf(e, "%d,%d", &cc, &ss) == 2 && cc >= 1 && ss >= 1 && ss <= kMaxStages) { c = cc; forced = ss; }
    }
    while (c > 1 && c * (smem_without_ring + 2 * (size_t)kTileBytes + 1024) > budget) c--;
    int st = (int)((budget / c - 1024 - smem_without_ring) / kTileBytes);
    if (forced && forced < st) st = forced;
    *per_sm = c;
    *nstages = st < 1 ? 1 : st > kMaxStages ? kMaxStages : st;
}

int sm_count() {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

// One sample per stratum of `stride` consecutive inputs, at a hashed position inside the stratum (a fixed
// position would alias with anything periodic in the data, e.g. 48 workers writing round-robin).
__global__ void __launch_bounds__(256)
sample_kernel(const long long* __restrict__ x, size_t stride, long long* __restrict__ out, size_t s,
              unsigned long long* __restrict__ outside, LatMoments* __restrict__ bounds) {
    long long mn = 0x7FFFFFFFFFFFFFFFll, mx = (long long)0x8000000000000000ull;
    unsigned wide = 0;
    auto where = [&](size_t i) -> const long long* {
        unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        return x + i * stride + z % stride;
    };
    auto take = [&](size_t i, long long v) {
        out[i] = v;
        wide += ((unsigned long long)v >> 32) != 0;  // negative or >= 2^32: not on the histogram kernel's fast path
        mn = v < mn ? v : mn;
        mx = v > mx ? v : mx;
    };
    // every load is a DRAM access of its own with nothing to hide behind: four in flight per thread
    const size_t step = (size_t)gridDim.x * 256;
    size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    for (; i + 3 * step < s; i += 4 * step) {
        const long long a = __ldg(where(i)), b = __ldg(where(i + step)), c = __ldg(where(i + 2 * step)),
                        d = __ldg(where(i + 3 * step));
        take(i, a); take(i + step, b); take(i + 2 * step, c); take(i + 3 * step, d);
    }
    for (; i < s; i += step) take(i, __ldg(where(i)));
    // how much of the sample the 32-bit fast path cannot take (the caller decides whether a histogram pass pays),
    // and the sample's extremes (the bounds of the select that follows): one set of atomics per CTA
    __shared__ long long smn[8], smx[8];
    __shared__ unsigned swide[8];
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        const long long omn = __shfl_xor_sync(0xFFFFFFFFu, mn, d), omx = __shfl_xor_sync(0xFFFFFFFFu, mx, d);
        mn = omn < mn ? omn : mn;
        mx = omx > mx ? omx : mx;
        wide += __shfl_xor_sync(0xFFFFFFFFu, wide, d);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { smn[warp] = mn; smx[warp] = mx; swide[warp] = wide; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) {
            mn = smn[w] < mn ? smn[w] : mn;
            mx = smx[w] > mx ? smx[w] : mx;
            wide += swide[w];
        }
        if (outside && wide) atomicAdd(outside, (unsigned long long)wide);
        if (bounds && mn <= mx) {
            atomicMin(&bounds->min_ns, mn);
            atomicMax(&bounds->max_ns, mx);
        }
    }
}

}  // namespace

cudaError_t launch_sample(const long long* d_samples, size_t stride, long long* d_out, size_t s, cudaStream_t stream,
                          unsigned long long* d_outside, LatMoments* d_bounds, const LatScratch* results) {
    if (results) {  // outside counter and bounds both live in the results block: one copy resets them
        cudaError_t e = lat_reset_results(*results, stream);
        if (e != cudaSuccess) return e;
    } else if (d_outside) {
        cudaError_t e = cudaMemsetAsync(d_outside, 0, sizeof *d_outside, stream);
        if (e != cudaSuccess) return e;
    }
    if (!s) return cudaSuccess;
    const unsigned grid = (unsigned)std::min<size_t>((s + 255) / 256, (size_t)sm_count() * 8);
    sample_kernel<<<grid, 256, 0, stream>>>(d_samples, stride, d_out, s, d_outside, d_bounds);
    return cudaGetLastError();
}

// Where a kernel's dynamic shared memory starts when it has no static shared memory: the histogram and CRC row
// kernels lay their tables out at absolute shared addresses and need it to be at most 1 KB (the driver's
// reserved block).  Asked once per device at set-up, so that a driver that reserves more fails b200_init
// with a message instead of trapping inside a kernel later.
__global__ void smem_base_probe_kernel(unsigned long long* out) {
    extern __shared__ __align__(1024) unsigned char probe_smem[];
    *out = (unsigned long long)__cvta_generic_to_shared(probe_smem);
}

cudaError_t smem_base_probe(unsigned long long* d_tmp) {
    unsigned long long base = ~0ull;
    smem_base_probe_kernel<<<1, 1, 1024>>>(d_tmp);
    const cudaError_t e = cudaMemcpy(&base, d_tmp, sizeof base, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return e;
    return base > 1024 ? cudaErrorInvalidConfiguration : cudaSuccess;  // see smem_base_probe_kernel
}

cudaError_t lat_scratch_init(LatScratch* s) {
    memset(s, 0, sizeof *s);
    cudaError_t e = cudaMalloc(&s->d_results, sizeof(LatResults));
    if (e != cudaSuccess) return e;
    s->d_counts = s->d_results->counts;  // (addresses only: nothing is read through the device pointer here)
    s->d_mom = &s->d_results->mom;
    s->d_compact_n = &s->d_results->compact_n;
    s->d_sel = &s->d_results->sel;
    e = smem_base_probe(s->d_compact_n);
    if (e != cudaSuccess) return e;
    e = cudaMalloc(&s->d_results_init, sizeof(LatResults));
    if (e == cudaSuccess) e = cudaHostAlloc(&s->h_results, sizeof(LatResults), cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaHostAlloc(&s->h_params, sizeof(HistParams), cudaHostAllocDefault);
    if (e == cudaSuccess) {
        memset(s->h_results, 0, sizeof(LatResults));
        s->h_results->mom.min_ns = 0x7FFFFFFFFFFFFFFFll;
        s->h_results->mom.max_ns = (long long)0x8000000000000000ull;
        e = cudaMemcpy(s->d_results_init, s->h_results, sizeof(LatResults), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMalloc(&s->d_params, sizeof(HistParams));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_radix, (size_t)kMaxSelect * kRadixBins * sizeof(unsigned long long));
    if (e == cudaSuccess)
        e = cudaHostAlloc(&s->h_radix, (size_t)kMaxSelect * kRadixBins * sizeof(unsigned long long), cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaMalloc(&s->d_prefix, 2 * kMaxSelect * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_sel_in, sizeof(SelDeferred));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_sel_bar, 2 * sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMemset(s->d_sel_bar, 0, 2 * sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_sel_hist, (size_t)kSelMax * kRadixBins * sizeof(unsigned long long));
    if (e == cudaSuccess) {
        s->sel_hist_clean = new (std::nothrow) bool(false);
        if (!s->sel_hist_clean) e = cudaErrorMemoryAllocation;
    }
    return e;
}
cudaError_t lat_reset_results(const LatScratch& s, cudaStream_t stream) {
    return cudaMemcpyAsync(s.d_results, s.d_results_init, sizeof(LatResults), cudaMemcpyDeviceToDevice, stream);
}
cudaError_t lat_read_results(const LatScratch& s, cudaStream_t stream) {
    return cudaMemcpyAsync(s.h_results, s.d_results, sizeof(LatResults), cudaMemcpyDeviceToHost, stream);
}
void lat_scratch_free(LatScratch* s) {
    cudaFree(s->d_params);
    cudaFree(s->d_results);
    cudaFree(s->d_results_init);
    cudaFreeHost(s->h_results);
    cudaFreeHost(s->h_params);
    cudaFree(s->d_radix);
    cudaFreeHost(s->h_radix);
    cudaFree(s->d_prefix);
    cudaFree(s->d_compact);
    cudaFree(s->d_sel_in);
    cudaFree(s->d_sel_bar);
    cudaFree(s->d_sel_hist);
    delete s->sel_hist_clean;
    cudaFree(s->d_sample);
    cudaFree(s->d_collect);
    memset(s, 0, sizeof *s);
}

static cudaError_t reset_moments(const LatScratch& s, cudaStream_t stream) {
    return cudaMemcpyAsync(s.d_mom, &s.d_results_init->mom, sizeof(LatMoments), cudaMemcpyDeviceToDevice, stream);
}

cudaError_t launch_hist(const long long* d_samples, size_t n, const LatScratch& s, int want_moments,
                        cudaStream_t stream, const HistParams& P, const HistCollect* collect) {
    // (the pinned staging is free again when the stream has been waited for, which every caller does before its
    // next launch; the device's lock is held throughout)
    memcpy(s.h_params, &P, sizeof P);
    cudaError_t e = cudaMemcpyAsync(s.d_params, s.h_params, sizeof P, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
    e = lat_reset_results(s, stream);  // counts, moments and the collect counter (s.d_compact_n)
    if (e != cudaSuccess) return e;
    if (collect && collect->count != s.d_compact_n) {
        e = cudaMemsetAsync(collect->count, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    if (!n) return cudaSuccess;
    using Kernel = void (*)(const long long*, size_t, const HistParams*, unsigned long long*, LatMoments*, HistCollect, int);
    // without moments the quotient is only formed as the table index, so the shift-indexed modes share an
    // instantiation — but for mode 3, whose samples are taken shifted by pre32 (longer reach)
    static const Kernel by_shift[2][4] = {
        {hist_kernel<false, 0, false, false>, hist_kernel<false, 0, false, false>, hist_kernel<false, 0, false, false>,
         hist_kernel<false, 3, false, false>},
        {hist_kernel<true, 0, false, false>, hist_kernel<true, 1, false, false>, hist_kernel<true, 2, false, false>,
         hist_kernel<true, 3, false, false>}};
    static const Kernel collecting[2][4] = {  // (a collecting pass is never indexed by quotient: its cells are parked)
        {hist_kernel<false, 0, false, true>, hist_kernel<false, 0, false, true>, hist_kernel<false, 0, false, true>,
         hist_kernel<false, 3, false, true>},
        {hist_kernel<true, 0, false, true>, hist_kernel<true, 1, false, true>, hist_kernel<true, 2, false, true>,
         hist_kernel<true, 3, false, true>}};
    static const Kernel collecting_no_squares[4] = {hist_kernel<2, 0, false, true>, hist_kernel<2, 1, false, true>,
                                                    hist_kernel<2, 2, false, true>, hist_kernel<2, 3, false, true>};
    static const Kernel by_q[2] = {hist_kernel<false, 3, true, false>, hist_kernel<true, 3, true, false>};
    if (collect && P.by_q) return cudaErrorInvalidValue;  // the builder never pairs the two
    const Kernel k = collect ? (want_moments == 2 ? collecting_no_squares[P.mode32] : collecting[want_moments ? 1 : 0][P.mode32])
                             : P.by_q ? by_q[want_moments ? 1 : 0] : by_shift[want_moments ? 1 : 0][P.mode32];
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    // resident CTAs per SM: kHistCtasPerSm by registers, fewer when the private counter tables are large
    const int nbins = P.nthr + 1;
    int per_sm, nstages;
    hist_shape(hist_smem(nbins, P.lut_bytes, 0, collect != nullptr), hist_ctas_by_regs(want_moments != 0, P.by_q != 0), &per_sm,
               &nstages);
    const size_t smem = hist_smem(nbins, P.lut_bytes, nstages, collect != nullptr);
    const int max_grid = sm_count() * per_sm;
    const size_t want_ctas = (n / 2 + kHistThreads - 1) / kHistThreads;
    const int grid = (int)std::min<size_t>((size_t)max_grid, want_ctas ? want_ctas : 1);
    k<<<grid, kHistThreads, smem, stream>>>(d_samples, n, s.d_params, s.d_counts, s.d_mom,
                                            collect ? *collect : HistCollect{nullptr, nullptr, 0}, nstages);
    return cudaGetLastError();
}

cudaError_t launch_minmax(const long long* d_samples, size_t n, const LatScratch& s, cudaStream_t stream) {
    cudaError_t e = reset_moments(s, stream);
    if (e != cudaSuccess || !n) return e;
    int grid = (int)std::min<size_t>((size_t)sm_count() * 8, (n + kMinmaxThreads - 1) / kMinmaxThreads);
    minmax_kernel<<<grid, kMinmaxThreads, 0, stream>>>(d_samples, n, s.d_mom);
    return cudaGetLastError();
}

cudaError_t launch_select_resident(const long long* d_samples, size_t n, const SelRanks& ranks, bool have_bounds,
                                   long long lo, long long hi, const LatScratch& s, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(s.d_sel_hist, 0, (size_t)kSelMax * kRadixBins * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    if (!have_bounds) {
        e = launch_minmax(d_samples, n, s, stream);
        if (e != cudaSuccess) return e;
    }
    sel_init_kernel<<<1, 32, 0, stream>>>(s.d_sel, s.d_mom, ranks, have_bounds ? 1 : 0, lo, hi, (unsigned long long)n);
    const size_t smem = (size_t)kSelMax * kRadixBins * sizeof(unsigned);
    e = cudaFuncSetAttribute(sel_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const size_t want_ctas = (n / 4 + kRadixThreads - 1) / kRadixThreads;
    const int grid = (int)std::min<size_t>((size_t)sm_count(), want_ctas ? want_ctas : 1);
    // with bounds the number of passes is known; without, the kernels past the last needed pass return at once
    int rounds = kSelRounds;
    if (have_bounds && lo <= hi) {
        const unsigned long long range = ((unsigned long long)hi ^ 0x8000000000000000ull) - ((unsigned long long)lo ^ 0x8000000000000000ull);
        int bits = 0;
        while (bits < 64 && (range >> bits)) bits++;
        rounds = (bits + kRadixBits - 1) / kRadixBits;
    }
    for (int r = 0; r < rounds; r++) {
        sel_pass_kernel<<<grid, kRadixThreads, smem, stream>>>(d_samples, n, s.d_sel, s.d_sel_hist);
        sel_scan_kernel<<<1, 1024, 0, stream>>>(s.d_sel, s.d_sel_hist);
    }
    return cudaGetLastError();
}

namespace {

struct WantedRanks {
    unsigned long long k[kSelMax];
};
__global__ void bracket_ranks_kernel(const unsigned long long* __restrict__ counts, int nb, WantedRanks want, int nk,
                                     unsigned long long cap, SelDeferred* __restrict__ out) {
    if (threadIdx.x || blockIdx.x) return;
    unsigned long long inside = 0, miss = 0;
    for (int j = 0; j < nb; j++) inside += counts[2 * j + 1];
    for (int i = 0; i < nk; i++) {
        unsigned long long below = 0, kept_before = 0, rank = 0;
        int j = 0;
        for (;; j++) {
            if (j == nb) { miss = 1; break; }            // beyond the last bracket
            below += counts[2 * j];
            if (want.k[i] < below) { miss = 1; break; }  // between brackets
            if (want.k[i] < below + counts[2 * j + 1]) { rank = kept_before + (want.k[i] - below); break; }
            below += counts[2 * j + 1];
            kept_before += counts[2 * j + 1];
        }
        out->ranks.rank[i] = rank;
    }
    for (int i = nk; i < kSelMax; i++) out->ranks.rank[i] = 0;
    out->ranks.nk = nk;
    if (inside == 0 || inside > cap) miss = 1;
    out->n = inside;
    out->miss = miss;
}

}  // namespace

cudaError_t launch_bracket_ranks(const LatScratch& s, int nbrackets, const unsigned long long* k, int nk,
                                 unsigned long long cap, cudaStream_t stream) {
    WantedRanks w{};
    for (int i = 0; i < nk && i < kSelMax; i++) w.k[i] = k[i];
    bracket_ranks_kernel<<<1, 32, 0, stream>>>(s.d_counts, nbrackets, w, nk, cap, s.d_sel_in);
    return cudaGetLastError();
}

// 1 when this device can run the one-kernel select (cooperative launch, a CTA per SM with the histograms of
// kSelMax ranks in shared memory), 0 when not or when B200_SELECT_FUSED=0 asks for the separate launches
int select_fused_grid() {
    static const int grid = [] {
        const char* e = getenv("B200_SELECT_FUSED");
        if (e && *e == '0') return 0;
        int dev = 0, coop = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (!coop) return 0;
        const size_t smem = (size_t)kSelMax * kRadixBins * sizeof(unsigned);
        if (cudaFuncSetAttribute(sel_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sel_fused_kernel, kFusedThreads, smem) != cudaSuccess ||
            per_sm < 1) {
            cudaGetLastError();
            return 0;
        }
        return sm_count();  // (devices of one process are alike: B200s)
    }();
    return grid;
}

cudaError_t launch_select_fused(const long long* d_samples, size_t n, const SelRanks& ranks, int have_bounds,
                                long long lo, long long hi, unsigned long long approx_max, unsigned ceil_mask,
                                const LatScratch& s, cudaStream_t stream, bool deferred) {
    // the kernel leaves the histograms cleared: they are only cleared here after a run that did not end well
    // (*s.sel_hist_clean is set by the caller once it has seen the run's result)
    cudaError_t e = cudaSuccess;
    if (!*s.sel_hist_clean) {
        e = cudaMemsetAsync(s.d_sel_hist, 0, (size_t)kSelMax * kRadixBins * sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    *s.sel_hist_clean = false;
    if (!have_bounds) {
        e = reset_moments(s, stream);
        if (e != cudaSuccess) return e;
    }
    const size_t want_ctas = (n / 4 + kFusedThreads - 1) / kFusedThreads;
    const int grid = (int)std::min<size_t>((size_t)select_fused_grid(), want_ctas ? want_ctas : 1);
    const size_t smem = (size_t)kSelMax * kRadixBins * sizeof(unsigned);
    SelState* st = s.d_sel;
    unsigned long long* hist = s.d_sel_hist;
    LatMoments* mom = s.d_mom;
    SelRanks r = ranks;
    int hb = have_bounds;
    const SelDeferred* in = deferred ? s.d_sel_in : nullptr;
    unsigned* bar = s.d_sel_bar;
    void* args[] = {(void*)&d_samples, (void*)&n, (void*)&st, (void*)&hist, (void*)&mom, (void*)&r, (void*)&hb,
                    (void*)&lo, (void*)&hi, (void*)&approx_max, (void*)&ceil_mask, (void*)&in, (void*)&bar};
    return cudaLaunchCooperativeKernel((const void*)sel_fused_kernel, dim3(grid), dim3(kFusedThreads), args, smem, stream);
}

cudaError_t launch_compact(const long long* d_samples, size_t n, unsigned long long key_base, int top, int nprefix,
                           const LatScratch& s, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(s.d_compact_n, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess || !n) return e;
    int grid = (int)std::min<size_t>((size_t)sm_count() * 4, (n + kRadixThreads - 1) / kRadixThreads);
    compact_kernel<<<grid, kRadixThreads, 0, stream>>>(d_samples, n, key_base, top, nprefix, s.d_prefix, s.d_compact,
                                                      s.d_compact_n, s.compact_cap);
    return cudaGetLastError();
}

cudaError_t launch_radix_pass(const long long* d_samples, size_t n, unsigned long long key_base, int shift, int top,
                              int nk, const LatScratch& s, cudaStream_t stream, bool own_bases) {
    cudaError_t e = cudaMemsetAsync(s.d_radix, 0, (size_t)kMaxSelect * kRadixBins * sizeof(unsigned long long), stream);
    if (e != cudaSuccess || !n) return e;
    size_t smem = (size_t)nk * kRadixBins * sizeof(unsigned);
    e = cudaFuncSetAttribute(radix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)(kMaxSelect * kRadixBins * sizeof(unsigned)));
    if (e != cudaSuccess) return e;
    size_t want_ctas = (n / 2 + kRadixThreads - 1) / kRadixThreads;
    int grid = (int)std::min<size_t>((size_t)sm_count() * 2, want_ctas ? want_ctas : 1);
    radix_kernel<<<grid, kRadixThreads, smem, stream>>>(d_samples, n, key_base, shift, top, nk, s.d_prefix,
                                                        own_bases ? s.d_prefix + kMaxSelect : nullptr, s.d_radix);
    return cudaGetLastError();
}

// ---- host: float edges -> int64 nanosecond thresholds --------------------------------------------------

namespace {

// smallest raw n with trunc(n / d) >= E   (E integer, d >= 1), saturating
long long first_n_with_quotient_at_least(double E, long long d) {
    const double lim = 9.2e18;
    if (E > 0) {
        double t = E * (double)d;
        if (t >= lim) return 0x7FFFFFFFFFFFFFFFll;
        return (long long)E * d;
    }
    double t = (E - 1) * (double)d;
    if (t <= -lim) return (long long)0x8000000000000000ull;
    return ((long long)E - 1) * d + 1;
}

void finish_hist_params(HistParams* P, long long divisor, bool collect);

}  // namespace

int host_edges_to_thresholds(const double* edges, int nedges, int mode, long long divisor, long long* thr) {
    if (divisor < 1 || nedges < 1 || nedges > kMaxThresholds - 2) return -1;
    for (int i = 0; i < nedges; i++) {
        if (std::isnan(edges[i])) return -1;
        if (i && !(edges[i] > edges[i - 1])) return -1;
    }
    // v = trunc(n/d) is an integer, so  v < e  <=>  v < ceil(e)  and  v <= e  <=>  v < floor(e) + 1
    int m = 0;
    if (mode == 0) {  // numpy: below edges[0] -> count 0 (dropped); last edge inclusive; above -> count nedges
        if (nedges < 2) return -1;
        for (int j = 0; j < nedges - 1; j++) thr[m++] = first_n_with_quotient_at_least(std::ceil(edges[j]), divisor);
        thr[m++] = first_n_with_quotient_at_least(std::floor(edges[nedges - 1]) + 1, divisor);
    } else {  // opencensus: bucket = #{bounds <= v}
        for (int j = 0; j < nedges; j++) thr[m++] = first_n_with_quotient_at_least(std::ceil(edges[j]), divisor);
    }
    return m;
}

int host_build_hist_params(const double* edges, int nedges, int mode, long long divisor, HistParams* P) {
    memset(P, 0, sizeof *P);
    const int m = host_edges_to_thresholds(edges, nedges, mode, divisor, P->thr);
    if (m < 0) return -1;
    P->nthr = m;
    finish_hist_params(P, divisor, false);
    return 0;
}

int host_build_hist_params_thr(const long long* thr, int nthr, long long divisor, bool collect, HistParams* P) {
    memset(P, 0, sizeof *P);
    if (nthr < 1 || nthr > kMaxThresholds || divisor < 1) return -1;
    for (int j = 0; j < nthr; j++) {
        if (j && !(thr[j] > thr[j - 1])) return -1;
        P->thr[j] = thr[j];
    }
    P->nthr = nthr;
    finish_hist_params(P, divisor, collect);
    return 0;
}

namespace {

// thresholds (P->thr, P->nthr) -> look-up table and division magics
void finish_hist_params(HistParams* P, long long divisor, bool collect) {
    const int m = P->nthr;
    // 32-bit dividends (Granlund-Montgomery, N = 32): l = ceil(log2 d), m' = floor(2^32 (2^l - d) / d) + 1,
    // q = (t + ((n - t) >> 1)) >> (l - 1) with t = umulhi(m', n)
    // (the search below is a few hundred 128-bit divisions: what it finds for a divisor is remembered)
    struct Magic32 {
        long long divisor = 0;
        unsigned magic32 = 0;
        int shift32 = 0, pre32 = 0, mode32 = 0, reach_bits = 32;
    };
    static thread_local Magic32 last;
    if (last.divisor != divisor) {
        Magic32 g;
        g.divisor = divisor;
        if (divisor == 1) {
            g.mode32 = 0;
        } else if ((unsigned long long)divisor > 0xFFFFFFFFull) {
            g.mode32 = 2;
        } else {
            unsigned long long d32 = (unsigned long long)divisor;
            int l = 0;
            while ((1ull << l) < d32) l++;
            g.magic32 = (unsigned int)((((unsigned __int128)1 << 32) * ((1ull << l) - d32)) / d32 + 1);
            g.shift32 = l - 1;
            g.mode32 = 1;
            // Cheaper when it exists: take the divisor's trailing zero bits out of divisor and dividend alike (d = 2^pre
            // d', u = n >> pre, trunc(n / d) = trunc(u / d')); a 32-bit magic M = ceil(2^k / d') without fix-up is exact
            // for all u < 2^B as soon as (M d' - 2^k) (2^B - 1) < 2^k.  The largest such B <= 32 is taken: the kernel's
            // fast path then reaches n < 2^(B + pre) (2^37 ns = 137 s for milliseconds, 2^34 ns = 17 s for
            // microseconds) instead of 2^32.  d' >= 16 also bounds the quotient below 2^28, which lets the kernel sum
            // four quotients in 32 bits and four squares in 64.
            int pre = 0;
            while (!((d32 >> pre) & 1)) pre++;
            const unsigned long long dp = d32 >> pre;
            if (dp >= 16) {
                for (int B = 32; B >= 32 - pre && g.mode32 != 3; B--) {
                    const unsigned __int128 nmax = ((unsigned __int128)1 << B) - 1;
                    for (int post = 0; post < 32; post++) {
                        const unsigned __int128 two_k = (unsigned __int128)1 << (32 + post);
                        const unsigned __int128 M = (two_k + dp - 1) / dp;
                        if (M >> 32) break;
                        if ((M * dp - two_k) * nmax < two_k) {
                            g.magic32 = (unsigned int)M;
                            g.shift32 = post;
                            g.pre32 = pre;
                            g.mode32 = 3;
                            g.reach_bits = B + pre;
                            break;
                        }
                    }
                }
            }
        }
        if (g.mode32 != 3) g.reach_bits = 32;
        last = g;
    }
    P->magic32 = last.magic32;
    P->shift32 = last.shift32;
    P->pre32 = last.pre32;
    P->mode32 = last.mode32;
    P->reach_bits = last.reach_bits;
    // The look-up table (see hist_kernel).  Cell c covers the samples [lo_c, hi_c]; its entry is 1 + the number of
    // thresholds <= lo_c — the counter row of every sample in the cell — unless a threshold lies inside (lo_c, hi_c]:
    // then 0, the sample is parked.  (A collecting kernel also parks the cells of kept bins, which are the even rows:
    // it tells them by the row's parity, and having the row it counts them on its fast path.)
    // Filled threshold by threshold — runs of equal entries between the cells that hold one — not cell by cell: this
    // runs on the host before every launch.
    const unsigned long long reach = 1ull << P->reach_bits;  // the fast path takes 0 <= n < reach
    // cell_of(t) for 0 < t < reach: the cell t falls into and whether t is that cell's lowest sample
    auto fill = [&](auto cell_of, auto cell_lo) {
        int parked = 0, below = 0, pos = 0;  // cells [0, pos) are done
        while (below < m && P->thr[below] <= 0) below++;
        for (int j = below; j < m && (unsigned long long)P->thr[j] < reach; j++) {
            bool exact;
            const int ct = cell_of((unsigned long long)P->thr[j], &exact);
            if (ct >= pos) {
                memset(P->lut + pos, below + 1, (size_t)(ct - pos));
                pos = ct;
                if (!exact) {
                    P->lut[ct] = 0;
                    parked++;
                    pos = ct + 1;
                }
            }  // (else: one more threshold in a cell that is parked already)
            below++;
        }
        if (pos < kLutSize) memset(P->lut + pos, below + 1, (size_t)(kLutSize - pos));
        for (int c = kLutSize - 1; c >= 0 && cell_lo(c) >= reach; c--) P->lut[c] = 0;  // beyond the reach: never looked up
        return parked;
    };
    const unsigned long long d32 = (unsigned long long)divisor;
    P->by_q = 0;
    P->lut_bytes = kLutSize;
    if (P->mode32 == 3 && !collect) {
        // index by quotient (clamped to the last cell): no cell is parked when every threshold in reach is a multiple
        // of the divisor and lies inside the table; thresholds beyond it (the OpenCensus bounds from 10 s up, in
        // milliseconds) park the last cell only, i.e. the samples of 8.19 s and more — still the better table.  (A
        // collecting pass is never indexed by quotient: its brackets are far narrower than a quotient.)
        const int parked = fill(
            [&](unsigned long long t, bool* exact) {
                const unsigned long long c = t / d32;
                *exact = c < (unsigned long long)kLutSize && t % d32 == 0;
                return (int)std::min<unsigned long long>(c, kLutSize - 1);
            },
            [&](int c) { return (unsigned long long)c * d32; });
        P->by_q = parked <= 1;
        if (P->by_q) P->lut_bytes = (int)std::min<unsigned long long>(kLutSize, ((reach - 1) / d32 + 1 + 127) & ~127ull);
    }
    if (!P->by_q) {
        // index by shift: the smallest shift (not below pre32: the kernel shifts the pre-shifted sample) that brings
        // the largest threshold in reach inside the table — finer cells park fewer samples; at most reach_bits - 13,
        // where the table spans the whole reach
        long long top = 0;
        for (int j = 0; j < m; j++)
            if (P->thr[j] >= 0 && (unsigned long long)P->thr[j] < reach) top = std::max(top, P->thr[j]);
        int sh = P->mode32 == 3 ? P->pre32 : 0;
        while (sh < P->reach_bits - 13 && ((unsigned long long)top >> sh) >= (unsigned long long)kLutSize - 1) sh++;
        P->lut_shift = sh;
        fill(
            [&](unsigned long long t, bool* exact) {
                const unsigned long long c = t >> sh;
                *exact = c < (unsigned long long)kLutSize && (t & ((1ull << sh) - 1)) == 0;
                return (int)std::min<unsigned long long>(c, kLutSize - 1);
            },
            [&](int c) { return (unsigned long long)c << sh; });
    }
    // division magic for unsigned 64-bit dividends (round-up method with optional add fix-up)
    P->divisor = divisor;
    unsigned long long d = (unsigned long long)divisor;
    if ((d & (d - 1)) == 0) {
        P->magic = 0;
        int l = 0;
        while ((1ull << l) < d) l++;
        P->magic_shift = l;
    } else {
        int l = 63;
        while (!((d >> l) & 1)) l--;  // floor(log2 d)
        unsigned __int128 num = (unsigned __int128)1 << (64 + l);
        unsigned long long q = (unsigned long long)(num / d);
        unsigned long long rem = (unsigned long long)(num % d);
        unsigned long long e = d - rem;
        if (e < (1ull << l)) {
            P->magic = q + 1;
            P->magic_add = 0;
            P->magic_shift = l;
        } else {
            unsigned long long q2 = q * 2, rem2 = rem * 2;
            if (rem2 >= d || rem2 < rem) q2 += 1;
            P->magic = q2 + 1;
            P->magic_add = 1;
            P->magic_shift = l;
        }
    }
}

}  // namespace

}  // namespace b200
