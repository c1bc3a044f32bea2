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
// per-sample division or float compare is needed for binning.  A log-linear LUT (32 cells per power of
// two) gives the count at the low end of the sample's cell plus the 0..few thresholds inside it.
// All kernels are HBM-bound streams of 8 bytes per sample with 128-bit loads; counters are private per
// (warp, lane) in shared memory so the skewed latency distribution causes no atomic contention.
#include "lat_kernels.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

namespace {

constexpr int kHistThreads = 256;
constexpr int kHistWarps = kHistThreads / 32;
#ifndef B200_HIST_CTAS
#define B200_HIST_CTAS 3
#endif
constexpr int kHistCtasPerSm = B200_HIST_CTAS;  // 3 x 256 threads: 85 registers each, room for the prefetched tile
constexpr int kRadixThreads = 512;
constexpr int kTrashRows = 2;   // counter rows past the real bins: the EVEN one of them is where the updates of
                                // parked samples land (odd rows are the ones a collecting kernel keeps)

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

// Generic path: any int64 sample.  Out of line, so that the fast loop keeps its registers and instruction cache.
// s == nullptr: count (and collect) only, no moments.
__device__ __noinline__ void slow_sample(long long v, const SlowParams* __restrict__ P, unsigned* mine,
                                         SlowMoments* s, const HistCollect* col) {
    int lo = 0, hi = P->nthr;  // c = #{j : thr[j] <= v}: thr[j] <= v for j < lo, thr[j] > v for j >= hi
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (P->thr[mid] <= v) lo = mid + 1; else hi = mid;
    }
    const int c = lo;
    mine[c * 32]++;
    if (col) {  // odd bins are kept; one global atomic for all the lanes that are here together
        const bool hit = c & 1;
        const unsigned peers = __ballot_sync(__activemask(), hit);
        if (hit) {
            const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
            unsigned long long at = 0;
            if (lane == leader) at = atomicAdd(col->count, (unsigned long long)__popc(peers));
            at = __shfl_sync(peers, at, leader) + __popc(peers & ((1u << lane) - 1));
            if (at < col->cap) col->buf[at] = v;
        }
    }
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

// Shared memory map of the histogram kernels, in ABSOLUTE shared addresses so that the hot loop addresses the
// cell table and the counters as register + immediate (the compiler's own shared pointers carry a window base
// it keeps recomputing under register pressure).  The driver's reserved 1 KB sits below the dynamic window;
// lat_scratch_init checks the window start once per device (smem_base_probe).
constexpr unsigned kSlowAbs = 0x400;                                    // SlowParams
constexpr int kParkCap = 128;                                           // parked samples per warp
constexpr int kParkDrain = 64;                                           // drain from this many on
constexpr unsigned kParkNAbs = kSlowAbs + 640;                          // unsigned[kHistWarps]
constexpr unsigned kParkQAbs = kParkNAbs + 64;                          // long long[kHistWarps][kParkCap]
constexpr unsigned kFastAbs = kParkQAbs + kHistWarps * kParkCap * 8;    // uint2 [kFastCells] (+ unsigned[kFastCells])
// then unsigned[kHistWarps][rows][32] counters, then (collecting kernels) long long[kLaneQueue][kHistThreads]
constexpr int kLaneQueue = 16;                                          // collected samples a lane can hold
constexpr unsigned kLaneQueueStride = kHistThreads * 8;                 // bytes between a lane's slots
constexpr unsigned kQLutAbs = kFastAbs;                                 // quotient kernel: unsigned char[kQLutSize]
static_assert(sizeof(SlowParams) <= 640, "SlowParams slot");
constexpr unsigned kQCntAbs = kQLutAbs + kQLutSize;                     //   then its counters

extern __shared__ __align__(1024) unsigned char hist_smem_raw[];

// ---- tile loop: a ring of tiles in shared memory, filled by 1-D bulk copies (TMA) ---------------------------
// A CTA's tiles are contiguous 16 KB pieces of the sample array, so the operands of a bulk copy are uniform and
// one thread can issue it.  Bandwidth here is (bytes in flight) / (loaded HBM latency, ~1.4 us): with the next
// tile prefetched into registers a thread has 64 bytes in flight, 48 KB per SM, which is where the register
// version stopped (4.85 TB/s; ncu: warps waiting on their loads, DRAM 54 %, issue 42 %).  The ring has nstages
// tiles per CTA in flight and costs no registers: every round the threads copy their samples of the current
// stage into registers, the LAST warp to have done so (a shared-memory counter per stage says who) re-arms the
// stage's barrier and issues the copy of the tile nstages rounds on, and only then are the samples counted.
// (An L2 bulk prefetch running ahead of register loads was measured too: slower than no prefetch at all.)
constexpr int kTileBytes = 4 * kHistThreads * 16;  // 16 KB: four 128-bit pairs per thread
constexpr int kMaxStages = 4;
constexpr int kRingHeader = 128;                   // full barriers (8 B each), then the per-stage counters (4 B each)

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
    longlong2 v;
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr) : "memory");
    return v;
}

// Walks x[0, n): whole tiles of 4 x kHistThreads 128-bit pairs go to eight(a, b, c, d) — tile t belongs to CTA
// t mod gridDim.x — everything else (unaligned head, remainder, odd tail) goes to one(v).  after_tile() runs once
// per tile with the whole warp converged.  ring_abs: absolute shared address (128-byte aligned) of kRingHeader
// bytes followed by nstages tile buffers.  Every thread of the CTA must call this (it synchronises the CTA once).
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
    constexpr size_t kTilePairs = 4 * kHistThreads;
    const size_t ntiles = npairs / kTilePairs;
    const unsigned mine = (size_t)blockIdx.x < ntiles ? (unsigned)((ntiles - 1 - blockIdx.x) / gridDim.x) + 1 : 0;
    const unsigned full0 = ring_abs, cnt0 = ring_abs + 8 * kMaxStages, buf0 = ring_abs + kRingHeader;
    const int lane = threadIdx.x & 31;
    const unsigned char* const tile0 = reinterpret_cast<const unsigned char*>(x2) + (size_t)blockIdx.x * kTileBytes;
    const size_t tile_step = (size_t)gridDim.x * kTileBytes;
    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; s++) {
            mbar_init(full0 + 8 * s, 1);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(cnt0 + 4 * s), "r"(0u) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (unsigned j = 0; j < (unsigned)nstages && j < mine; j++) {
            mbar_expect_tx(full0 + 8 * j, kTileBytes);
            bulk_load(buf0 + j * kTileBytes, tile0 + (size_t)j * tile_step, kTileBytes, full0 + 8 * j);
        }
    }
    __syncthreads();
    unsigned s = 0, ph = 0;  // tile j sits in stage j mod nstages, in that stage's (j div nstages)-th use
    for (unsigned j = 0; j < mine; j++) {
        mbar_wait(full0 + 8 * s, ph);
        const unsigned at = buf0 + s * kTileBytes + threadIdx.x * 16;
        const longlong2 a = lds_pair(at), b = lds_pair(at + 4096), c = lds_pair(at + 8192), d = lds_pair(at + 12288);
        __syncwarp();
        if (lane == 0) {
            // the operands tie this atomic to the four loads above: it is issued once their data has arrived, so
            // when the last warp's count shows up every thread of the CTA has its samples out of the stage
            unsigned seen;
            asm volatile("atom.shared.add.u32 %0, [%1], %2;"
                         : "=r"(seen) : "r"(cnt0 + 4 * s), "r"(1u), "l"(a.x), "l"(b.x), "l"(c.x), "l"(d.x) : "memory");
            if (seen == kHistWarps - 1) {
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(cnt0 + 4 * s), "r"(0u) : "memory");
                if (j + nstages < mine) {
                    mbar_expect_tx(full0 + 8 * s, kTileBytes);
                    bulk_load(buf0 + s * kTileBytes, tile0 + (size_t)(j + nstages) * tile_step, kTileBytes, full0 + 8 * s);
                }
            }
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
            for (int l = 0; l < 32; l++) t += tbl[((size_t)w * rows + b) * 32 + ((l + b) & 31)];
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
// Returns false when the queue is full: the caller then takes the generic path on the spot.
__device__ __forceinline__ bool park1(unsigned char* sm0, int warp, long long v) {
    unsigned* const n = reinterpret_cast<unsigned*>(sm0 + kParkNAbs) + warp;
    const unsigned at = atomicAdd(n, 1u);
    if (at >= (unsigned)kParkCap) return false;
    reinterpret_cast<long long*>(sm0 + kParkQAbs)[(size_t)warp * kParkCap + at] = v;
    return true;
}
// whole warp, converged: hand the parked samples to slow(v), 32 at a time
template <typename Slow>
__device__ __forceinline__ void drain_parked(unsigned char* sm0, int warp, int lane, Slow slow) {
    unsigned* const n = reinterpret_cast<unsigned*>(sm0 + kParkNAbs) + warp;
    const long long* q = reinterpret_cast<const long long*>(sm0 + kParkQAbs) + (size_t)warp * kParkCap;
    __syncwarp();
    const unsigned cnt = min(*n, (unsigned)kParkCap);
    for (unsigned k = lane; k < cnt; k += 32) slow(q[k]);
    __syncwarp();
    if (lane == 0) *n = 0;
    __syncwarp();
}
// after a tile round, warp converged: drain when the queue is filling up (every lane reads the same counter)
template <typename Slow>
__device__ __forceinline__ void maybe_drain(unsigned char* sm0, int warp, int lane, Slow slow) {
    __syncwarp();
    if (*(reinterpret_cast<const unsigned*>(sm0 + kParkNAbs) + warp) >= (unsigned)kParkDrain)
        drain_parked(sm0, warp, lane, slow);
}

// the two halves of a sample, split where the compiler cannot merge the halves' arithmetic back into 64-bit
// operations (an OR of eight high words must not become eight 64-bit ORs)
__device__ __forceinline__ void split64(long long v, unsigned& lo, unsigned& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}

// the lane's private counter at absolute shared address `addr` += 1
__device__ __forceinline__ void bump_abs(unsigned addr) {
#ifdef B200_HIST_RED
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(1u) : "memory");
    return;
#endif
    unsigned c;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c) : "r"(addr) : "memory");
    c += 1u;
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(c) : "memory");
}

// ---- quotient kernel ------------------------------------------------------------------------------------
// When every threshold is a multiple of the divisor (edges given in the divisor's unit: the README bins and the
// OpenCensus bounds in milliseconds) and the quotients of 32-bit samples are few (divisor >= 2^19), the bin is a
// function of q = trunc(ns / divisor), which the moments need anyway: one byte-table look-up indexed by q
// replaces the float-exponent cell, its compare and the masks.  Per sample: 3 instructions for q, 1 look-up,
// 1 address, 3 for the counter, ~3 for the moments.  Samples outside 0 <= ns < 2^32 take slow_sample().
template <bool MOMENTS>
__global__ void __launch_bounds__(kHistThreads, kHistCtasPerSm)
hist_qlut_kernel(const long long* __restrict__ x, size_t n, const HistParams* __restrict__ Pg,
                 unsigned long long* __restrict__ counts, LatMoments* __restrict__ mom, int nstages) {
    const unsigned smem_base = (unsigned)__cvta_generic_to_shared(hist_smem_raw);
    unsigned char* const sm0 = hist_smem_raw - smem_base;  // generic pointer to absolute shared address 0
    unsigned* const tbl = reinterpret_cast<unsigned*>(sm0 + kQCntAbs);  // [warp][bin][lane] private counters
    SlowParams* const sp = reinterpret_cast<SlowParams*>(sm0 + kSlowAbs);
    load_slow_params(sp, Pg);
    if (threadIdx.x < kHistWarps) reinterpret_cast<unsigned*>(sm0 + kParkNAbs)[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < kQLutSize / 4; i += kHistThreads)
        reinterpret_cast<unsigned*>(sm0 + kQLutAbs)[i] = reinterpret_cast<const unsigned*>(Pg->qlut)[i];
    const int nbins = Pg->nthr + 1, rows = nbins;
    for (int i = threadIdx.x; i < kHistWarps * rows * 32; i += kHistThreads) tbl[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned* const mine = tbl + (size_t)warp * rows * 32 + lane;
    const unsigned mine_abs = kQCntAbs + (unsigned)(warp * rows * 32 + lane) * 4u;
    const unsigned magic32 = Pg->magic32;
    const int shift32 = Pg->shift32, pre32 = Pg->pre32;
    // q <= 8191: a thread's whole share of quotients sums in 32 bits (the launcher bounds n), eight squares too
    unsigned mn32 = 0xFFFFFFFFu, mx32 = 0, fsum = 0;
    unsigned long long fsq = 0;
    SlowMoments sm_slow;
    auto quotient = [&](unsigned u) -> unsigned { return __umulhi(u >> pre32, magic32) >> shift32; };
    // (the table is read-only while the kernel counts: a plain asm, so that the eight look-ups of a round are
    // issued together instead of each waiting for the previous sample's counter update)
    auto bin_of = [&](unsigned q) -> unsigned {
        unsigned b;
        asm("ld.shared.u8 %0, [%1+%2];" : "=r"(b) : "r"(q), "n"(kQLutAbs));
        return b;
    };
    auto bump_q = [&](unsigned q) { bump_abs(mine_abs + (bin_of(q) << 7)); };
    // A sample outside 0 <= ns < 2^32 is parked; in the common code it is replaced by its clamp to that range (0
    // or 2^32 - 1).  Such a surrogate lands in a known bin (nlo / nhi count how many this thread made; the bins are
    // corrected at the end) and adds a known quotient to the sums (0, or qmax: subtracted at the end).  It cannot
    // spoil the extremes either: a 0 can lower the minimum only when a negative sample exists, which is lower
    // still, and 2^32 - 1 can raise the maximum only when a sample >= 2^32 exists; a thread that saw nothing but
    // surrogates leaves its fast-path extremes out.
    unsigned nlo = 0, nhi = 0, nproc = 0;
    auto slow = [&](long long v) { slow_sample(v, sp, mine, MOMENTS ? &sm_slow : nullptr, nullptr); };
    auto one = [&](long long v) {
        if ((unsigned long long)v >> 32) {  // negative or >= 2^32 ns
            slow(v);
            return;
        }
        const unsigned u = (unsigned)v, q = quotient(u);
        bump_q(q);
        nproc++;
        if (MOMENTS) {
            mn32 = min(mn32, u);
            mx32 = max(mx32, u);
            fsum += q;
            fsq += (unsigned long long)q * q;
        }
    };
    auto eight = [&](const longlong2 a, const longlong2 b, const longlong2 c, const longlong2 d) {
        const long long v[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
        unsigned u[8], hi[8], any = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            split64(v[k], u[k], hi[k]);
            any |= hi[k];
        }
        if (any) {  // rare: a few lanes of three rounds in four
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (hi[k]) {
                    if (!park1(sm0, warp, v[k])) slow(v[k]);
                    const bool neg = (int)hi[k] < 0;
                    u[k] = neg ? 0u : 0xFFFFFFFFu;
                    nlo += neg;
                    nhi += !neg;
                }
        }
        nproc += 8;
        unsigned q[8];
#pragma unroll
        for (int k = 0; k < 8; k++) q[k] = quotient(u[k]);
        unsigned bin[8];
#pragma unroll
        for (int k = 0; k < 8; k++) bin[k] = bin_of(q[k]);
#pragma unroll
        for (int k = 0; k < 8; k++) bump_abs(mine_abs + (bin[k] << 7));
        if (MOMENTS) {
            unsigned s = 0, sq = 0;
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
                mn32 = min(min(mn32, u[k]), u[k + 1]);
                mx32 = max(max(mx32, u[k]), u[k + 1]);
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                s += q[k];
                sq += q[k] * q[k];
            }
            fsum += s;
            fsq += sq;
        }
    };
    const unsigned ring_abs = (kQCntAbs + (unsigned)(kHistWarps * rows * 32) * 4u + 127u) & ~127u;
    for_each_sample(x, n, ring_abs, nstages, eight, one, [&] { maybe_drain(sm0, warp, lane, slow); });
    drain_parked(sm0, warp, lane, slow);

    // take the surrogates out again: their bins, their quotients
    const unsigned qhi = quotient(0xFFFFFFFFu);
    mine[bin_of(0u) * 32] -= nlo;
    mine[bin_of(qhi) * 32] -= nhi;
    long long mn = sm_slow.mn, mx = sm_slow.mx;
    if (nproc > nlo + nhi) {  // this thread saw genuine fast-path samples
        mn = (long long)mn32 < mn ? (long long)mn32 : mn;
        mx = (long long)mx32 > mx ? (long long)mx32 : mx;
    }
    U128 sum = sm_slow.sum, sq = sm_slow.sq;
    add128(sum, (unsigned long long)fsum - (unsigned long long)nhi * qhi, 0);
    add128(sq, fsq - (unsigned long long)nhi * qhi * qhi, 0);
    hist_epilogue<MOMENTS>(tbl, rows, nbins, n, counts, mom, mn, mx, sum, sq);
}

// ---- cell kernel (any thresholds) ---------------------------------------------------------------------------
// DIVMODE: how the fast path divides a 32-bit sample by the divisor: 0 divisor is 1, 1 multiply-high magic with
// the add fix-up, 2 divisor >= 2^32 (quotient 0), 3 pre-shift + multiply-high without fix-up and quotient < 2^28.
// COLLECT: cells may hold two thresholds and the samples of odd bins are appended to col.buf.
template <bool MOMENTS, int DIVMODE, bool COLLECT>
__global__ void __launch_bounds__(kHistThreads, (MOMENTS && COLLECT) ? kHistCtasPerSm - 1 : kHistCtasPerSm)
hist_kernel(const long long* __restrict__ x, size_t n, const HistParams* __restrict__ Pg,
            unsigned long long* __restrict__ counts, LatMoments* __restrict__ mom, const HistCollect col, int nstages) {
    // cell entries are 8 bytes; collecting kernels keep the cells' second compare values in a table of their own,
    // read (under a predicate) only by the lanes whose cell has one: 16-byte entries cost twice the shared-memory
    // wavefronts on every look-up and made the pass LSU-bound
    constexpr unsigned kFast2Abs = kFastAbs + kFastCells * 8;
    constexpr unsigned kCntAbs = kFast2Abs + (COLLECT ? kFastCells * 4 : 0);
    constexpr int kFastBias = (int)kFastAbs - kFastFirst * 8;   // cell address = bias + 8 * (bits >> 18)
    constexpr int kFast2Bias = (int)kFast2Abs - kFastFirst * 4;
    // (set-up has checked once per device that the dynamic window starts at or below kFastAbs: smem_base_probe)
    const unsigned smem_base = (unsigned)__cvta_generic_to_shared(hist_smem_raw);
    unsigned char* const sm0 = hist_smem_raw - smem_base;  // generic pointer to absolute shared address 0
    unsigned* const tbl = reinterpret_cast<unsigned*>(sm0 + kCntAbs);  // [warp][row][lane] private counters
    SlowParams* const sp = reinterpret_cast<SlowParams*>(sm0 + kSlowAbs);
    load_slow_params(sp, Pg);
    if (threadIdx.x < kHistWarps) reinterpret_cast<unsigned*>(sm0 + kParkNAbs)[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < kFastCells; i += kHistThreads) {
        const uint4 c = Pg->fast[i];
        // bit 3 of a collecting kernel's entry: the cell has a second compare value
        reinterpret_cast<uint2*>(sm0 + kFastAbs)[i] =
            COLLECT ? make_uint2(c.x, c.y | (c.z != 0xFFFFFFFFu && !(c.y & 4u) ? 8u : 0u)) : make_uint2(c.x, c.w);
        if (COLLECT) reinterpret_cast<unsigned*>(sm0 + kFast2Abs)[i] = c.z;
    }
    const int nbins = Pg->nthr + 1, rows = nbins + kTrashRows;
    for (int i = threadIdx.x; i < kHistWarps * rows * 32; i += kHistThreads) tbl[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned* const mine = tbl + (size_t)warp * rows * 32 + lane;
    const unsigned mine_abs = kCntAbs + (unsigned)(warp * rows * 32 + lane) * 4u;
    Moments m;
    const unsigned magic32 = Pg->magic32;
    const int shift32 = Pg->shift32, pre32 = Pg->pre32;

    // collecting kernels: a lane's pending samples sit in its own slots until some lane of the warp runs short
    const unsigned q_base = kCntAbs + (unsigned)(kHistWarps * rows * 32) * 4u + threadIdx.x * 8u;
    unsigned q_addr = q_base;

    SlowMoments sm_slow;
    // Generic path (slow_sample, 32 parked samples at a time): samples >= 2^32 ns or negative, and samples in a
    // cell holding more thresholds than the kernel compares with.  The latter are ordinary for the moments — the
    // fast path keeps those — and only their bin comes from the generic path (count_only).
    auto slow = [&](long long v) {
        const bool count_only = ((unsigned long long)v >> 32) == 0;
        slow_sample(v, sp, mine, MOMENTS && !count_only ? &sm_slow : nullptr, COLLECT ? &col : nullptr);
    };
    // Fast path: 0 <= v < 2^32 (4.29 s in nanoseconds), 32-bit arithmetic only.  lookup() fetches the sample's
    // cell (w.z is the cell's second compare value, used by collecting kernels only); bit 2 of w.y says the cell
    // holds more thresholds than we compare with (its counter offset then points at a scratch row).
    auto lookup = [&](unsigned u, uint4& w) {
        const unsigned bits = __float_as_uint(__uint2float_rz(u | 1u));
        const unsigned idx = (bits >> 15) & 0x3FFF8u;
        asm("ld.shared.v2.u32 {%0, %1}, [%2+%3];" : "=r"(w.x), "=r"(w.y) : "r"(idx), "n"(kFastBias));
        if (COLLECT) {
            asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\tmov.u32 %0, 0xFFFFFFFF;\n\t"
                "@p ld.shared.u32 %0, [%1+%3];\n\t}"
                : "=r"(w.z) : "r"(idx >> 1), "r"(w.y & 8u), "n"(kFast2Bias));
        }
    };
    // returns the byte offset of the counter row: bit 7 = the sample's bin is odd
    auto bump = [&](unsigned u, const uint4 w) -> unsigned {
        unsigned off = (COLLECT ? (w.y & ~8u) : w.y) + (u > w.x ? 128u : 0u);  // (flagged cells never get here)
        if (COLLECT) off += u > w.z ? 128u : 0u;
        bump_abs(mine_abs + off);
        return off;
    };
    // keep the sample if `in` (0 or 128) is set: the slot is written under a predicate, the cursor moves by in * 16
    auto collect = [&](long long v, unsigned in) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.shared.v2.u32 [%0], {%1, %2};\n\t}"
                     ::"r"(q_addr), "r"((unsigned)v), "r"((unsigned)((unsigned long long)v >> 32)), "r"(in) : "memory");
        static_assert(kLaneQueueStride == 128u << 4, "cursor step");
        q_addr += in << 4;
    };
    auto flush = [&]() {  // whole warp: one global atomic for everything the 32 lanes hold
        const unsigned cnt = (q_addr - q_base) / kLaneQueueStride;
        unsigned incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += t;
        }
        const unsigned total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        unsigned long long at = 0;
        if (lane == 31 && total) at = atomicAdd(col.count, (unsigned long long)total);
        at = __shfl_sync(0xFFFFFFFFu, at, 31) + (incl - cnt);
        for (unsigned e = 0; e < cnt; e++) {
            unsigned lo, hi;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(q_base + e * kLaneQueueStride) : "memory");
            if (at + e < col.cap) col.buf[at + e] = (long long)(((unsigned long long)hi << 32) | lo);
        }
        q_addr = q_base;
    };
    auto quotient = [&](unsigned u) -> unsigned {
        if (DIVMODE == 1) {
            const unsigned t = __umulhi(magic32, u);
            return (t + ((u - t) >> 1)) >> shift32;
        }
        if (DIVMODE == 3) return __umulhi(u >> pre32, magic32) >> shift32;
        return DIVMODE == 0 ? u : 0u;
    };
    auto accumulate = [&](unsigned q) {  // running 64-bit sums; q*q wraps at most once per add
        m.fsum += q;
        const unsigned long long sq = (unsigned long long)q * q + m.fsq;
        m.fsq_carry += sq < m.fsq;
        m.fsq = sq;
    };
    // In the common code a parked sample is replaced by a surrogate: its clamp to [0, 2^32) when it lies outside
    // (nlo / nhi count these: the quotient of 2^32 - 1 is taken out of the sums at the end; the extremes cannot be
    // spoiled, see hist_qlut_kernel), itself when only its cell is the problem; either way its counter update is
    // pointed at a scratch row, which is even and so never collected.
    unsigned nlo = 0, nhi = 0, nproc = 0;
    const unsigned scratch_row = (unsigned)((nbins + 1) & ~1) * 128u;  // even, past the last bin
    auto one = [&](long long v) {
        if ((unsigned long long)v >> 32) { slow(v); return; }
        const unsigned u = (unsigned)v;
        uint4 w;
        lookup(u, w);
        if (w.y & 4u) {
            slow(v);  // count only
            w = make_uint4(0xFFFFFFFFu, scratch_row, 0xFFFFFFFFu, 0);
        }
        const unsigned off = bump(u, w);
        if (COLLECT) collect(v, off & 128u);
        nproc++;
        if (MOMENTS) {
            m.mn32 = min(m.mn32, u);
            m.mx32 = max(m.mx32, u);
            accumulate(quotient(u));
        }
    };
    // Eight samples at a time.  Whether any of them needs the generic path is ONE test (the OR of the high words
    // and of the cells' flag bits); the common work carries no masks.
    auto eight = [&](const longlong2 a, const longlong2 b, const longlong2 c, const longlong2 d) {
        const long long v[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
        uint4 w[8];
        unsigned u[8], hi[8], any_hi = 0, any_y = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            split64(v[k], u[k], hi[k]);
            lookup(u[k], w[k]);
            any_hi |= hi[k];
            any_y |= w[k].y;
        }
        if (any_hi | (any_y & 4u)) {  // rare: a few lanes of three rounds in four
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (hi[k] | (w[k].y & 4u)) {
                    if (!park1(sm0, warp, v[k])) slow(v[k]);
                    if (hi[k]) {
                        const bool neg = (int)hi[k] < 0;
                        u[k] = neg ? 0u : 0xFFFFFFFFu;
                        nlo += neg;
                        nhi += !neg;
                    }
                    w[k] = make_uint4(0xFFFFFFFFu, scratch_row, 0xFFFFFFFFu, 0);
                }
        }
        nproc += 8;
        unsigned s = 0;
        unsigned long long sq = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const unsigned off = bump(u[k], w[k]);
            if (COLLECT) collect(v[k], off & 128u);
            if (MOMENTS) {
                const unsigned q = quotient(u[k]);
                if (DIVMODE == 3) {  // quotients < 2^28: eight of them fit 32 bits, eight squares fit 64
                    s += q;
                    sq += (unsigned long long)q * q;
                } else {
                    accumulate(q);
                }
            }
        }
        if (MOMENTS) {
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
                m.mn32 = min(min(m.mn32, u[k]), u[k + 1]);
                m.mx32 = max(max(m.mx32, u[k]), u[k + 1]);
            }
            if (DIVMODE == 3) {
                m.fsum += s;
                sq += m.fsq;
                m.fsq_carry += sq < m.fsq;
                m.fsq = sq;
            }
        }
    };
    // Collecting kernels: a round adds at most 8 samples to a lane's slots (7 + 8 < kLaneQueue); the generic
    // path appends to the output buffer directly.
    const unsigned ring_abs = (kCntAbs + (unsigned)(kHistWarps * rows * 32) * 4u +
                               (COLLECT ? (unsigned)kLaneQueue * kLaneQueueStride : 0u) + 127u) & ~127u;
    for_each_sample(x, n, ring_abs, nstages, eight, one, [&] {
        maybe_drain(sm0, warp, lane, slow);
        if (COLLECT && __any_sync(0xFFFFFFFFu, q_addr >= q_base + (kLaneQueue - 8) * kLaneQueueStride)) flush();
    });
    drain_parked(sm0, warp, lane, slow);
    if (COLLECT) flush();

    long long mn = sm_slow.mn, mx = sm_slow.mx;
    if (nproc > nlo + nhi) {  // this thread saw genuine fast-path samples
        mn = (long long)m.mn32 < mn ? (long long)m.mn32 : mn;
        mx = (long long)m.mx32 > mx ? (long long)m.mx32 : mx;
    }
    U128 sum = sm_slow.sum, sq = sm_slow.sq;
    add128(sum, m.fsum, 0);
    add128(sq, m.fsq, m.fsq_carry);
    if (MOMENTS && nhi) {  // take the surrogates of the samples >= 2^32 out of the sums again: nhi * q(2^32 - 1)
        const unsigned long long qhi = quotient(0xFFFFFFFFu), q2 = qhi * qhi;
        const unsigned long long s_lo = (unsigned long long)nhi * qhi;  // < 2^64
        const unsigned long long p_lo = (unsigned long long)nhi * q2, p_hi = __umul64hi((unsigned long long)nhi, q2);
        const unsigned long long b1 = sum.lo < s_lo;
        sum.lo -= s_lo;
        sum.hi -= b1;
        const unsigned long long b2 = sq.lo < p_lo;
        sq.lo -= p_lo;
        sq.hi -= p_hi + b2;
    }
    hist_epilogue<MOMENTS>(tbl, rows, nbins, n, counts, mom, mn, mx, sum, sq);
}

__global__ void __launch_bounds__(kHistThreads)
minmax_kernel(const long long* __restrict__ x, size_t n, LatMoments* __restrict__ mom) {
    long long mn = 0x7FFFFFFFFFFFFFFFll, mx = (long long)0x8000000000000000ull;
    auto one = [&](long long v) {
        mn = v < mn ? v : mn;
        mx = v > mx ? v : mx;
    };
    size_t head = ((reinterpret_cast<uintptr_t>(x) & 15) ? 1 : 0);
    if (head > n) head = n;
    const size_t tid = (size_t)blockIdx.x * kHistThreads + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * kHistThreads;
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
             const unsigned long long* __restrict__ prefix, unsigned long long* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char sm[];
    unsigned* hist = reinterpret_cast<unsigned*>(sm);  // [nk][kRadixBins]
    for (int i = threadIdx.x; i < nk * kRadixBins; i += kRadixThreads) hist[i] = 0;
    __shared__ unsigned long long pf[kMaxSelect];
    if (threadIdx.x < nk) pf[threadIdx.x] = top >= 64 ? 0 : prefix[threadIdx.x];
    __syncthreads();
    const unsigned mask = (top - shift >= kRadixBits) ? (kRadixBins - 1) : ((1u << (top - shift)) - 1);
    const bool all = top >= 64;
    const size_t nthreads = (size_t)gridDim.x * kRadixThreads;
    auto one = [&](long long v) {
        unsigned long long key = ((unsigned long long)v ^ 0x8000000000000000ull) - key_base;
        unsigned dgt = (unsigned)(key >> shift) & mask;
        unsigned long long hi = all ? 0 : key >> top;
#pragma unroll
        for (int k = 0; k < kMaxSelect; k++)
            if (k < nk && hi == pf[k]) atomicAdd(hist + k * kRadixBins + dgt, 1u);
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

// dynamic window: everything up to the end of the counters / lane slots (absolute address, see the map above),
// then the tile ring
size_t hist_smem(int nbins, bool collect, int nstages) {
    size_t end = kFastAbs + (size_t)kFastCells * (collect ? 12 : 8) +
                 (size_t)kHistWarps * (nbins + kTrashRows) * 32 * sizeof(unsigned);
    if (collect) end += (size_t)kLaneQueue * kLaneQueueStride;
    return ((end + 127) & ~(size_t)127) + kRingHeader + (size_t)nstages * kTileBytes;
}
size_t hist_qlut_smem(int nbins, int nstages) {
    const size_t end = kQCntAbs + (size_t)kHistWarps * nbins * 32 * sizeof(unsigned);
    return ((end + 127) & ~(size_t)127) + kRingHeader + (size_t)nstages * kTileBytes;
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
              unsigned long long* __restrict__ outside) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    bool wide = false;
    if (i < s) {
        unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        const long long v = __ldg(x + i * stride + z % stride);
        out[i] = v;
        wide = ((unsigned long long)v >> 32) != 0;  // negative or >= 2^32: not on the histogram kernel's fast path
    }
    // how much of the sample the 32-bit fast path cannot take: the caller decides whether a histogram pass pays
    const unsigned m = __ballot_sync(0xFFFFFFFFu, wide);
    if (outside && m && (threadIdx.x & 31) == 0) atomicAdd(outside, (unsigned long long)__popc(m));
}

}  // namespace

cudaError_t launch_sample(const long long* d_samples, size_t stride, long long* d_out, size_t s, cudaStream_t stream,
                          unsigned long long* d_outside) {
    if (d_outside) {
        cudaError_t e = cudaMemsetAsync(d_outside, 0, sizeof *d_outside, stream);
        if (e != cudaSuccess) return e;
    }
    if (!s) return cudaSuccess;
    sample_kernel<<<(unsigned)((s + 255) / 256), 256, 0, stream>>>(d_samples, stride, d_out, s, d_outside);
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
    cudaError_t e = cudaMalloc(&s->d_compact_n, sizeof(unsigned long long));
    if (e != cudaSuccess) return e;
    e = smem_base_probe(s->d_compact_n);
    if (e != cudaSuccess) return e;
    e = cudaMalloc(&s->d_params, sizeof(HistParams));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_counts, (kMaxThresholds + 1) * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_mom, sizeof(LatMoments));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_radix, (size_t)kMaxSelect * kRadixBins * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_prefix, kMaxSelect * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_sel, sizeof(SelState));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_sel_hist, (size_t)kSelMax * kRadixBins * sizeof(unsigned long long));
    return e;
}

void lat_scratch_free(LatScratch* s) {
    cudaFree(s->d_params);
    cudaFree(s->d_counts);
    cudaFree(s->d_mom);
    cudaFree(s->d_radix);
    cudaFree(s->d_prefix);
    cudaFree(s->d_compact);
    cudaFree(s->d_compact_n);
    cudaFree(s->d_sel);
    cudaFree(s->d_sel_hist);
    cudaFree(s->d_sample);
    cudaFree(s->d_collect);
    memset(s, 0, sizeof *s);
}

static cudaError_t reset_moments(const LatScratch& s, cudaStream_t stream) {
    LatMoments init{};
    init.min_ns = 0x7FFFFFFFFFFFFFFFll;
    init.max_ns = (long long)0x8000000000000000ull;
    return cudaMemcpyAsync(s.d_mom, &init, sizeof init, cudaMemcpyHostToDevice, stream);
}

cudaError_t launch_hist(const long long* d_samples, size_t n, const LatScratch& s, bool want_moments,
                        cudaStream_t stream, int nbins, int mode32, const HistCollect* collect, bool qlut) {
    cudaError_t e = cudaMemsetAsync(s.d_counts, 0, (kMaxThresholds + 1) * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    e = reset_moments(s, stream);
    if (e != cudaSuccess) return e;
    if (collect) {
        e = cudaMemsetAsync(collect->count, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
    }
    if (!n) return cudaSuccess;
    // the quotient kernel keeps a thread's whole sum of quotients (each <= 8191) in 32 bits: fine up to 2^35 samples
    if (qlut && !collect && n < ((size_t)1 << 35)) {
        int per_sm, nstages;
        hist_shape(hist_qlut_smem(nbins, 0), kHistCtasPerSm, &per_sm, &nstages);
        const size_t smem = hist_qlut_smem(nbins, nstages);
        auto k = want_moments ? hist_qlut_kernel<true> : hist_qlut_kernel<false>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        const size_t want_ctas = (n / 2 + kHistThreads - 1) / kHistThreads;
        const int grid = (int)std::min<size_t>((size_t)sm_count() * per_sm, want_ctas ? want_ctas : 1);
        k<<<grid, kHistThreads, smem, stream>>>(d_samples, n, s.d_params, s.d_counts, s.d_mom, nstages);
        return cudaGetLastError();
    }
    using Kernel = void (*)(const long long*, size_t, const HistParams*, unsigned long long*, LatMoments*, HistCollect, int);
    // without moments the quotient is never formed, so one instantiation serves all modes
    static const Kernel kernels[2][4] = {
        {hist_kernel<false, 0, false>, hist_kernel<false, 0, false>, hist_kernel<false, 0, false>,
         hist_kernel<false, 0, false>},
        {hist_kernel<true, 0, false>, hist_kernel<true, 1, false>, hist_kernel<true, 2, false>,
         hist_kernel<true, 3, false>}};
    static const Kernel collecting[2][4] = {
        {hist_kernel<false, 0, true>, hist_kernel<false, 0, true>, hist_kernel<false, 0, true>,
         hist_kernel<false, 0, true>},
        {hist_kernel<true, 0, true>, hist_kernel<true, 1, true>, hist_kernel<true, 2, true>,
         hist_kernel<true, 3, true>}};
    const Kernel k = (collect ? collecting : kernels)[want_moments ? 1 : 0][mode32];
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    // resident CTAs per SM: kHistCtasPerSm by registers, fewer when the private counter tables are large
    // (the kernel that keeps moments AND collects needs more registers than a third of the file)
    const int by_regs = collect && want_moments ? kHistCtasPerSm - 1 : kHistCtasPerSm;
    int per_sm, nstages;
    hist_shape(hist_smem(nbins, collect != nullptr, 0), by_regs, &per_sm, &nstages);
    const size_t smem = hist_smem(nbins, collect != nullptr, nstages);
    size_t want_ctas = (n / 2 + kHistThreads - 1) / kHistThreads;
    int grid = (int)std::min<size_t>((size_t)sm_count() * per_sm, want_ctas ? want_ctas : 1);
    k<<<grid, kHistThreads, smem, stream>>>(d_samples, n, s.d_params, s.d_counts, s.d_mom,
                                            collect ? *collect : HistCollect{nullptr, nullptr, 0}, nstages);
    return cudaGetLastError();
}

cudaError_t launch_minmax(const long long* d_samples, size_t n, const LatScratch& s, cudaStream_t stream) {
    cudaError_t e = reset_moments(s, stream);
    if (e != cudaSuccess || !n) return e;
    int grid = (int)std::min<size_t>((size_t)sm_count() * 8, (n + kHistThreads - 1) / kHistThreads);
    minmax_kernel<<<grid, kHistThreads, 0, stream>>>(d_samples, n, s.d_mom);
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
                              int nk, const LatScratch& s, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(s.d_radix, 0, (size_t)kMaxSelect * kRadixBins * sizeof(unsigned long long), stream);
    if (e != cudaSuccess || !n) return e;
    size_t smem = (size_t)nk * kRadixBins * sizeof(unsigned);
    e = cudaFuncSetAttribute(radix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)(kMaxSelect * kRadixBins * sizeof(unsigned)));
    if (e != cudaSuccess) return e;
    size_t want_ctas = (n / 2 + kRadixThreads - 1) / kRadixThreads;
    int grid = (int)std::min<size_t>((size_t)sm_count() * 2, want_ctas ? want_ctas : 1);
    radix_kernel<<<grid, kRadixThreads, smem, stream>>>(d_samples, n, key_base, shift, top, nk, s.d_prefix, s.d_radix);
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

void finish_hist_params(HistParams* P, long long divisor);

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
    finish_hist_params(P, divisor);
    return 0;
}

int host_build_hist_params_thr(const long long* thr, int nthr, long long divisor, HistParams* P) {
    memset(P, 0, sizeof *P);
    if (nthr < 1 || nthr > kMaxThresholds || divisor < 1) return -1;
    for (int j = 0; j < nthr; j++) {
        if (j && !(thr[j] > thr[j - 1])) return -1;
        P->thr[j] = thr[j];
    }
    P->nthr = nthr;
    finish_hist_params(P, divisor);
    return 0;
}

namespace {

// thresholds (P->thr, P->nthr) -> cell tables and division magics
void finish_hist_params(HistParams* P, long long divisor) {
    const int m = P->nthr;
    // log-linear LUT over n >= 0
    for (int idx = 0; idx < kLutCells; idx++) {
        unsigned long long lo, hi;
        if (idx < 32) lo = hi = (unsigned long long)idx;
        else {
            int e = idx / 32 + 4, mant = idx % 32;
            if (e > 62) { P->lut[idx] = (unsigned short)(m << 8); continue; }
            lo = (1ull << e) + ((unsigned long long)mant << (e - 5));
            hi = lo + (1ull << (e - 5)) - 1;
        }
        int base = 0, extra = 0;
        for (int j = 0; j < m; j++) {
            if (P->thr[j] <= (long long)lo) base++;
            else if (P->thr[j] <= (long long)hi) extra++;
        }
        P->lut[idx] = (unsigned short)((base << 8) | (extra > 255 ? 255 : extra));
    }
    // fast-path cells (see HistParams::fast): the cell of n is read off float_rz(n | 1)
    for (int t = 0; t < kFastCells; t++) {
        unsigned long long lo, hi;
        bool empty = false;
        const int e = t / 32, mant = t % 32;
        if (e >= 6) {  // bit 0 lies below the five mantissa bits: the | 1 does not move the cell
            lo = (1ull << e) + ((unsigned long long)mant << (e - 5));
            hi = lo + (1ull << (e - 5)) - 1;
        } else {  // n | 1 is the odd number 2^e + (mant >> (5 - e)); the cell is that number and the even one below
            const unsigned long long odd = (1ull << e) + (unsigned long long)(mant >> (5 - e));
            empty = (mant & ((1 << (5 - e)) - 1)) != 0 || !(odd & 1);
            lo = odd - 1;
            hi = odd;
        }
        int base = 0, extra = 0;
        for (int j = 0; j < m && !empty; j++) {
            if (P->thr[j] <= (long long)lo) base++;
            else if (P->thr[j] <= (long long)hi) extra++;
        }
        // thresholds inside a cell are > lo >= 0 and <= hi < 2^32, so thr - 1 fits 32 bits
        P->fast[t].x = extra >= 1 ? (unsigned int)(P->thr[base] - 1) : 0xFFFFFFFFu;
        P->fast[t].z = extra >= 2 ? (unsigned int)(P->thr[base + 1] - 1) : 0xFFFFFFFFu;
        // A cell with more thresholds than the kernel compares with sends the sample to the generic path; its
        // (no-op) counter update is pointed at the trash rows past the last bin, and bit 2 keeps the address
        // word-aligned.
        const unsigned int row = (unsigned int)base * 128u, generic = (unsigned int)(m + 1) * 128u | 4u;
        P->fast[t].w = extra >= 2 ? generic : row;
        P->fast[t].y = extra >= 3 ? generic : row;
    }
    // 32-bit dividends (Granlund-Montgomery, N = 32): l = ceil(log2 d), m' = floor(2^32 (2^l - d) / d) + 1,
    // q = (t + ((n - t) >> 1)) >> (l - 1) with t = umulhi(m', n)
    if (divisor == 1) {
        P->mode32 = 0;
    } else if ((unsigned long long)divisor > 0xFFFFFFFFull) {
        P->mode32 = 2;
    } else {
        unsigned long long d32 = (unsigned long long)divisor;
        int l = 0;
        while ((1ull << l) < d32) l++;
        P->magic32 = (unsigned int)((((unsigned __int128)1 << 32) * ((1ull << l) - d32)) / d32 + 1);
        P->shift32 = l - 1;
        P->mode32 = 1;
        // Cheaper when it exists: strip the divisor's trailing zero bits first (d = 2^pre * d'), then the dividend
        // has only 32 - pre bits and a 32-bit magic M = ceil(2^k / d') without fix-up is exact as soon as
        // (M d' - 2^k) (2^(32-pre) - 1) < 2^k.  Divisors >= 16 also bound the quotient below 2^28, which lets the
        // kernel sum eight quotients in 32 bits and eight squares in 64.
        int pre = 0;
        while (!((d32 >> pre) & 1)) pre++;
        const unsigned long long dp = d32 >> pre;
        if (d32 >= 16 && dp > 1) {
            const unsigned __int128 nmax = ((unsigned __int128)1 << (32 - pre)) - 1;
            for (int post = 0; post < 32; post++) {
                const unsigned __int128 two_k = (unsigned __int128)1 << (32 + post);
                const unsigned __int128 M = (two_k + dp - 1) / dp;
                if (M >> 32) break;
                if ((M * dp - two_k) * nmax < two_k) {
                    P->magic32 = (unsigned int)M;
                    P->shift32 = post;
                    P->pre32 = pre;
                    P->mode32 = 3;
                    break;
                }
            }
        }
    }
    // Quotient table (hist_qlut_kernel): usable when every threshold a 32-bit sample can reach is a multiple of
    // the divisor — then the bin is a function of q = trunc(n / divisor) alone — and the quotients are few.
    P->qlut_ok = 0;
    if (P->mode32 == 3) {
        const unsigned long long d32 = (unsigned long long)divisor, qmax = 0xFFFFFFFFull / d32;
        bool aligned = qmax < (unsigned long long)kQLutSize;
        for (int j = 0; j < m && aligned; j++)
            if (P->thr[j] > 0 && P->thr[j] <= 0xFFFFFFFFll && (unsigned long long)P->thr[j] % d32 != 0) aligned = false;
        if (aligned) {
            int c = 0;
            for (unsigned long long q = 0; q < (unsigned long long)kQLutSize; q++) {
                if (q <= qmax)
                    while (c < m && P->thr[c] <= (long long)(q * d32)) c++;
                P->qlut[q] = (unsigned char)c;
            }
            P->qlut_ok = 1;
        }
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
