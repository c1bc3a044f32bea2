// CRC-32C kernels for sm_100a.
//
// What is computed: Go's crc32.Update(crc, MakeTable(Castagnoli), p) — the per-chunk statement inside
// storage.Reader.Read that the reference's read loop drives (main.go:140) — for a whole batch of
// device-resident objects at once, bit-exactly.
//
// K1 "row kernel" (HBM-bound by design, shared-memory-lookup-bound in practice):
//   An object body is viewed as rows of 512 bytes = 128 little-endian words.  A warp owns a segment of
//   consecutive rows; lane l issues one coalesced 128-bit load per row and thereby owns word columns
//   4l..4l+3.  Each column is an independent CRC register that is advanced by a whole row per step:
//       a_c <- A^512(a_c) ^ w(r, c)          (A = "feed one zero byte", a GF(2)-linear map)
//   A^512 on a 32-bit register is four byte-indexed lookups.  The four 256-entry tables live in shared
//   memory replicated once per bank (entry v for lane l at byte v*256 + l*4 (+128 for the odd table of a
//   pair)), so lane l only ever touches bank l: lookups are bank-conflict free whatever the data, and the
//   address is a single PRMT of (state byte, lane*4).  128 KB of tables -> one 1024-thread CTA per SM.
//   At the end of a segment the 128 column registers are folded with their in-row weights
//   x^(32*(128-c)) and xor-reduced over the warp into one raw register per segment.
// Finalize kernel: per object, Horner over the segment registers with the per-segment shift constant,
//   unaligned head and sub-row tail bytes, init/final inversion, comparison with want[].
// Whole segments of large objects, and batches of many small objects (a lane per object), are the fold kernel's:
//   crc_fold_kernel.cu.  launch_batch below picks the road for a batch.
#include "crc_kernels.h"

#ifdef B200_DEBUG_ASSERT
#include <cassert>
#endif

#include <cstdlib>

namespace b200 {

namespace {

// Shared memory map of the row kernel.  The replicated tables start at ABSOLUTE shared address 0x10000 so
// that a lookup address is a pure byte permutation of (state byte, lane word): 0x00TTVVLL with
// TT = 1 + pair, VV = state byte, LL = lane*4 (+0x80 for the odd table of a pair, as an immediate).
constexpr uint32_t kTabAbs = 0x10000;     // absolute shared address of table pair {0,1}; {2,3} follow at +64 KB
constexpr uint32_t kPairStride = 65536;
constexpr uint32_t kTabBytes = 4 * 256 * 32 * 4;
constexpr uint32_t kAdv4Abs = 0x1000;     // absolute shared address of the 4 KB x^32 tables (inside the gap)

extern __shared__ __align__(1024) uint8_t smem_raw[];

// B200_DEBUG_ASSERT builds: the absolute shared-memory addresses of the table look-ups are checked against the
// CTA's dynamic window (compute-sanitizer is closed on this pool).
#ifdef B200_DEBUG_ASSERT
__device__ __forceinline__ void check_tab(uint32_t addr) {
    uint32_t size;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(size));
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem_raw);
    assert(addr >= kTabAbs && addr + 4 <= kTabAbs + kTabBytes && addr + 4 <= base + size && (addr & 3) == 0);
}
#else
__device__ __forceinline__ void check_tab(uint32_t) {}
#endif

// A^512 applied to a column register.  lanew = 0x00010000 | lane*4: byte 0 = lane*4, byte 2 = 1, byte 3 = 0.
__device__ __forceinline__ uint32_t adv_row(uint32_t x, uint32_t lanew) {
    uint32_t o0 = __byte_perm(x, lanew, 0x7604);  // [lane*4, byte0(x), 0x01, 0x00]
    uint32_t o1 = __byte_perm(x, lanew, 0x7614);
    uint32_t o2 = __byte_perm(x, lanew, 0x7624);
    uint32_t o3 = __byte_perm(x, lanew, 0x7634);
    check_tab(o0); check_tab(o1 + 128); check_tab(o2 + 65536); check_tab(o3 + 65664);
    uint32_t t0, t1, t2, t3;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t0) : "r"(o0));
    asm volatile("ld.shared.u32 %0, [%1+128];" : "=r"(t1) : "r"(o1));
    asm volatile("ld.shared.u32 %0, [%1+65536];" : "=r"(t2) : "r"(o2));
    asm volatile("ld.shared.u32 %0, [%1+65664];" : "=r"(t3) : "r"(o3));
    return t0 ^ t1 ^ t2 ^ t3;
}

// v * x^32 with the 4 KB (un-replicated) tables at kAdv4Abs; used a handful of times per segment only
__device__ __forceinline__ uint32_t mul_x32(uint32_t v) {
    uint32_t t0, t1, t2, t3;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t0) : "r"(kAdv4Abs + ((v & 0xFF) << 2)));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t1) : "r"(kAdv4Abs + 1024 + (((v >> 8) & 0xFF) << 2)));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t2) : "r"(kAdv4Abs + 2048 + (((v >> 16) & 0xFF) << 2)));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t3) : "r"(kAdv4Abs + 3072 + ((v >> 24) << 2)));
    return t0 ^ t1 ^ t2 ^ t3;
}

__device__ __forceinline__ uint4 ldg_row(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}

__device__ __forceinline__ uint32_t find_object(const uint32_t* __restrict__ seg_start, uint32_t nobj, uint32_t task) {
    uint32_t lo = 0, hi = nobj;  // invariant: seg_start[lo] <= task < seg_start[hi]
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(seg_start + mid) <= task) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(kRowsThreads, 1)
crc_rows_kernel(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off, const uint64_t* __restrict__ len,
                TaskTable tt, SegPlan plan,
                uint64_t single_off, uint64_t single_len, uint32_t skip_kinds, const uint32_t* __restrict__ adv_tab,
                const uint32_t* __restrict__ col_k, const uint32_t* __restrict__ adv4_tab, uint32_t* __restrict__ seg_raw,
                uint32_t partial_first) {
    // skip_kinds: bit k set = segments of SegKind k are the fold kernel's, not ours; partial_first: so are the
    // partial segments below this task number (equal blocks hashed a lane per object)
    // replicate the compact 4x256 table once per bank, at absolute shared address kTabAbs
    {
        uint32_t i = threadIdx.x;  // kRowsThreads == 4*256
        uint32_t t = i >> 8, v = i & 255;
        uint32_t val = __ldg(adv_tab + i);
        uint32_t dst = kTabAbs + (t >> 1) * kPairStride + v * 256 + (t & 1) * 128;
#pragma unroll
        for (int k = 0; k < 8; k++)
            asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(dst + 16 * k), "r"(val));
    }
    // (the dynamic window must cover [0x10000, 0x30000): it starts at most 1 KB up, which set-up has checked
    // once per device with smem_base_probe — b200_init and the no-alloc entry point fail there otherwise)
    // the x^32 slicing tables for the segment epilogue live in the otherwise unused gap below the big tables
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(kAdv4Abs + threadIdx.x * 4), "r"(__ldg(adv4_tab + threadIdx.x)));
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lane4 = lane * 4;
    const uint32_t lanew = 0x00010000u | lane4;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t k3 = __ldg(col_k + lane4 + 3);  // x^(32*(128-c)) of this lane's last column c = 4*lane+3
    const uint32_t nwarps = gridDim.x * kRowsWarps;
    // Kind by kind, every warp takes a CONTIGUOUS run of that kind's tasks, so the object a task belongs to is
    // found by one binary search at the start and a forward walk afterwards (a search per task is a ~20-deep
    // chain of dependent loads, which is what bounds batches of small objects).  Warps are numbered CTA-major.
    const uint32_t gw = warp * gridDim.x + blockIdx.x;
    for (uint32_t kind = 0; kind < 3; kind++) {
    if ((skip_kinds >> kind) & 1) continue;  // these segments belong to the fold kernel
    const uint32_t first = kind == kSegPartial ? partial_first : 0u;
    const uint32_t ntasks = tt.count[kind] > first ? tt.count[kind] - first : 0u;  // ours: [first, count)
    const uint32_t* __restrict__ seg_start = off ? tt.prefix(kind) : nullptr;
    const uint32_t raw_base = tt.base(kind);
    const uint32_t task_lo = first + (uint32_t)((uint64_t)ntasks * gw / nwarps);
    const uint32_t task_hi = first + (uint32_t)((uint64_t)ntasks * (gw + 1) / nwarps);
    uint32_t obj = 0, obj_first = 0, obj_next = first + ntasks;  // tasks [obj_first, obj_next) belong to obj
    if (off && task_lo < task_hi) {
        obj = find_object(seg_start, tt.nobj, task_lo);
        obj_first = __ldg(seg_start + obj);
        obj_next = __ldg(seg_start + obj + 1);
    }

    for (uint32_t task = task_lo; task < task_hi; task++) {
        uint32_t seg = task;
        uint64_t o = single_off, n = single_len;
        if (off) {
            while (task >= obj_next) {  // objects without segments of this kind are stepped over
                obj++;
                obj_first = obj_next;
                obj_next = __ldg(seg_start + obj + 1);
            }
            seg = task - obj_first;
            o = __ldg(off + obj);
            n = __ldg(len + obj);
        }
        const ObjectCut cut = cut_object((uint64_t)(uintptr_t)base + o, n);
        const Seg sg = seg_of(cut.rows, plan, kind, seg);
        const uint64_t row0 = sg.row0;
        const uint32_t R = sg.rows;
        const uint4* p = reinterpret_cast<const uint4*>(base + o + cut.head + row0 * kRowBytes) + lane;

        uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#define B200_STEP(ROW)                         \
    a0 = adv_row(a0, lanew) ^ (ROW).x;         \
    a1 = adv_row(a1, lanew) ^ (ROW).y;         \
    a2 = adv_row(a2, lanew) ^ (ROW).z;         \
    a3 = adv_row(a3, lanew) ^ (ROW).w;
        // Main loop: 8 rows per trip in two register sets so that four 128-bit loads are always in flight
        // and nothing is copied.  The prefetch of the trip after the last one is clamped to valid rows.
        uint32_t r = 0;
        if (R >= 8) {
            const uint32_t last = R - 4;
            uint4 x0 = ldg_row(p), x1 = ldg_row(p + 32), x2 = ldg_row(p + 64), x3 = ldg_row(p + 96);
            for (; r + 8 <= R; r += 8) {
                const uint4* qb = p + (size_t)(r + 4) * 32;
                uint4 y0 = ldg_row(qb), y1 = ldg_row(qb + 32), y2 = ldg_row(qb + 64), y3 = ldg_row(qb + 96);
                B200_STEP(x0) B200_STEP(x1) B200_STEP(x2) B200_STEP(x3)
                const uint32_t nx = r + 8 < last ? r + 8 : last;
                const uint4* qa = p + (size_t)nx * 32;
                x0 = ldg_row(qa); x1 = ldg_row(qa + 32); x2 = ldg_row(qa + 64); x3 = ldg_row(qa + 96);
                B200_STEP(y0) B200_STEP(y1) B200_STEP(y2) B200_STEP(y3)
            }
        }
        for (; r < R; r++) {  // at most 7 rows (only the last segment of an object)
            uint4 x = ldg_row(p + (size_t)r * 32);
            B200_STEP(x)
        }
#undef B200_STEP
        // fold the 128 columns: column c still has to pass over 4*(128-c) bytes of its row.  In-lane the four
        // columns are 4 bytes apart (three x^32 table steps), then one multiply by the lane's constant.
        uint32_t u = mul_x32(a0) ^ a1;
        u = mul_x32(u) ^ a2;
        u = mul_x32(u) ^ a3;
        uint32_t s = gf2_mul(k3, u);
#pragma unroll
        for (int d = 16; d; d >>= 1) s ^= __shfl_xor_sync(0xFFFFFFFFu, s, d);
        if (lane == 0) seg_raw[raw_base + task] = s;
    }
    }  // kind
}

__device__ __forceinline__ uint32_t adv_tab4(const uint32_t* __restrict__ t, uint32_t s) {
    return __ldg(t + (s & 0xFF)) ^ __ldg(t + 256 + ((s >> 8) & 0xFF)) ^ __ldg(t + 512 + ((s >> 16) & 0xFF)) ^
           __ldg(t + 768 + (s >> 24));
}

// After crc_fold_objects_kernel: a thread per object adds what the stream of whole 16-byte words left out — the
// head up to the first 16-byte boundary, the tail of fewer than 16 bytes — and the initial value, and gives the
// verdict.  s <- bytes(head); s <- x^(8 m) s ^ raw; s <- bytes(tail).
__global__ void __launch_bounds__(kFinalThreads)
crc_finalize_objects_kernel(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off,
                            const uint64_t* __restrict__ len, const uint32_t* __restrict__ want,
                            const uint32_t* __restrict__ raw, const uint32_t* __restrict__ xpow128,
                            const uint32_t* __restrict__ byte_tab, uint32_t init_crc, uint32_t nobj,
                            uint32_t* __restrict__ got, uint8_t* __restrict__ ok, unsigned long long* nbad) {
    __shared__ uint32_t bt[256];
    for (int i = threadIdx.x; i < 256; i += kFinalThreads) bt[i] = __ldg(byte_tab + i);
    __syncthreads();
    const uint32_t obj = blockIdx.x * kFinalThreads + threadIdx.x;
    if (obj >= nobj) return;
    const uint64_t o = off[obj], n = len[obj];
    const uint8_t* p = base + o;
    const uint64_t h = (16 - (reinterpret_cast<uintptr_t>(p) & 15)) & 15;
    const uint32_t head = (uint32_t)(h < n ? h : n);
    const uint64_t m = (n - head) & ~(uint64_t)15;
    const uint32_t tail = (uint32_t)(n - head - m);
    // (an aligned 16-byte word that holds a byte of the object cannot cross a page: reading all of it is safe)
    auto bytes_of = [&](const uint4& q, uint32_t from, uint32_t count, uint32_t st) {
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (uint32_t k = 0; k < 16; k++)
            if (k >= from && k < from + count) st = (st >> 8) ^ bt[(st ^ (w[k >> 2] >> (8 * (k & 3)))) & 0xFF];
        return st;
    };
    uint4 hq = make_uint4(0, 0, 0, 0), tq = hq;
    if (head) hq = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)15));
    if (tail) tq = __ldg(reinterpret_cast<const uint4*>(p + head + m));
    uint32_t s = ~init_crc;
    if (head) s = bytes_of(hq, (uint32_t)(reinterpret_cast<uintptr_t>(p) & 15), head, s);
    if (m) s = gf2_mul(__ldg(xpow128 + (m >> 4)), s) ^ raw[obj];  // (m <= 32 KiB: the shift comes from a table)
    if (tail) s = bytes_of(tq, 0, tail, s);
    const uint32_t crc = ~s;
    got[obj] = crc;
    const bool good = want ? (crc == want[obj]) : true;
    ok[obj] = good ? 1 : 0;
    if (!good) atomicAdd(nbad, 1ull);
}

// One CTA per object.
__global__ void __launch_bounds__(kFinalThreads)
crc_finalize_kernel(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off, const uint64_t* __restrict__ len,
                    TaskTable tt, const uint32_t* __restrict__ want,
                    const uint32_t* __restrict__ seg_raw, const uint32_t* __restrict__ adv_seg,
                    const uint32_t* __restrict__ adv_seg2, const uint32_t* __restrict__ xsq,
                    const uint32_t* __restrict__ byte_tab, SegPlan plan, uint32_t init_crc, const uint32_t* init_ptr,
                    uint64_t single_off, uint64_t single_len, uint32_t* got, uint8_t* __restrict__ ok,
                    unsigned long long* nbad) {
    __shared__ uint32_t red[kFinalThreads / 32];
    const uint32_t obj = blockIdx.x;
    const uint64_t o = off ? off[obj] : single_off;
    const uint64_t n = off ? len[obj] : single_len;
    const ObjectCut cut = cut_object((uint64_t)(uintptr_t)base + o, n);
    // where this object's raw segment registers sit: one run per kind
    const uint32_t first = off ? tt.prefix(kSegLong)[obj] : 0;
    const uint32_t first_short = tt.base(kSegShort) + (off ? tt.prefix(kSegShort)[obj] : 0);
    const uint32_t first_part = tt.base(kSegPartial) + (off ? tt.prefix(kSegPartial)[obj] : 0);
    const uint64_t nfull = cut.rows / plan.a;                       // long segments
    const uint64_t rem = cut.rows - nfull * plan.a;
    const uint32_t nshort = plan.b ? (uint32_t)(rem / plan.b) : 0;  // short segments (< a/b of them)
    const uint32_t last_rows = (uint32_t)(rem - (uint64_t)nshort * plan.b);
    const uint64_t seg_bytes = (uint64_t)plan.a * kRowBytes;
    const uint8_t* p = base + o;

    // F = sum over long segments j of x^(8*seg_bytes*(nfull-1-j)) * raw_j
    uint32_t F = 0;
    if (nfull > 64) {
        const uint64_t per = (nfull + kFinalThreads - 1) / kFinalThreads;
        uint64_t lo = (uint64_t)threadIdx.x * per, hi = lo + per;
        if (hi > nfull) hi = nfull;
        uint32_t part = 0;
        if (lo < hi) {
            for (uint64_t j = lo; j < hi; j++) part = adv_tab4(adv_seg, part) ^ seg_raw[first + j];
            uint64_t after = (nfull - hi) * seg_bytes;
            if (after) part = gf2_mul(gf2_xpow8(after, xsq), part);
        }
#pragma unroll
        for (int d = 16; d; d >>= 1) part ^= __shfl_xor_sync(0xFFFFFFFFu, part, d);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
        __syncthreads();
        if (threadIdx.x == 0)
            for (int w = 0; w < kFinalThreads / 32; w++) F ^= red[w];
    } else if (threadIdx.x == 0) {
        for (uint64_t j = 0; j < nfull; j++) F = adv_tab4(adv_seg, F) ^ seg_raw[first + j];
    }
    if (threadIdx.x != 0) return;

    uint32_t s = ~(init_ptr ? *init_ptr : init_crc);
    for (uint32_t i = 0; i < cut.head; i++) s = (s >> 8) ^ __ldg(byte_tab + ((s ^ p[i]) & 0xFF));
    if (nfull) s = gf2_mul(gf2_xpow8(nfull * seg_bytes, xsq), s) ^ F;
    for (uint32_t j = 0; j < nshort; j++) s = adv_tab4(adv_seg2, s) ^ seg_raw[first_short + j];
    if (last_rows) s = gf2_mul(gf2_xpow8((uint64_t)last_rows * kRowBytes, xsq), s) ^ seg_raw[first_part];
    const uint8_t* t = p + cut.head + cut.rows * kRowBytes;  // 16-byte aligned
    uint32_t i = 0;
    for (; i + 4 <= cut.tail; i += 4) {  // sub-row tail, a word at a time
        uint32_t x = *reinterpret_cast<const uint32_t*>(t + i) ^ s;
#pragma unroll
        for (int k = 0; k < 4; k++) x = (x >> 8) ^ __ldg(byte_tab + (x & 0xFF));
        s = x;
    }
    for (; i < cut.tail; i++) s = (s >> 8) ^ __ldg(byte_tab + ((s ^ t[i]) & 0xFF));
    const uint32_t crc = ~s;
    got[obj] = crc;
    const bool good = want ? (crc == want[obj]) : true;
    ok[obj] = good ? 1 : 0;
    if (!good) atomicAdd(nbad, 1ull);
}

// Thread-per-object variant for batches whose objects all have few segments (the many-small-objects case:
// a CTA per object would be a million nearly empty CTAs).
__global__ void __launch_bounds__(kFinalThreads)
crc_finalize_small_kernel(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off,
                          const uint64_t* __restrict__ len, TaskTable tt,
                          const uint32_t* __restrict__ want, const uint32_t* __restrict__ seg_raw,
                          const uint32_t* __restrict__ adv_seg, const uint32_t* __restrict__ adv_seg2,
                          const uint32_t* __restrict__ xsq, const uint32_t* __restrict__ byte_tab,
                          const uint32_t* __restrict__ adv4_tab, SegPlan plan, uint32_t init_crc, uint32_t nobj,
                          uint32_t* __restrict__ got, uint8_t* __restrict__ ok, unsigned long long* nbad,
                          uint32_t first_obj) {  // objects below first_obj were finished by the kernel that hashed them
    __shared__ uint32_t bt[256];
    __shared__ uint32_t w4[1024];  // x^32 slicing tables: a whole tail word per step instead of four byte steps
    for (int i = threadIdx.x; i < 256; i += kFinalThreads) bt[i] = __ldg(byte_tab + i);
    for (int i = threadIdx.x; i < 1024; i += kFinalThreads) w4[i] = __ldg(adv4_tab + i);
    __syncthreads();
    const uint32_t obj = first_obj + blockIdx.x * kFinalThreads + threadIdx.x;
    if (obj >= nobj) return;
    const uint64_t o = off[obj], n = len[obj];
    const ObjectCut cut = cut_object((uint64_t)(uintptr_t)base + o, n);
    const uint32_t first = tt.prefix(kSegLong)[obj];
    const uint32_t first_short = tt.base(kSegShort) + tt.prefix(kSegShort)[obj];
    const uint32_t first_part = tt.base(kSegPartial) + tt.prefix(kSegPartial)[obj];
    const uint64_t nfull = cut.rows / plan.a;
    const uint64_t rem = cut.rows - nfull * plan.a;
    const uint32_t nshort = plan.b ? (uint32_t)(rem / plan.b) : 0;
    const uint32_t last_rows = (uint32_t)(rem - (uint64_t)nshort * plan.b);
    const uint8_t* p = base + o;
    // One thread walks the whole object, so every dependent global load is ~a microsecond of latency chain:
    // the head comes from the ONE aligned 16-byte word that contains it, the tail is read in 16-byte words one
    // ahead of use.  (Reading the rest of an aligned 16-byte word past either end of the object is safe: such a
    // word holds at least one byte of the object and cannot cross a page, so it is mapped.)
    auto bytes_of = [&](const uint4& q, uint32_t from, uint32_t count, uint32_t st) {
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (uint32_t k = 0; k < 16; k++)
            if (k >= from && k < from + count) st = (st >> 8) ^ bt[(st ^ (w[k >> 2] >> (8 * (k & 3)))) & 0xFF];
        return st;
    };
    auto word_step = [&](uint32_t st, uint32_t w) {
        const uint32_t x = w ^ st;
        return w4[x & 0xFF] ^ w4[256 + ((x >> 8) & 0xFF)] ^ w4[512 + ((x >> 16) & 0xFF)] ^ w4[768 + (x >> 24)];
    };
    const uint8_t* t = p + cut.head + cut.rows * kRowBytes;  // 16-byte aligned start of the tail
    const uint32_t nvec = (cut.tail + 15) / 16;
    uint4 q = make_uint4(0, 0, 0, 0), hq = q;
    if (cut.head) hq = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)15));
    if (nvec) q = __ldg(reinterpret_cast<const uint4*>(t));
    uint32_t s = ~init_crc;
    if (cut.head) s = bytes_of(hq, (uint32_t)(reinterpret_cast<uintptr_t>(p) & 15), cut.head, s);
    // the state walks over the segments in order: s <- x^(8*segment bytes) * s ^ raw(segment)
    for (uint64_t j = 0; j < nfull; j++) s = adv_tab4(adv_seg, s) ^ seg_raw[first + j];
    for (uint32_t j = 0; j < nshort; j++) s = adv_tab4(adv_seg2, s) ^ seg_raw[first_short + j];
    if (last_rows) s = gf2_mul(gf2_xpow8((uint64_t)last_rows * kRowBytes, xsq), s) ^ seg_raw[first_part];
    for (uint32_t v = 0; v < nvec; v++) {
        const uint4 cur = q;
        if (v + 1 < nvec) q = __ldg(reinterpret_cast<const uint4*>(t) + v + 1);
        const uint32_t left = cut.tail - v * 16;
        if (left >= 16) {
            s = word_step(word_step(word_step(word_step(s, cur.x), cur.y), cur.z), cur.w);
        } else {
            s = bytes_of(cur, 0, left, s);
        }
    }
    const uint32_t crc = ~s;
    got[obj] = crc;
    const bool good = want ? (crc == want[obj]) : true;
    ok[obj] = good ? 1 : 0;
    if (!good) atomicAdd(nbad, 1ull);
}

__device__ __forceinline__ uint64_t splitmix64_dev(uint64_t seed, uint64_t index) {
    uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void fill_payload_kernel(uint64_t* __restrict__ dst, uint64_t first_word, size_t nwords, uint64_t seed) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += stride)
        dst[i] = splitmix64_dev(seed, first_word + i);
}

__global__ void fill_latency_kernel(int64_t* __restrict__ dst, uint64_t first, size_t n, uint64_t seed) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t idx = first + i;
        uint64_t z = splitmix64_dev(seed, 2 * idx), z2 = splitmix64_dev(seed, 2 * idx + 1);
        uint64_t s = (z & 0xFFFF) + ((z >> 16) & 0xFFFF) + ((z >> 32) & 0xFFFF) + (z >> 48);
        uint64_t ns = 20000000ull + (s * 75000000ull) / 262141ull;
        uint64_t q = z2 & 0xFF;
        if (q < 64) ns = (ns / 1000000ull) * 1000000ull;
        else if (q < 128) ns = (ns / 1000ull) * 1000ull;
        uint64_t sel = (z2 >> 8) % 100;
        if (sel == 0) ns = 95000000ull + (z2 >> 20) % (10000000000ull - 95000000ull);
        else if (sel == 1) ns = (z2 >> 20) % 20000000ull;
        dst[i] = (int64_t)ns;
    }
}

}  // namespace

// dynamic shared memory: everything up to the end of the tables at absolute 0x10000 + 128 KB (the driver's
// reserved 1 KB sits below the dynamic window; set-up checks the window start once per device)
bool fold_enabled(uint32_t rows_per_seg) {
    static const bool disabled = [] {  // A/B switch, read once
        const char* e = getenv("B200_NO_FOLD");
        return e && *e && *e != '0';
    }();
    // a lane owns rows_per_seg * 16 bytes of the segment and consumes them in whole stages
    return !disabled && rows_per_seg >= kFoldMinRows && rows_per_seg % 16 == 0 &&
           (rows_per_seg * 16u) % fold_lane_bytes() == 0;
}

size_t rows_kernel_smem() { return kTabAbs + kTabBytes; }

int rows_kernel_grid(int dev) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;  // persistent: one 1024-thread CTA per SM (128 KB of tables each)
}

cudaError_t launch_batch(const BatchArgs& a, const DeviceTables& t, int grid, cudaStream_t stream,
                         cudaEvent_t ev_begin, cudaEvent_t ev_end, const SideStream* side) {
    static thread_local int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(crc_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)rows_kernel_smem());
        if (e != cudaSuccess) return e;
        attr_dev = dev;
    }
    cudaError_t e = cudaMemsetAsync(a.nbad, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    if (ev_begin) cudaEventRecord(ev_begin, stream);
    if (!a.no_fold && a.off && a.lane_objects && a.small_objects && !a.uniform_objs && t.xpow128 &&
        fold_enabled(kFoldMinRows)) {
        // many small ragged objects: a lane per object on the fold kernel, then a thread per object for heads and tails
        e = launch_fold_objects(a.base, a.off, a.len, a.nobj, a.lane_order, t, a.seg_raw, grid, stream);
        if (e != cudaSuccess) return e;
        if (ev_end) cudaEventRecord(ev_end, stream);
        crc_finalize_objects_kernel<<<(a.nobj + kFinalThreads - 1) / kFinalThreads, kFinalThreads, 0, stream>>>(
            a.base, a.off, a.len, a.want, a.seg_raw, t.xpow128, t.byte_tab, a.init_crc, a.nobj, a.got, a.ok, a.nbad);
        return cudaGetLastError();
    }
    const SegPlan plan{a.rows_per_seg, a.rows_per_seg2};
    const TaskTable tt{a.seg_start, a.nobj, {a.nfull_tasks, a.nshort_tasks, a.npartial_tasks}};
    const bool fold_long = !a.no_fold && fold_enabled(plan.a) && a.nfull_tasks > 0;
    const bool fold_short = !a.no_fold && plan.b && fold_enabled(plan.b) && a.nshort_tasks > 0;
    const uint32_t skip_kinds = (fold_long ? 1u << kSegLong : 0u) | (fold_short ? 1u << kSegShort : 0u);
    // equal blocks back to back, one partial segment each: a lane per object on the fold kernel for every whole
    // group of 32 (the row kernel pays a table look-up per byte and ~450 instructions of set-up and epilogue per object)
    uint32_t lane_objs = 0;
    if (!a.no_fold && a.off && a.small_objects && a.uniform_objs && a.nfull_tasks == 0 && a.nshort_tasks == 0 &&
        a.npartial_tasks == a.nobj &&
        fold_lane_block_ok(a.uniform_len) && ((uintptr_t)(a.base + a.uniform_off0) & 15) == 0)
        lane_objs = a.uniform_objs;
    const bool rows = a.ntasks && (a.npartial_tasks > lane_objs || (!fold_long && a.nfull_tasks > 0) ||
                                   (!fold_short && a.nshort_tasks > 0));
    // The long segments run on `stream`; the short ones and the row kernel run next to them on the side stream
    // (when there is one), so that their CTAs fill the SMs the long fold kernel's tail leaves idle.
    cudaStream_t second = stream;
    const bool forked = fold_long && (fold_short || rows) && side;
    if (forked) {
        e = cudaEventRecord(side->fork, stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(side->stream, side->fork, 0);
        if (e != cudaSuccess) return e;
        second = side->stream;
    }
    if (fold_long) {
        e = launch_fold(a, t, grid, stream, kSegLong);
        if (e != cudaSuccess) return e;
    }
    if (fold_short) {
        e = launch_fold(a, t, grid, second, kSegShort);
        if (e != cudaSuccess) return e;
    }
    if (lane_objs) {
        e = launch_fold_per_lane(a, t, grid, second, lane_objs);
        if (e != cudaSuccess) return e;
    }
    if (rows) {
        uint32_t mine = 0;  // tasks of the kinds the row kernel takes
        for (uint32_t k = 0; k < 3; k++) {
            const uint32_t cnt = k == kSegPartial ? tt.count[k] - lane_objs : tt.count[k];
            if (!((skip_kinds >> k) & 1)) mine = mine > cnt ? mine : cnt;
        }
        uint32_t need = (mine + kRowsWarps - 1) / kRowsWarps;
        int g = (int)(need < (uint32_t)grid ? need : (uint32_t)grid);
        crc_rows_kernel<<<g ? g : 1, kRowsThreads, rows_kernel_smem(), second>>>(
            a.base, a.off, a.len, tt, plan, a.single_off, a.single_len, skip_kinds, t.adv_row, t.col_k, t.adv4, a.seg_raw,
            lane_objs);
    }
    if (forked) {  // join
        e = cudaEventRecord(side->join, second);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(stream, side->join, 0);
        if (e != cudaSuccess) return e;
    }
    if (ev_end) cudaEventRecord(ev_end, stream);
    if (a.nobj && a.off && a.small_objects) {
        if (a.nobj > lane_objs)
            crc_finalize_small_kernel<<<(a.nobj - lane_objs + kFinalThreads - 1) / kFinalThreads, kFinalThreads, 0, stream>>>(
                a.base, a.off, a.len, tt, a.want, a.seg_raw, a.adv_seg, a.adv_seg2, t.xpow8_sq, t.byte_tab, t.adv4, plan,
                a.init_crc, a.nobj, a.got, a.ok, a.nbad, lane_objs);
    } else if (a.nobj)
        crc_finalize_kernel<<<a.nobj, kFinalThreads, 0, stream>>>(a.base, a.off, a.len, tt, a.want, a.seg_raw,
                                                                  a.adv_seg, a.adv_seg2, t.xpow8_sq, t.byte_tab, plan,
                                                                  a.init_crc, a.init_ptr, a.single_off, a.single_len, a.got,
                                                                  a.ok, a.nbad);
    return cudaGetLastError();
}

cudaError_t launch_fill_payload(uint64_t* dst, uint64_t first_word, size_t nwords, uint64_t seed, cudaStream_t s) {
    if (!nwords) return cudaSuccess;
    fill_payload_kernel<<<148 * 8, 256, 0, s>>>(dst, first_word, nwords, seed);
    return cudaGetLastError();
}

cudaError_t launch_fill_latency(int64_t* dst, uint64_t first, size_t n, uint64_t seed, cudaStream_t s) {
    if (!n) return cudaSuccess;
    fill_latency_kernel<<<148 * 8, 256, 0, s>>>(dst, first, n, seed);
    return cudaGetLastError();
}

}  // namespace b200
