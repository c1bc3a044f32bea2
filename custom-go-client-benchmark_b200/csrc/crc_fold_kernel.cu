// K2 "fold kernel": CRC-32C of full segments with NO table lookups in the hot loop.
//
// Same contract as the row kernel (crc_kernels.cu): for every full segment (rows_per_seg rows of 512 bytes)
// write the raw CRC register of that segment to seg_raw[task].  Partial segments are left to the row kernel.
//
// Idea.  The row kernel pays one shared-memory lookup per byte and is bound by the L1/shared data path
// (128 B/clk/SM) — and by the power those lookups draw.  Here each lane instead reduces its own contiguous
// stream modulo a SPARSE multiple of the CRC polynomial,
//     Q'(x) = y^64 + y^48 + y^5 + x^2 (y^54 + y^37 + y^26),   y = x^32
// ( = y^5 * (y^59 + y^43 + x^2 (y^49 + y^32 + y^21) + 1), found by tools/find_word_aligned_multiples.c; P | Q').
// Reducing modulo Q' is a linear feedback shift register over 32-bit WORDS: when a new word enters, the word
// T that leaves the 64-word window is xor-ed back at five places, two of them word aligned and three of them
// shifted by the same 2 bits (one funnel shift of (T_prev, T) serves all three).  That is 1 SHF + 5 LOP3 per
// 32-bit word on the ALU pipe and nothing else — no table, no multiply.  The 64-word state lives in 64
// registers; the window slides by renaming (the loop is unrolled 64 steps, so every index is static).
// Only when a lane has consumed its whole stream is the 64-word state reduced modulo P itself (256 byte-table
// lookups per lane, ~3 % of the work at 16 KB per lane), weighted by x^(8*bytes_after_this_lane) and
// xor-reduced over the warp.
//
// Data movement.  A lane needs a CONTIGUOUS stream, which coalesced register loads cannot give; so the warp
// stages it through shared memory with cp.async (LDGSTS, 16 B per lane, no registers).  In the shipped shape
// <8 warps, 2 stages, 256 B per lane per stage> one instruction moves two lanes' next 256 bytes fully coalesced
// (16 lanes x 16 B per row), sixteen instructions fill a 32-row x 256-byte stage (8 KB).  Rows are stored with
// a 128-byte XOR swizzle (chunk j of row l at chunk j ^ (l & 7)), which keeps the per-lane 128-bit reads and
// most of the copies off each other's banks (ncu: 5.9 % of the shared-memory wavefronts conflict).  Two stages
// per warp are in flight (commit/wait groups): 8 warps x 2 x 8 KB + 4 KB of tables = 132 KB of shared memory.
// The other shapes in kVariants (B200_FOLD_VARIANT) are kept for A/B runs; their measured rates are listed there.
// (TMA bulk copies were tried first: UBLKCP takes uniform operands only, so per-lane copies serialise into a
// 32-iteration loop per warp.)
#include "crc_kernels.h"

#include <cstdlib>

namespace b200 {

namespace {

// lane-per-object mode: where a lane leaves its object's checksum and verdict
struct BlockOut {
    const uint32_t* want;
    uint32_t* got;
    uint8_t* ok;
    unsigned long long* nbad;
    uint32_t init_crc;
};

constexpr int kFoldSmemTab = 4 * 256 * 4;  // A^4 slicing tables for the final reduction (not replicated)

extern __shared__ __align__(1024) uint8_t fold_smem[];

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// (.L2::128B / .L2::256B prefetch hints on this copy were measured: no difference.)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// One LFSR step at body position T (0..63).  Logical word k of the window at step t is r[(k - t) & 63]; the
// word leaving the window and the word entering it share one register.
#define FOLD_STEP(T, W)                                           \
    {                                                             \
        const uint32_t top = r[(63 - (T)) & 63];                  \
        r[(63 - (T)) & 63] = (W);                                 \
        r[(48 - (T)-1) & 63] ^= top;                              \
        r[(5 - (T)-1) & 63] ^= top;                               \
        const uint32_t f = __funnelshift_r(tprev, top, 2);        \
        r[(55 - (T)-1) & 63] ^= f;                                \
        r[(38 - (T)-1) & 63] ^= f;                                \
        r[(27 - (T)-1) & 63] ^= f;                                \
        tprev = top;                                              \
    }
#define FOLD_QUAD(T, V) FOLD_STEP((T), (V).x) FOLD_STEP((T) + 1, (V).y) FOLD_STEP((T) + 2, (V).z) FOLD_STEP((T) + 3, (V).w)
// one stage = 32 words = 8 quads, at body offset B (0 or 32)
#define FOLD_STAGE(B, ADDR)                                                                       \
    {                                                                                             \
        uint4 q0 = lds128((ADDR) + (sw ^ 0)), q1 = lds128((ADDR) + (sw ^ 16)), q2 = lds128((ADDR) + (sw ^ 32)),   \
              q3 = lds128((ADDR) + (sw ^ 48)), q4 = lds128((ADDR) + (sw ^ 64)), q5 = lds128((ADDR) + (sw ^ 80)),  \
              q6 = lds128((ADDR) + (sw ^ 96)), q7 = lds128((ADDR) + (sw ^ 112));                              \
        FOLD_QUAD((B) + 0, q0) FOLD_QUAD((B) + 4, q1) FOLD_QUAD((B) + 8, q2) FOLD_QUAD((B) + 12, q3)   \
        FOLD_QUAD((B) + 16, q4) FOLD_QUAD((B) + 20, q5) FOLD_QUAD((B) + 24, q6) FOLD_QUAD((B) + 28, q7) \
    }

__device__ __forceinline__ uint32_t find_object2(const uint32_t* __restrict__ seg_start, uint32_t nobj, uint32_t task) {
    uint32_t lo = 0, hi = nobj;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(seg_start + mid) <= task) lo = mid; else hi = mid;
    }
    return lo;
}

// WARPS warps per CTA (one CTA per SM), STAGES staging buffers per warp, LANE_BYTES (128 or 256) bytes per lane
// per stage.  Larger LANE_BYTES means longer contiguous bursts per stream in DRAM but fewer warps fit.
template <int WARPS, int STAGES, int LANE_BYTES>
__global__ void __launch_bounds__(WARPS * 32, 1)
crc_fold_kernel(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off, const uint64_t* __restrict__ len,
                TaskTable tt, SegPlan plan, uint32_t kind,
                uint64_t single_off, uint64_t single_len, const uint32_t* __restrict__ adv4_tab,
                const uint32_t* __restrict__ xsq, uint32_t* __restrict__ seg_raw, uint32_t block_rows,
                uint32_t block_tasks, BlockOut bo) {
    // block_rows != 0: a lane per OBJECT.  The batch is equal blocks of block_rows * 16 bytes back to back from
    // base + single_off; a task is 32 consecutive blocks, lane l folds block 32 * task + l exactly as it folds its
    // share of a segment; its raw register, unweighted, is what the row kernel would have written for that object's
    // only segment, and since the object is nothing but that segment the lane finishes it on the spot (BlockOut):
    // crc = ~(x^(8 len) * ~init ^ raw), compared with want[], no seg_raw round trip and no finalize launch.  A
    // million 4 KiB blocks then cost 256 table look-ups per object instead of the row kernel's 4096.
    constexpr int kThreads = WARPS * 32;
    constexpr int kWarpStageBytes = 32 * LANE_BYTES;
    constexpr int kLanesPerRow = LANE_BYTES / 16;      // copy role: this many lanes move one row's stage
    constexpr int kRowsPerInstr = 32 / kLanesPerRow;
    constexpr int kCopyInstrs = 32 / kRowsPerInstr;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* tab = reinterpret_cast<uint32_t*>(fold_smem);  // [4][256]
    uint8_t* stage_base = fold_smem + kFoldSmemTab;          // [warps][stages][32 rows x LANE_BYTES]
    for (uint32_t i = threadIdx.x; i < 1024; i += kThreads) tab[i] = __ldg(adv4_tab + i);
    __syncthreads();

    const uint32_t rows_per_seg = block_rows ? block_rows : kind == kSegLong ? plan.a : plan.b;  // this launch's length
    const uint32_t lane_bytes = rows_per_seg * 16;   // S: bytes of the segment owned by one lane
    const uint32_t nstage = lane_bytes / LANE_BYTES;  // exact: fold_enabled() admits only lengths with S % LANE_BYTES == 0
    // weight of this lane's register: the (31 - lane) lanes after it still have S bytes each to pass
    const uint32_t lane_k = gf2_xpow8((uint64_t)lane_bytes * (31 - lane), xsq);
    // lane-per-object mode: the initial register carried over the block, the same for every block
    const uint32_t block_init = block_rows ? gf2_mul(gf2_xpow8((uint64_t)lane_bytes, xsq), ~bo.init_crc) : 0u;
    const uint32_t warp_stage = smem_u32(stage_base + (size_t)warp * STAGES * kWarpStageBytes);
    const uint32_t my_row = warp_stage + lane * LANE_BYTES;  // this lane's row inside a stage
    const uint32_t sw = (lane & 7) << 4;                     // its swizzle (chunk j lives at chunk j ^ (row & 7))
    const uint32_t crow = lane / kLanesPerRow, cchunk = lane % kLanesPerRow;
    const uint32_t nwarps = gridDim.x * WARPS;
    // this launch's tasks: the segments of kind `kind`, numbered 0..ntasks-1 over the whole batch, dealt round
    // robin so every warp gets the same number (+-1)
    const uint32_t ntasks = block_rows ? block_tasks : tt.count[kind], raw_base = block_rows ? 0 : tt.base(kind);
    const uint32_t* __restrict__ seg_start = off && !block_rows ? tt.prefix(kind) : nullptr;

    for (uint32_t task = warp * gridDim.x + blockIdx.x; task < ntasks; task += nwarps) {
        const uint8_t* src;  // lane l's stream starts at src + l*S
        if (block_rows) {
            src = base + single_off + (size_t)task * 32 * lane_bytes;
        } else {
            uint32_t seg = task;
            uint64_t o = single_off, n = single_len;
            if (seg_start) {
                const uint32_t obj = find_object2(seg_start, tt.nobj, task);
                seg = task - __ldg(seg_start + obj);
                o = __ldg(off + obj);
                n = __ldg(len + obj);
            }
            const ObjectCut cut = cut_object((uint64_t)(uintptr_t)base + o, n);
            const Seg sg = seg_of(cut.rows, plan, kind, seg);
            src = base + o + cut.head + sg.row0 * kRowBytes;
        }
        const uint8_t* csrc = src + (size_t)crow * lane_bytes + cchunk * 16;

        uint32_t r[64];
#pragma unroll
        for (int k = 0; k < 64; k++) r[k] = 0;
        uint32_t tprev = 0;

        uint32_t ibuf = 0;  // staging buffer the next issue() fills
        auto issue = [&](uint32_t st) {
            if (st < nstage) {
                const uint32_t buf = warp_stage + ibuf * kWarpStageBytes;
                const uint8_t* g = csrc + (size_t)st * LANE_BYTES;
#pragma unroll
                for (int k = 0; k < kCopyInstrs; k++) {
                    const uint32_t row = crow + kRowsPerInstr * k;
                    cp_async16(buf + row * LANE_BYTES + ((cchunk ^ (row & 7)) << 4),
                               g + (size_t)(kRowsPerInstr * k) * lane_bytes);
                }
            }
            cp_async_commit();  // always commit so that group counting stays uniform
            ibuf = ibuf + 1 == STAGES ? 0 : ibuf + 1;
        };
        __syncwarp();  // previous task's readers are done with the staging buffers
#pragma unroll
        for (int s0 = 0; s0 < STAGES - 1; s0++) issue(s0);
        uint32_t cbuf = 0;  // staging buffer being consumed
        if (LANE_BYTES >= 256) {
            for (uint32_t st = 0; st < nstage; st++) {
                issue(st + STAGES - 1);
                cp_async_wait<STAGES - 1>();  // stage st has landed (this lane's part)
                __syncwarp();                 // ... and everybody else's
                const uint32_t a = my_row + cbuf * kWarpStageBytes;
#pragma unroll 1
                for (int h = 0; h < LANE_BYTES / 256; h++) {
                    FOLD_STAGE(0, a + h * 256)
                    FOLD_STAGE(32, a + h * 256 + 128)
                }
                __syncwarp();                 // all lanes done reading before the buffer is refilled
                cbuf = cbuf + 1 == STAGES ? 0 : cbuf + 1;
            }
        } else {
            for (uint32_t st = 0; st < nstage; st += 2) {
                issue(st + STAGES - 1);
                cp_async_wait<STAGES - 1>();
                __syncwarp();
                FOLD_STAGE(0, my_row + cbuf * kWarpStageBytes)
                __syncwarp();
                cbuf = cbuf + 1 == STAGES ? 0 : cbuf + 1;
                issue(st + STAGES);
                cp_async_wait<STAGES - 1>();
                __syncwarp();
                FOLD_STAGE(32, my_row + cbuf * kWarpStageBytes)
                __syncwarp();
                cbuf = cbuf + 1 == STAGES ? 0 : cbuf + 1;
            }
        }
        // the low part of the last leaving word has not been applied yet (it waits for a partner that never comes)
        r[54] ^= tprev >> 2;
        r[37] ^= tprev >> 2;
        r[26] ^= tprev >> 2;
        // reduce the 64-word state modulo P: feed words y^63 .. y^0 through the register, 4 bytes at a time
        uint32_t reg = 0;
#pragma unroll
        for (int k = 63; k >= 0; k--) {
            const uint32_t x = reg ^ r[k];
            reg = tab[x & 0xFF] ^ tab[256 + ((x >> 8) & 0xFF)] ^ tab[512 + ((x >> 16) & 0xFF)] ^ tab[768 + (x >> 24)];
        }
        if (block_rows) {
            const uint32_t obj = 32 * task + lane, crc = ~(block_init ^ reg);
            bo.got[obj] = crc;
            const bool good = bo.want ? crc == __ldg(bo.want + obj) : true;
            bo.ok[obj] = good ? 1 : 0;
            if (!good) atomicAdd(bo.nbad, 1ull);
            continue;
        }
        uint32_t s = gf2_mul(lane_k, reg);
#pragma unroll
        for (int d = 16; d; d >>= 1) s ^= __shfl_xor_sync(0xFFFFFFFFu, s, d);
        if (lane == 0) seg_raw[raw_base + task] = s;
    }
}

// ---- ragged small objects: a lane per object, any offset, any length ------------------------------------------
// The row kernel pays a table look-up per byte and ~450 instructions of set-up and segment end per object, which is
// what a million objects of a few KB cost; here lane l of a warp-task folds object 32 * task + l as a stream of its
// own.  An object is cut into a head (to the next 16-byte boundary), a run of whole 16-byte words — the stream — and
// a tail of fewer than 16 bytes; the kernel leaves the raw register of the stream in raw[object] and
// crc_finalize_objects_kernel (crc_kernels.cu) adds head, tail, initial value and verdict, a thread per object.
// Lanes of a warp have streams of different lengths, and a stream need not be whole stages: every stream is
// therefore RIGHT-aligned to the longest one of its warp-task by zeros in front — zero words entering a zero
// register leave it zero, so the padding costs cycles but no correction — which keeps the loop free of branches
// (all lanes run the stages of the longest stream; objects sorted by length would waste none).  The zeros are never
// read from memory: the copy is a cp.async with a source size of 0 for every 16-byte chunk before the stream's start.
template <int WARPS, int STAGES>
__global__ void __launch_bounds__(WARPS * 32, 1)
crc_fold_objects_kernel(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off, const uint64_t* __restrict__ len,
                        uint32_t nobj, const uint32_t* __restrict__ order, const uint32_t* __restrict__ adv4_tab,
                        uint32_t* __restrict__ raw) {
    // order (may be null): the objects sorted by the length of their streams in stages, so that the 32 lanes of a
    // task have streams of (nearly) the same length and none of them idles through another's stages
    constexpr int LANE_BYTES = 256, kThreads = WARPS * 32, kWarpStageBytes = 32 * LANE_BYTES;
    constexpr int kLanesPerRow = LANE_BYTES / 16, kRowsPerInstr = 32 / kLanesPerRow, kCopyInstrs = 32 / kRowsPerInstr;
    struct RowInfo {
        const uint8_t* vbase;  // address of the stream's virtual byte 0 (its real start minus the zeros in front)
        uint32_t zeros;        // virtual bytes in front of the real start (a multiple of 16)
        uint32_t pad;
    };
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* tab = reinterpret_cast<uint32_t*>(fold_smem);  // [4][256]
    RowInfo* info = reinterpret_cast<RowInfo*>(fold_smem + kFoldSmemTab) + warp * 32;
    uint8_t* stage_base = fold_smem + kFoldSmemTab + WARPS * 32 * sizeof(RowInfo);
    for (uint32_t i = threadIdx.x; i < 1024; i += kThreads) tab[i] = __ldg(adv4_tab + i);
    __syncthreads();
    const uint32_t warp_stage = smem_u32(stage_base + (size_t)warp * STAGES * kWarpStageBytes);
    const uint32_t my_row = warp_stage + lane * LANE_BYTES;
    const uint32_t sw = (lane & 7) << 4;
    const uint32_t crow = lane / kLanesPerRow, cchunk = lane % kLanesPerRow;
    const uint32_t nwarps = gridDim.x * WARPS, ntasks = (nobj + 31) / 32;

    // an object's number, offset and length are two dependent global loads (~2 us under load): the next task's are
    // fetched while this one is folded
    auto fetch = [&](uint32_t task, uint32_t& obj, uint64_t& o, uint64_t& n) {
        const uint32_t slot = 32 * task + lane;
        obj = task < ntasks && slot < nobj ? (order ? __ldg(order + slot) : slot) : nobj;
        o = n = 0;
        if (obj < nobj) {
            o = __ldg(off + obj);
            n = __ldg(len + obj);
        }
    };
    uint32_t obj_next;
    uint64_t o_next, n_next;
    fetch(warp * gridDim.x + blockIdx.x, obj_next, o_next, n_next);
    for (uint32_t task = warp * gridDim.x + blockIdx.x; task < ntasks; task += nwarps) {
        const uint32_t obj = obj_next;
        const uint64_t o = o_next, n = n_next;
        fetch(task + nwarps, obj_next, o_next, n_next);
        const uint8_t* p = base + o;
        const uint64_t h = (16 - (reinterpret_cast<uintptr_t>(p) & 15)) & 15;
        const uint64_t head = h < n ? h : n;
        const uint32_t m = (uint32_t)((n - head) & ~(uint64_t)15);  // the stream: whole 16-byte words (objects <= 4 GiB)
        uint32_t nstage = (m + LANE_BYTES - 1) / LANE_BYTES;
#pragma unroll
        for (int d = 16; d; d >>= 1) nstage = max(nstage, __shfl_xor_sync(0xFFFFFFFFu, nstage, d));  // the longest
        __syncwarp();  // the previous task's copies have read their row info
        info[lane].zeros = nstage * LANE_BYTES - m;
        info[lane].vbase = p + head - (size_t)(nstage * LANE_BYTES - m);
        __syncwarp();

        // the sixteen rows this lane copies for (its chunk of each), kept in registers for the whole task: where the
        // lane's chunk of the row's NEXT stage lies, and how many zero bytes the row still has in front of it
        const uint8_t* src_of[kCopyInstrs];
        uint32_t zeros_of[kCopyInstrs];
#pragma unroll
        for (int k = 0; k < kCopyInstrs; k++) {
            const RowInfo ri = info[crow + kRowsPerInstr * k];
            src_of[k] = ri.vbase + cchunk * 16;
            zeros_of[k] = ri.zeros;
        }
        uint32_t r[64];
#pragma unroll
        for (int k = 0; k < 64; k++) r[k] = 0;
        uint32_t tprev = 0;
        uint32_t ibuf = 0;
        auto issue = [&](uint32_t st) {
            if (st < nstage) {
                const uint32_t buf = warp_stage + ibuf * kWarpStageBytes;
                const uint32_t u = st * LANE_BYTES + cchunk * 16;  // virtual offset of this lane's chunk in every row
#pragma unroll
                for (int k = 0; k < kCopyInstrs; k++) {
                    const uint32_t row = crow + kRowsPerInstr * k;
                    const bool real = u >= zeros_of[k];
                    const uint32_t dst = buf + row * LANE_BYTES + ((cchunk ^ (row & 7)) << 4);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst),
                                 "l"(real ? src_of[k] : base), "r"(real ? 16u : 0u) : "memory");
                    src_of[k] += LANE_BYTES;
                }
            }
            cp_async_commit();
            ibuf = ibuf + 1 == STAGES ? 0 : ibuf + 1;
        };
#pragma unroll
        for (int s0 = 0; s0 < STAGES - 1; s0++) issue(s0);
        uint32_t cbuf = 0;
        for (uint32_t st = 0; st < nstage; st++) {
            issue(st + STAGES - 1);
            cp_async_wait<STAGES - 1>();
            __syncwarp();
            const uint32_t a = my_row + cbuf * kWarpStageBytes;
            FOLD_STAGE(0, a)
            FOLD_STAGE(32, a + 128)
            __syncwarp();
            cbuf = cbuf + 1 == STAGES ? 0 : cbuf + 1;
        }
        r[54] ^= tprev >> 2;
        r[37] ^= tprev >> 2;
        r[26] ^= tprev >> 2;
        uint32_t reg = 0;
#pragma unroll
        for (int k = 63; k >= 0; k--) {
            const uint32_t x = reg ^ r[k];
            reg = tab[x & 0xFF] ^ tab[256 + ((x >> 8) & 0xFF)] ^ tab[512 + ((x >> 16) & 0xFF)] ^ tab[768 + (x >> 24)];
        }
        if (obj < nobj) raw[obj] = reg;
    }
}

constexpr int kObjWarps = 8, kObjStages = 2;  // (12 warps were measured: 0.98 ms against 0.955 on the ragged 1 Mi shape)
constexpr size_t kObjSmem = kFoldSmemTab + kObjWarps * 32 * 16 + (size_t)kObjWarps * kObjStages * 32 * 256;

struct FoldVariant {
    int warps, stages, lane_bytes;
    void (*kernel)(const uint8_t*, const uint64_t*, const uint64_t*, TaskTable, SegPlan, uint32_t, uint64_t, uint64_t,
                   const uint32_t*, const uint32_t*, uint32_t*, uint32_t, uint32_t, BlockOut);
    size_t smem() const { return kFoldSmemTab + (size_t)warps * stages * 32 * lane_bytes; }
};
// Measured on B200 (4096 x 2 MiB, 300 back-to-back steps, GB/s): <8,2,256> 6878 | <7,2,256> 6796 | <4,3,512> 6780 |
// <6,3,256> 6758 | <9,2,256> 6658 | <8,3,256> 6568 | <12,2,256> 6499 | <4,2,256> 6139 | <16,3,128> 5563.
// 256-byte bursts per stream matter most (DRAM page locality); beyond ~64 KB in flight per SM more hurts.
const FoldVariant kVariants[] = {
    {8, 2, 256, crc_fold_kernel<8, 2, 256>},    // default
    {8, 3, 256, crc_fold_kernel<8, 3, 256>},
    {16, 3, 128, crc_fold_kernel<16, 3, 128>},
    {4, 3, 512, crc_fold_kernel<4, 3, 512>},
    {7, 2, 256, crc_fold_kernel<7, 2, 256>},
};

const FoldVariant& fold_variant() {
    static const int v = [] {  // tuning override, read once
        const char* e = getenv("B200_FOLD_VARIANT");
        const int x = e ? atoi(e) : 0;
        return (x < 0 || x >= (int)(sizeof kVariants / sizeof kVariants[0])) ? 0 : x;
    }();
    return kVariants[v];
}

}  // namespace

size_t fold_kernel_smem() { return fold_variant().smem(); }
uint32_t fold_lane_bytes() { return (uint32_t)fold_variant().lane_bytes; }

static cudaError_t fold_attr(const FoldVariant& fv) {
    static thread_local int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(fv.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fv.smem());
        if (e != cudaSuccess) return e;
        attr_dev = dev;
    }
    return cudaSuccess;
}

cudaError_t launch_fold(const BatchArgs& a, const DeviceTables& t, int grid, cudaStream_t stream, uint32_t kind) {
    const FoldVariant& fv = fold_variant();
    if (cudaError_t e = fold_attr(fv)) return e;
    const TaskTable tt{a.seg_start, a.nobj, {a.nfull_tasks, a.nshort_tasks, a.npartial_tasks}};
    if (!tt.count[kind]) return cudaSuccess;
    uint32_t need = (tt.count[kind] + fv.warps - 1) / fv.warps;
    int g = (int)(need < (uint32_t)grid ? need : (uint32_t)grid);
    fv.kernel<<<g, fv.warps * 32, fv.smem(), stream>>>(a.base, a.off, a.len, tt, SegPlan{a.rows_per_seg, a.rows_per_seg2},
                                                       kind, a.single_off, a.single_len, t.adv4, t.xpow8_sq, a.seg_raw,
                                                       0, 0, BlockOut{});
    return cudaGetLastError();
}

cudaError_t launch_fold_objects(const uint8_t* base, const uint64_t* off, const uint64_t* len, uint32_t nobj,
                                const uint32_t* order, const DeviceTables& t, uint32_t* raw, int grid, cudaStream_t stream) {
    static thread_local int attr_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (attr_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(crc_fold_objects_kernel<kObjWarps, kObjStages>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kObjSmem);
        if (e != cudaSuccess) return e;
        attr_dev = dev;
    }
    if (!nobj) return cudaSuccess;
    const uint32_t tasks = (nobj + 31) / 32, need = (tasks + kObjWarps - 1) / kObjWarps;
    const int g = (int)(need < (uint32_t)grid ? need : (uint32_t)grid);
    crc_fold_objects_kernel<kObjWarps, kObjStages><<<g, kObjWarps * 32, kObjSmem, stream>>>(base, off, len, nobj, order,
                                                                                            t.adv4, raw);
    return cudaGetLastError();
}

bool fold_lane_block_ok(uint32_t block_bytes) {
    // whole 512-byte rows (no tail for a finalize kernel to pick up), whole stages per lane, and long enough that the
    // 256 look-ups of the final reduction are a minor part (1 KiB: a quarter of the row kernel's look-ups)
    return fold_enabled(kFoldMinRows) && block_bytes >= 1024 && block_bytes <= 32768 && block_bytes % 512 == 0 &&
           block_bytes % fold_lane_bytes() == 0;
}

cudaError_t launch_fold_per_lane(const BatchArgs& a, const DeviceTables& t, int grid, cudaStream_t stream, uint32_t nobjs) {
    const FoldVariant& fv = fold_variant();
    if (cudaError_t e = fold_attr(fv)) return e;
    const uint32_t tasks = nobjs / 32;
    if (!tasks) return cudaSuccess;
    const TaskTable tt{a.seg_start, a.nobj, {a.nfull_tasks, a.nshort_tasks, a.npartial_tasks}};
    uint32_t need = (tasks + fv.warps - 1) / fv.warps;
    int g = (int)(need < (uint32_t)grid ? need : (uint32_t)grid);
    fv.kernel<<<g, fv.warps * 32, fv.smem(), stream>>>(a.base, nullptr, nullptr, tt, SegPlan{a.rows_per_seg, a.rows_per_seg2},
                                                       kSegPartial, a.uniform_off0, (uint64_t)nobjs * a.uniform_len, t.adv4,
                                                       t.xpow8_sq, a.seg_raw, a.uniform_len / 16, tasks,
                                                       BlockOut{a.want, a.got, a.ok, a.nbad, a.init_crc});
    return cudaGetLastError();
}

}  // namespace b200
