// Internal (C++) launch interface of the CRC-32C kernels.  The C ABI in api.cu is built on this.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gf2.h"

namespace b200 {

constexpr int kRowBytes = 512;    // one row = one coalesced 128-bit load per lane of a warp
constexpr int kRowWords = 128;
constexpr int kRowsThreads = 1024;
constexpr int kRowsWarps = kRowsThreads / 32;
constexpr int kFinalThreads = 128;
constexpr int kFoldWarps = 8;  // fold kernel (crc_fold_kernel.cu): default variant runs 8 warps per SM
constexpr uint32_t kFoldMinRows = 128;  // segments shorter than this stay on the row kernel

// How an object [addr, addr+len) is cut: an unaligned head (< 16 B) so the body starts on a 16-byte
// boundary, whole 512-byte rows, and a tail (< 512 B).  Rows go to the row kernel, head and tail are
// folded in by the finalize kernel.
struct ObjectCut {
    uint32_t head;
    uint32_t tail;
    uint64_t rows;
};
B200_HD ObjectCut cut_object(uint64_t addr, uint64_t len) {
    ObjectCut c;
    uint64_t h = (16 - (addr & 15)) & 15;
    c.head = (uint32_t)(h < len ? h : len);
    uint64_t body = len - c.head;
    c.rows = body / kRowBytes;
    c.tail = (uint32_t)(body % kRowBytes);
    return c;
}

// How the rows of an object body are cut into segments (= tasks).  Two lengths: as many LONG segments of
// `a` rows as fit, then as many SHORT segments of `b` rows as fit in what is left (b == 0: none), then one
// partial segment with the rest.  Long and short segments are "full" (the fold kernel can take them); the
// partial one always goes to the row kernel.  With b == 0 this is the plain one-length scheme.
enum SegKind : uint32_t { kSegLong = 0, kSegShort = 1, kSegPartial = 2 };
struct SegPlan {
    uint32_t a, b;
};
struct Seg {
    uint64_t row0;  // first row of the segment inside the body
    uint32_t rows;  // its length
    uint32_t kind;
};
// number of segments of each kind in a body of `rows` rows
B200_HD void seg_counts(uint64_t rows, SegPlan p, uint64_t n[3]) {
    n[kSegLong] = rows / p.a;
    const uint64_t rem = rows - n[kSegLong] * p.a;
    n[kSegShort] = p.b ? rem / p.b : 0;
    n[kSegPartial] = (rem - n[kSegShort] * p.b) ? 1 : 0;
}
// the j-th segment of kind `kind`
B200_HD Seg seg_of(uint64_t rows, SegPlan p, uint32_t kind, uint64_t j) {
    if (kind == kSegLong) return Seg{j * p.a, p.a, kSegLong};
    const uint64_t na = rows / p.a, rem = rows - na * p.a;
    if (kind == kSegShort) return Seg{na * p.a + j * p.b, p.b, kSegShort};
    const uint64_t nb = p.b ? rem / p.b : 0;
    return Seg{na * p.a + nb * p.b, (uint32_t)(rem - nb * p.b), kSegPartial};
}

// Tasks are numbered kind by kind: all long segments of the batch first, then the short ones, then the
// partial ones, so that a kernel that takes one kind can hand every warp the same number of them.  For kind
// k, start[k * (nobj + 1) + i] is the number of kind-k segments of objects 0..i-1 (a prefix sum with a final
// total); seg_raw[] is indexed by the global id base(k) + local id.  In single-object mode start is null and
// object 0 owns all count[k] segments of each kind.
struct TaskTable {
    const uint32_t* start;
    uint32_t nobj;
    uint32_t count[3];
    B200_HD uint32_t base(uint32_t kind) const {
        return kind == kSegLong ? 0 : kind == kSegShort ? count[0] : count[0] + count[1];
    }
    B200_HD const uint32_t* prefix(uint32_t kind) const { return start + (size_t)kind * (nobj + 1); }
};

// Constant tables a device needs (built once per context on the host, see tables.cu).
struct DeviceTables {
    uint32_t* adv_row;   // [4][256] multiply-by-x^(8*512) slicing tables (compact; kernels replicate per bank)
    uint32_t* col_k;     // [128] x^(8*4*(128-c)): weight of column c inside a row
    uint32_t* xpow8_sq;  // [64]  x^(8*2^k)
    uint32_t* byte_tab;  // [256] classic one-byte table
    uint32_t* adv4;      // [4][256] multiply-by-x^32 slicing tables (fold kernel's final reduction)
    uint32_t* xpow128;   // [kXpow128Entries] x^(8*16*j): the shift over a stream of j 16-byte words (objects up to
                         // 32 KiB, crc_finalize_objects_kernel); null where the lane-per-object path is not offered
};
constexpr uint32_t kXpow128Entries = 32768 / 16 + 1;

struct BatchArgs {
    const uint8_t* base;        // device arena, 16-byte aligned
    const uint64_t* off;        // [nobj] device
    const uint64_t* len;        // [nobj] device
    const uint32_t* seg_start;  // [3][nobj+1] device: per-kind prefix sums of segment counts (see TaskTable)
    const uint32_t* want;       // [nobj] device or nullptr
    uint32_t* seg_raw;          // [ntasks] device scratch: raw CRC register of each segment
    uint32_t* got;              // [nobj] device
    uint8_t* ok;                // [nobj] device
    unsigned long long* nbad;   // device scalar (zeroed by the launcher)
    const uint32_t* adv_seg;    // [4][256] multiply-by-x^(8*512*rows_per_seg) tables (device)
    uint32_t nobj;
    uint32_t ntasks;
    uint32_t rows_per_seg;      // long segments
    uint32_t rows_per_seg2;     // short segments (0 = none)
    const uint32_t* adv_seg2;   // [4][256] multiply-by-x^(8*512*rows_per_seg2) tables (device), when rows_per_seg2
    uint32_t init_crc;          // crc32.Update's first argument, same for every object (0 = fresh checksum)
    const uint32_t* init_ptr;   // single-object mode: when set, the first argument is read from the device
                                // (may alias `got`: a running CRC that stays on the device between chunks)
    // single-object mode (off == nullptr): the object is base[single_off .. single_off+single_len) and owns
    // tasks 0..ntasks-1; no plan arrays are read.
    uint64_t single_off;
    uint64_t single_len;
    // how many of the ntasks segments are long / short / partial; the fold kernel takes the long and short
    // ones when it is in use for that length, the row kernel the rest
    uint32_t nfull_tasks;       // long
    uint32_t nshort_tasks;
    uint32_t npartial_tasks;
    // every object has at most a few full segments and there are many objects: finalize one object per THREAD
    uint32_t small_objects;
    uint32_t no_fold;           // keep every segment on the row kernel (the caller has no tables for the fold kernel)
    // Equal blocks back to back (object i = base[uniform_off0 + i * uniform_len ..), uniform_len a multiple of 512
    // between 1 and 32 KiB, every object one partial segment): the first uniform_objs objects (a multiple of 32) may
    // be hashed by the fold kernel with ONE LANE PER OBJECT when base + uniform_off0 is 16-byte aligned; 0 = not so.
    uint64_t uniform_off0;
    uint32_t uniform_len;
    uint32_t uniform_objs;
    // many small objects of any offset and length (all of them at most 32 KiB, seg_raw has room for one word per
    // object): the fold kernel may take them a lane per object, crc_finalize_objects_kernel finishes them
    uint32_t lane_objects;
    const uint32_t* lane_order;  // [nobj] device: the objects sorted by stream length (may be null)
};

// true when launch_batch will give full segments of this length to the fold kernel
bool fold_enabled(uint32_t rows_per_seg);

int rows_kernel_grid(int dev);
size_t rows_kernel_smem();
// Enqueue row kernel + finalize kernel on `stream`.  ev_* may be null; when given they bracket the row kernel.
// `side` (optional): a second stream plus two events; when both segment kernels are needed the row kernel
// runs there so that its CTAs fill the SMs the fold kernel's tail leaves idle.
struct SideStream {
    cudaStream_t stream;
    cudaEvent_t fork, join;
};
cudaError_t launch_batch(const BatchArgs& a, const DeviceTables& t, int grid, cudaStream_t stream,
                         cudaEvent_t ev_rows_begin, cudaEvent_t ev_rows_end, const SideStream* side = nullptr);

size_t fold_kernel_smem();
uint32_t fold_lane_bytes();  // bytes a lane consumes per stage in the variant in use; a lane's share of a segment must be a multiple
// kind: kSegLong or kSegShort — which of the two full-segment lengths this launch folds
cudaError_t launch_fold(const BatchArgs& a, const DeviceTables& t, int grid, cudaStream_t stream, uint32_t kind);
// The fold kernel with a lane per object: objects [0, nobjs) of a.uniform_* (nobjs a multiple of 32), finished on the
// spot (a.got / a.ok / a.nbad against a.want).  fold_lane_block_ok: block_bytes suits it.
// Ragged small objects, a lane per object (crc_fold_kernel.cu): raw[object] = raw register of the object's stream of
// whole 16-byte words (what lies between its head, up to the next 16-byte boundary, and its tail of < 16 bytes).
// order: nullptr, or the object numbers sorted by stream length (lanes of a task then finish together).
cudaError_t launch_fold_objects(const uint8_t* base, const uint64_t* off, const uint64_t* len, uint32_t nobj,
                                const uint32_t* order, const DeviceTables& t, uint32_t* raw, int grid, cudaStream_t stream);
bool fold_lane_block_ok(uint32_t block_bytes);
cudaError_t launch_fold_per_lane(const BatchArgs& a, const DeviceTables& t, int grid, cudaStream_t stream, uint32_t nobjs);

cudaError_t launch_fill_payload(uint64_t* dst, uint64_t first_word, size_t nwords, uint64_t seed, cudaStream_t s);
cudaError_t launch_fill_latency(int64_t* dst, uint64_t first, size_t n, uint64_t seed, cudaStream_t s);

// Host side table construction (tables.cu)
void host_build_adv_table(uint64_t nbytes, uint32_t out[4 * 256]);
void host_build_col_k(uint32_t out[128]);
void host_build_xpow8_sq(uint32_t out[64]);
void host_build_byte_tab(uint32_t out[256]);
void host_build_xpow128(uint32_t* out /* [kXpow128Entries] */);
uint32_t host_xpow8(uint64_t nbytes);


}  // namespace b200
