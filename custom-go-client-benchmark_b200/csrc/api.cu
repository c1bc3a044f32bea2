// C ABI of libb200crc.so (declared in include/b200crc.h).  Host-side plumbing only: contexts, plans,
// staging, streams.  Every message byte is hashed on the GPU; the only CRC arithmetic done here is
// GF(2) combination of finished per-shard / per-chunk CRCs (b200_crc32c_combine).
#include "b200crc.h"

#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "crc_kernels.h"
#include "lat_kernels.h"

using namespace b200;

namespace {

thread_local char g_err[512] = "";

int fail_cuda(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof g_err, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    cudaGetLastError();  // clear sticky-less error state
    return B200_ERR_CUDA;
}
int fail(int code, const char* what) {
    snprintf(g_err, sizeof g_err, "%s", what);
    return code;
}
#define CU(call)                                             \
    do {                                                     \
        cudaError_t e_ = (call);                             \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call);  \
    } while (0)

constexpr size_t kStageBytes = 32u << 20;   // staging slab for pageable host input
constexpr int kSingleSlots = 8192;          // result slots for chunked single-object work

constexpr size_t kSlabBytes = 256u << 20;  // device slab for VerifyObjects on host buffers

// Reusable plan + result buffers of one device slab (host mirrors are pinned so uploads are asynchronous).
struct HostSlab {
    size_t cap = 0, task_cap = 0;
    uint64_t *d_off = nullptr, *d_len = nullptr, *h_off = nullptr, *h_len = nullptr;
    uint32_t *d_seg = nullptr, *h_seg = nullptr, *d_want = nullptr, *h_want = nullptr;
    uint32_t *d_got = nullptr, *h_got = nullptr, *d_seg_raw = nullptr;
    uint8_t *d_ok = nullptr, *h_ok = nullptr;
    unsigned long long *d_nbad = nullptr, *h_nbad = nullptr;
    cudaEvent_t done = nullptr;
    bool busy = false;
    size_t first = 0, n = 0;
    void release();
    int reserve(size_t nobj);
    int reserve_tasks(size_t ntasks);
};

struct DevState {
    int dev = 0;
    int grid = 148;
    cudaStream_t stream = nullptr;       // compute
    cudaStream_t copy_stream = nullptr;  // H2D
    DeviceTables tabs{};
    std::mutex mu;                  // guards the streams and scratch of this device
    std::recursive_mutex stage_mu;  // guards the staging slabs (held across whole host-input calls)
    std::recursive_mutex select_mu; // one order-statistic query at a time per device (they share lat.d_compact);
                                    // recursive: the sampled select runs ordinary selects on its sample
    std::map<uint32_t, uint32_t*> adv_seg;  // rows_per_seg -> device [4][256]
    uint32_t* seg_raw = nullptr;            // scratch for single-object runs
    size_t seg_raw_cap = 0;
    uint32_t* d_got = nullptr;              // [kSingleSlots]
    uint8_t* d_ok = nullptr;                // [1]
    unsigned long long* d_nbad = nullptr;   // [1]
    uint8_t* d_stage[2] = {nullptr, nullptr};
    size_t d_stage_cap = 0;
    uint8_t* h_stage[2] = {nullptr, nullptr};
    uint32_t* h_got = nullptr;  // pinned [kSingleSlots]
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    cudaEvent_t ev_timer[2] = {nullptr, nullptr};
    SideStream side{};  // second compute stream: row kernel next to the fold kernel
    LatScratch lat{};
    unsigned long long* text_scratch = nullptr;  // b200_parse_latency_text: line ranks and start masks (grow-only)
    size_t text_scratch_cap = 0;                 // in 64-bit words
    HostSlab slab[2];
};

void HostSlab::release() {
    cudaFree(d_off); cudaFree(d_len); cudaFree(d_seg); cudaFree(d_want); cudaFree(d_got); cudaFree(d_ok);
    cudaFree(d_nbad); cudaFree(d_seg_raw);
    cudaFreeHost(h_off); cudaFreeHost(h_len); cudaFreeHost(h_seg); cudaFreeHost(h_want); cudaFreeHost(h_got);
    cudaFreeHost(h_ok); cudaFreeHost(h_nbad);
    if (done) cudaEventDestroy(done);
    *this = HostSlab();
}

int HostSlab::reserve(size_t nobj) {
    if (!done) CU(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    if (!d_nbad) {
        CU(cudaMalloc(&d_nbad, 8));
        CU(cudaHostAlloc(&h_nbad, 8, cudaHostAllocDefault));
    }
    if (nobj <= cap) return 0;
    size_t c = std::max<size_t>(nobj * 2, 1024);
    // free AND forget: if one of the allocations below fails, a later reserve() / release() must not free these again
    auto drop_dev = [](auto*& p) { cudaFree(p); p = nullptr; };
    auto drop_host = [](auto*& p) { cudaFreeHost(p); p = nullptr; };
    drop_dev(d_off); drop_dev(d_len); drop_dev(d_seg); drop_dev(d_want); drop_dev(d_got); drop_dev(d_ok);
    drop_host(h_off); drop_host(h_len); drop_host(h_seg); drop_host(h_want); drop_host(h_got); drop_host(h_ok);
    cap = 0;
    CU(cudaMalloc(&d_off, c * 8));
    CU(cudaMalloc(&d_len, c * 8));
    CU(cudaMalloc(&d_seg, 3 * (c + 1) * 4));
    CU(cudaMalloc(&d_want, c * 4));
    CU(cudaMalloc(&d_got, c * 4));
    CU(cudaMalloc(&d_ok, c));
    CU(cudaHostAlloc(&h_off, c * 8, cudaHostAllocDefault));
    CU(cudaHostAlloc(&h_len, c * 8, cudaHostAllocDefault));
    CU(cudaHostAlloc(&h_seg, 3 * (c + 1) * 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&h_want, c * 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&h_got, c * 4, cudaHostAllocDefault));
    CU(cudaHostAlloc(&h_ok, c, cudaHostAllocDefault));
    cap = c;
    return 0;
}

int HostSlab::reserve_tasks(size_t ntasks) {
    if (ntasks <= task_cap) return 0;
    size_t c = std::max<size_t>(ntasks * 2, 1 << 16);
    cudaFree(d_seg_raw);
    d_seg_raw = nullptr;
    task_cap = 0;
    CU(cudaMalloc(&d_seg_raw, c * 4));
    task_cap = c;
    return 0;
}

}  // namespace

struct b200_ctx {
    std::vector<DevState*> devs;
};

namespace {

DevState* get_dev(b200_ctx* ctx, int dev) {
    if (!ctx) return nullptr;
    for (DevState* d : ctx->devs)
        if (d->dev == dev) return d;
    return nullptr;
}

int upload(uint32_t** dst, const uint32_t* src, size_t n) {
    CU(cudaMalloc(dst, n * sizeof(uint32_t)));
    CU(cudaMemcpy(*dst, src, n * sizeof(uint32_t), cudaMemcpyHostToDevice));
    return 0;
}

int get_adv_seg(DevState* d, uint32_t rows_per_seg, const uint32_t** out) {
    auto it = d->adv_seg.find(rows_per_seg);
    if (it == d->adv_seg.end()) {
        uint32_t tab[1024];
        host_build_adv_table((uint64_t)rows_per_seg * kRowBytes, tab);
        uint32_t* dp = nullptr;
        if (int rc = upload(&dp, tab, 1024)) return rc;
        it = d->adv_seg.emplace(rows_per_seg, dp).first;
    }
    *out = it->second;
    return 0;
}

constexpr uint32_t kMaxRowsPerSeg = 2048;  // 1 MiB segments: 32 KB per lane in the fold kernel

uint32_t env_rows_per_seg() {  // test / tuning override, read once
    static const uint32_t v = [] {
        const char* e = getenv("B200_ROWS_PER_SEG");
        const int x = e ? atoi(e) : 0;
        return (uint32_t)(x > 0 ? x : 0);
    }();
    return v;
}

constexpr uint32_t kShortRows = 256;  // short segments: 128 KB, 4 KB per lane in the fold kernel
constexpr size_t kMaxBatchObjects = (size_t)1 << 29;  // task ids are 32-bit and an object has up to two non-long segments

// The segmentation of a batch (see SegPlan in crc_kernels.h) and its task counts.
struct HostPlan {
    SegPlan p{16, 0};
    uint64_t ntasks = 0, nlong = 0, nshort = 0, npartial = 0, max_long = 0;
};

// Long segments: the longest power of two (<= 2048 rows) that still gives every warp of the fold kernel
// (8 per SM) ~6, or at least 2, of them; what the long segments leave over in each object is cut into short
// 256-row segments, which the fold kernel also takes, so only < 256 rows per object reach the row kernel.
// Inputs too small for that use the row kernel's finer grain alone (32 warps per SM).
template <typename RowsOf>
HostPlan make_plan(size_t nobj, RowsOf rows_of, int grid) {
    HostPlan h;
    uint64_t total = 0;
    for (size_t i = 0; i < nobj; i++) total += rows_of(i);
    bool chosen = false;
    if (uint32_t e = env_rows_per_seg()) {
        h.p.a = e;
        h.p.b = (fold_enabled(e) && e > kShortRows && e % kShortRows == 0) ? kShortRows : 0;
        chosen = true;
    } else if (fold_enabled(kMaxRowsPerSeg)) {
        const uint64_t warps = (uint64_t)grid * kFoldWarps;
        for (int pass = 0; pass < 2 && !chosen; pass++) {
            const uint64_t want = warps * (pass == 0 ? 6 : 2);
            for (uint32_t r = kMaxRowsPerSeg; r >= kShortRows && !chosen; r /= 2) {
                uint64_t full = 0;
                for (size_t i = 0; i < nobj; i++) full += rows_of(i) / r;
                if (full >= want) {
                    h.p.a = r;
                    h.p.b = r > kShortRows ? kShortRows : 0;
                    chosen = true;
                }
            }
        }
    }
    if (!chosen) {
        uint64_t target = total / ((uint64_t)grid * kRowsWarps * 6);
        uint32_t r = 16;
        while ((uint64_t)r * 2 <= target && r < 1024) r *= 2;
        h.p.a = r;
        h.p.b = 0;
    }
    for (;;) {  // task ids are 32-bit
        h.ntasks = h.nlong = h.nshort = h.npartial = h.max_long = 0;
        for (size_t i = 0; i < nobj; i++) {
            const uint64_t rows = rows_of(i), na = rows / h.p.a, rem = rows - na * h.p.a;
            const uint64_t nb = h.p.b ? rem / h.p.b : 0;
            h.nlong += na;
            h.nshort += nb;
            h.npartial += (rem - nb * h.p.b) ? 1 : 0;
            h.max_long = std::max(h.max_long, na);
        }
        h.ntasks = h.nlong + h.nshort + h.npartial;
        // doubling the long length shrinks nlong only; a batch whose partial / short segments alone exceed 2^30
        // (more than 2^30 objects) is refused by the callers (kMaxBatchObjects), so this ends before p.a overflows
        if (h.ntasks <= (1u << 30) || h.p.a >= (1u << 30)) break;
        h.p.a *= 2;
    }
    return h;
}

// per-kind prefix sums of segment counts: out[k * (nobj + 1) + i] (TaskTable in crc_kernels.h)
template <typename RowsOf>
void fill_task_prefix(size_t nobj, RowsOf rows_of, SegPlan p, uint32_t* out) {
    uint64_t run[3] = {0, 0, 0};
    for (size_t i = 0; i <= nobj; i++) {
        for (int k = 0; k < 3; k++) out[(size_t)k * (nobj + 1) + i] = (uint32_t)run[k];
        if (i == nobj) break;
        uint64_t n[3];
        seg_counts(rows_of(i), p, n);
        for (int k = 0; k < 3; k++) run[k] += n[k];
    }
}

// fill the launch arguments that describe the plan; looks up (or builds) the shift tables of both lengths
int apply_plan(DevState* d, const HostPlan& h, BatchArgs* a);

int apply_plan(DevState* d, const HostPlan& h, BatchArgs* a) {
    a->rows_per_seg = h.p.a;
    a->rows_per_seg2 = h.p.b;
    a->ntasks = (uint32_t)h.ntasks;
    a->nfull_tasks = (uint32_t)h.nlong;
    a->nshort_tasks = (uint32_t)h.nshort;
    a->npartial_tasks = (uint32_t)h.npartial;
    if (int rc = get_adv_seg(d, h.p.a, &a->adv_seg)) return rc;
    a->adv_seg2 = a->adv_seg;
    if (h.p.b)
        if (int rc = get_adv_seg(d, h.p.b, &a->adv_seg2)) return rc;
    return 0;
}

int ensure_seg_raw(DevState* d, size_t ntasks) {
    if (ntasks <= d->seg_raw_cap) return 0;
    if (d->seg_raw) cudaFree(d->seg_raw);
    d->seg_raw = nullptr;
    d->seg_raw_cap = 0;
    size_t cap = std::max<size_t>(ntasks, 1 << 16);
    CU(cudaMalloc(&d->seg_raw, cap * sizeof(uint32_t)));
    d->seg_raw_cap = cap;
    return 0;
}

int ensure_stage(DevState* d, size_t bytes) {
    if (bytes <= d->d_stage_cap) return 0;
    for (int i = 0; i < 2; i++) {
        if (d->d_stage[i]) cudaFree(d->d_stage[i]);
        d->d_stage[i] = nullptr;
    }
    d->d_stage_cap = 0;
    for (int i = 0; i < 2; i++) CU(cudaMalloc(&d->d_stage[i], bytes));
    d->d_stage_cap = bytes;
    return 0;
}

// Enqueue CRC of one device buffer on d->stream; result lands in d->d_got[slot], or — when `running` is given —
// the CRC at *running (device) is continued over the buffer and written back there.  Caller holds d->mu.
int enqueue_single(DevState* d, const void* dptr, uint64_t n, uint32_t init_crc, int slot, uint32_t* running = nullptr) {
    uintptr_t addr = (uintptr_t)dptr;
    uintptr_t base = addr & ~(uintptr_t)15;
    BatchArgs a{};
    a.base = (const uint8_t*)base;
    a.single_off = addr - base;
    a.single_len = n;
    ObjectCut cut = cut_object(addr, n);
    const HostPlan h = make_plan(1, [&](size_t) { return cut.rows; }, d->grid);
    if (int rc = ensure_seg_raw(d, h.ntasks)) return rc;
    if (int rc = apply_plan(d, h, &a)) return rc;
    a.nobj = 1;
    a.seg_raw = d->seg_raw;
    a.got = running ? running : d->d_got + slot;
    a.init_ptr = running;
    a.ok = d->d_ok;
    a.nbad = d->d_nbad;
    a.init_crc = init_crc;
    CU(launch_batch(a, d->tabs, d->grid, d->stream, nullptr, nullptr, &d->side));
    return 0;
}

bool is_pinned(const void* p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost;
}

// CRC of n host bytes, chunked and double buffered: H2D of chunk k+1 overlaps the kernels of chunk k.
// Chunk CRCs (each a fresh checksum) are folded on the host with GF(2) combine.  Caller holds d->mu.
int stream_host_locked(DevState* d, uint32_t crc, const uint8_t* host, size_t n, size_t chunk, uint32_t* out) {
    if (n == 0) {
        *out = crc;
        return 0;
    }
    if (chunk == 0) chunk = kStageBytes;
    chunk = (chunk + 511) & ~(size_t)511;
    const bool pinned = is_pinned(host);
    if (!pinned && chunk > kStageBytes) chunk = kStageBytes;
    size_t nchunks = (n + chunk - 1) / chunk;
    while (nchunks > (size_t)kSingleSlots) {  // keep within the result slots: larger chunks (pinned only)
        if (!pinned) break;
        chunk *= 2;
        nchunks = (n + chunk - 1) / chunk;
    }
    if (int rc = ensure_stage(d, std::min(chunk, (n + 511) & ~(size_t)511))) return rc;
    size_t done_chunks = 0;
    uint32_t acc = crc;
    while (done_chunks < nchunks) {
        size_t batch = std::min(nchunks - done_chunks, (size_t)kSingleSlots);
        for (size_t k = 0; k < batch; k++) {
            size_t idx = done_chunks + k;
            int b = (int)(idx & 1);
            size_t o = idx * chunk;
            size_t len = std::min(chunk, n - o);
            if (idx >= 2) CU(cudaStreamWaitEvent(d->copy_stream, d->ev_done[b], 0));
            const uint8_t* src = host + o;
            if (!pinned) {
                if (idx >= 2) CU(cudaEventSynchronize(d->ev_copied[b]));  // pinned slab free again
                memcpy(d->h_stage[b], src, len);
                src = d->h_stage[b];
            }
            CU(cudaMemcpyAsync(d->d_stage[b], src, len, cudaMemcpyHostToDevice, d->copy_stream));
            CU(cudaEventRecord(d->ev_copied[b], d->copy_stream));
            CU(cudaStreamWaitEvent(d->stream, d->ev_copied[b], 0));
            if (int rc = enqueue_single(d, d->d_stage[b], len, 0, (int)k)) return rc;
            CU(cudaEventRecord(d->ev_done[b], d->stream));
        }
        CU(cudaMemcpyAsync(d->h_got, d->d_got, batch * sizeof(uint32_t), cudaMemcpyDeviceToHost, d->stream));
        CU(cudaStreamSynchronize(d->stream));
        for (size_t k = 0; k < batch; k++) {
            size_t idx = done_chunks + k;
            size_t len = std::min(chunk, n - idx * chunk);
            acc = b200_crc32c_combine(acc, d->h_got[k], len);
        }
        done_chunks += batch;
    }
    *out = acc;
    return 0;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------

extern "C" {

const char* b200_strerror(int status) {
    switch (status) {
        case B200_OK: return "ok";
        case B200_ERR_INVALID: return "invalid argument";
        case B200_ERR_CUDA: return "CUDA error";
        case B200_ERR_NOMEM: return "out of memory";
        case B200_ERR_EMPTY: return "empty sample set";
        case B200_ERR_TOO_LARGE: return "size beyond fixed capacity";
        default: return "unknown status";
    }
}

const char* b200_last_error(void) { return g_err; }

int b200_device_count(int* n) {
    if (!n) return fail(B200_ERR_INVALID, "n is null");
    CU(cudaGetDeviceCount(n));
    return 0;
}

int b200_device_pci_bus_id(int dev, char* buf, int len) {
    if (!buf || len < 16) return fail(B200_ERR_INVALID, "b200_device_pci_bus_id: buffer of at least 16 bytes");
    CU(cudaDeviceGetPCIBusId(buf, len, dev));
    return 0;
}

int b200_init(int ndev, const int* devs, b200_ctx** out) {
    if (!out || ndev < 1) return fail(B200_ERR_INVALID, "b200_init: ndev < 1 or out is null");
    int have = 0;
    CU(cudaGetDeviceCount(&have));
    b200_ctx* ctx = new (std::nothrow) b200_ctx;
    if (!ctx) return fail(B200_ERR_NOMEM, "b200_init");
    uint32_t adv_row[1024], col_k[128], xsq[64], byte_tab[256], adv4[1024];
    host_build_adv_table(kRowBytes, adv_row);
    host_build_adv_table(4, adv4);
    host_build_col_k(col_k);
    host_build_xpow8_sq(xsq);
    host_build_byte_tab(byte_tab);
    std::vector<uint32_t> xpow128(kXpow128Entries);
    host_build_xpow128(xpow128.data());
    for (int i = 0; i < ndev; i++) {
        int dev = devs ? devs[i] : i;
        if (dev < 0 || dev >= have) {
            b200_shutdown(ctx);
            return fail(B200_ERR_INVALID, "b200_init: no such device");
        }
        DevState* d = new (std::nothrow) DevState;
        if (!d) {
            b200_shutdown(ctx);
            return fail(B200_ERR_NOMEM, "b200_init");
        }
        d->dev = dev;
        ctx->devs.push_back(d);
        cudaError_t e = cudaSetDevice(dev);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d->side.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->side.fork, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->side.join, cudaEventDisableTiming);
        for (int b = 0; b < 2 && e == cudaSuccess; b++) {
            e = cudaEventCreateWithFlags(&d->ev_copied[b], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->ev_done[b], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaHostAlloc(&d->h_stage[b], kStageBytes, cudaHostAllocDefault);
            if (e == cudaSuccess) e = cudaEventCreate(&d->ev_timer[b]);
        }
        if (e == cudaSuccess) e = cudaHostAlloc(&d->h_got, kSingleSlots * sizeof(uint32_t), cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaMalloc(&d->d_got, kSingleSlots * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMalloc(&d->d_ok, 16);
        if (e == cudaSuccess) e = cudaMalloc(&d->d_nbad, sizeof(unsigned long long));
        if (e != cudaSuccess) {
            b200_shutdown(ctx);
            return fail_cuda(e, "b200_init: device setup");
        }
        d->grid = rows_kernel_grid(dev);
        int rc = upload(&d->tabs.adv_row, adv_row, 1024);
        if (!rc) rc = upload(&d->tabs.col_k, col_k, 128);
        if (!rc) rc = upload(&d->tabs.xpow8_sq, xsq, 64);
        if (!rc) rc = upload(&d->tabs.byte_tab, byte_tab, 256);
        if (!rc) rc = upload(&d->tabs.adv4, adv4, 1024);
        if (!rc) rc = upload(&d->tabs.xpow128, xpow128.data(), xpow128.size());
        if (!rc) {
            const cudaError_t e = lat_scratch_init(&d->lat);
            if (e == cudaErrorInvalidConfiguration)
                rc = fail(B200_ERR_CUDA, "this driver reserves more than 1 KB of shared memory per CTA: the kernels' "
                                         "absolute shared-memory maps do not fit");
            else if (e != cudaSuccess) rc = fail_cuda(e, "latency scratch");
        }
        if (rc) {
            b200_shutdown(ctx);
            return rc;
        }
    }
    *out = ctx;
    return 0;
}

int b200_shutdown(b200_ctx* ctx) {
    if (!ctx) return 0;
    for (DevState* d : ctx->devs) {
        cudaSetDevice(d->dev);
        if (d->stream) cudaStreamSynchronize(d->stream);
        if (d->copy_stream) cudaStreamSynchronize(d->copy_stream);
        for (auto& kv : d->adv_seg) cudaFree(kv.second);
        cudaFree(d->tabs.adv_row);
        cudaFree(d->tabs.col_k);
        cudaFree(d->tabs.xpow8_sq);
        cudaFree(d->tabs.byte_tab);
        cudaFree(d->tabs.adv4);
        cudaFree(d->tabs.xpow128);
        cudaFree(d->seg_raw);
        cudaFree(d->d_got);
        cudaFree(d->d_ok);
        cudaFree(d->d_nbad);
        cudaFreeHost(d->h_got);
        lat_scratch_free(&d->lat);
        cudaFree(d->text_scratch);
        d->slab[0].release();
        d->slab[1].release();
        for (int b = 0; b < 2; b++) {
            cudaFree(d->d_stage[b]);
            cudaFreeHost(d->h_stage[b]);
            if (d->ev_copied[b]) cudaEventDestroy(d->ev_copied[b]);
            if (d->ev_done[b]) cudaEventDestroy(d->ev_done[b]);
            if (d->ev_timer[b]) cudaEventDestroy(d->ev_timer[b]);
        }
        if (d->side.stream) cudaStreamDestroy(d->side.stream);
        if (d->side.fork) cudaEventDestroy(d->side.fork);
        if (d->side.join) cudaEventDestroy(d->side.join);
        if (d->stream) cudaStreamDestroy(d->stream);
        if (d->copy_stream) cudaStreamDestroy(d->copy_stream);
        delete d;
    }
    cudaGetLastError();
    delete ctx;
    return 0;
}

int b200_ctx_ndev(const b200_ctx* ctx) { return ctx ? (int)ctx->devs.size() : 0; }

int b200_timer_start(b200_ctx* ctx, int dev) {
    DevState* d = get_dev(ctx, dev);
    if (!d) return fail(B200_ERR_INVALID, "b200_timer_start");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    CU(cudaStreamSynchronize(d->stream));
    CU(cudaEventRecord(d->ev_timer[0], d->stream));
    return 0;
}

int b200_timer_stop(b200_ctx* ctx, int dev, float* ms) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !ms) return fail(B200_ERR_INVALID, "b200_timer_stop");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    CU(cudaEventRecord(d->ev_timer[1], d->stream));
    CU(cudaEventSynchronize(d->ev_timer[1]));
    CU(cudaEventElapsedTime(ms, d->ev_timer[0], d->ev_timer[1]));
    return 0;
}

int b200_alloc_pinned(b200_ctx* ctx, size_t n, void** host) {
    if (!ctx || !host) return fail(B200_ERR_INVALID, "b200_alloc_pinned");
    CU(cudaSetDevice(ctx->devs[0]->dev));
    cudaError_t e = cudaHostAlloc(host, n ? n : 1, cudaHostAllocPortable);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return fail(B200_ERR_NOMEM, "cudaHostAlloc");
    }
    CU(e);
    return 0;
}

int b200_free_pinned(b200_ctx* ctx, void* host) {
    if (!ctx) return fail(B200_ERR_INVALID, "b200_free_pinned");
    if (host) CU(cudaFreeHost(host));
    return 0;
}

int b200_alloc_device(b200_ctx* ctx, int dev, size_t n, void** dptr) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !dptr) return fail(B200_ERR_INVALID, "b200_alloc_device: unknown device or null out");
    CU(cudaSetDevice(dev));
    cudaError_t e = cudaMalloc(dptr, n ? n : 1);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return fail(B200_ERR_NOMEM, "cudaMalloc");
    }
    CU(e);
    return 0;
}

int b200_free_device(b200_ctx* ctx, int dev, void* dptr) {
    DevState* d = get_dev(ctx, dev);
    if (!d) return fail(B200_ERR_INVALID, "b200_free_device: unknown device");
    CU(cudaSetDevice(dev));
    if (dptr) CU(cudaFree(dptr));
    return 0;
}

int b200_memcpy_h2d(b200_ctx* ctx, int dev, void* dptr, const void* host, size_t n) {
    DevState* d = get_dev(ctx, dev);
    if (!d || (n && (!dptr || !host))) return fail(B200_ERR_INVALID, "b200_memcpy_h2d");
    CU(cudaSetDevice(dev));
    // on the device's own stream and waited for: a plain cudaMemcpy from pageable memory may return with the last
    // piece still on its way, and the kernels' non-blocking stream would not wait for it
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaMemcpyAsync(dptr, host, n, cudaMemcpyHostToDevice, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

int b200_memcpy_d2h(b200_ctx* ctx, int dev, void* host, const void* dptr, size_t n) {
    DevState* d = get_dev(ctx, dev);
    if (!d || (n && (!dptr || !host))) return fail(B200_ERR_INVALID, "b200_memcpy_d2h");
    CU(cudaSetDevice(dev));
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaMemcpyAsync(host, dptr, n, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

uint32_t b200_crc32c_combine(uint32_t crc1, uint32_t crc2, uint64_t len2) {
    return gf2_mul(host_xpow8(len2), crc1) ^ crc2;
}

int b200_crc32c_update_device(b200_ctx* ctx, int dev, uint32_t crc, const void* dptr, size_t n, uint32_t* out) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !out || (n && !dptr)) return fail(B200_ERR_INVALID, "b200_crc32c_update_device");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    if (int rc = enqueue_single(d, dptr, n, crc, 0)) return rc;
    CU(cudaMemcpyAsync(d->h_got, d->d_got, sizeof(uint32_t), cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    *out = d->h_got[0];
    return 0;
}

int b200_crc32c_stream_host(b200_ctx* ctx, int dev, uint32_t crc, const void* host, size_t n, size_t chunk,
                            uint32_t* out) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !out || (n && !host)) return fail(B200_ERR_INVALID, "b200_crc32c_stream_host");
    std::lock_guard<std::recursive_mutex> slk(d->stage_mu);
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    return stream_host_locked(d, crc, (const uint8_t*)host, n, chunk, out);
}

int b200_crc32c_update_host(b200_ctx* ctx, int dev, uint32_t crc, const void* p, size_t n, uint32_t* out) {
    return b200_crc32c_stream_host(ctx, dev, crc, p, n, 0, out);
}

int b200_crc32c_sharded(b200_ctx* ctx, const void* const* dptr_per_dev, const uint64_t* len_per_dev, int ndev,
                        uint32_t* out) {
    if (!ctx || !dptr_per_dev || !len_per_dev || !out || ndev < 1 || ndev > (int)ctx->devs.size())
        return fail(B200_ERR_INVALID, "b200_crc32c_sharded");
    // enqueue on every device first, then collect: the shards run concurrently
    for (int i = 0; i < ndev; i++) {
        DevState* d = ctx->devs[i];
        d->mu.lock();
    }
    int rc = 0;
    for (int i = 0; i < ndev && !rc; i++) {
        DevState* d = ctx->devs[i];
        cudaError_t e = cudaSetDevice(d->dev);
        if (e != cudaSuccess) { rc = fail_cuda(e, "cudaSetDevice"); break; }
        rc = enqueue_single(d, dptr_per_dev[i], len_per_dev[i], 0, 0);
        if (!rc) {
            e = cudaMemcpyAsync(d->h_got, d->d_got, sizeof(uint32_t), cudaMemcpyDeviceToHost, d->stream);
            if (e != cudaSuccess) rc = fail_cuda(e, "cudaMemcpyAsync");
        }
    }
    uint32_t acc = 0;
    for (int i = 0; i < ndev; i++) {
        DevState* d = ctx->devs[i];
        if (!rc) {
            cudaSetDevice(d->dev);
            cudaError_t e = cudaStreamSynchronize(d->stream);
            if (e != cudaSuccess) rc = fail_cuda(e, "cudaStreamSynchronize");
            else acc = b200_crc32c_combine(acc, d->h_got[0], len_per_dev[i]);
        }
    }
    for (int i = 0; i < ndev; i++) ctx->devs[i]->mu.unlock();
    if (!rc) *out = acc;
    return rc;
}

int b200_fill_payload_device(b200_ctx* ctx, int dev, void* dptr, uint64_t first_word, size_t nwords, uint64_t seed) {
    DevState* d = get_dev(ctx, dev);
    if (!d || (nwords && !dptr) || ((uintptr_t)dptr & 7)) return fail(B200_ERR_INVALID, "b200_fill_payload_device");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    CU(launch_fill_payload((uint64_t*)dptr, first_word, nwords, seed, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

int b200_fill_latency_ns_device(b200_ctx* ctx, int dev, int64_t* dptr, uint64_t first, size_t n, uint64_t seed) {
    DevState* d = get_dev(ctx, dev);
    if (!d || (n && !dptr)) return fail(B200_ERR_INVALID, "b200_fill_latency_ns_device");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    CU(launch_fill_latency(dptr, first, n, seed, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

}  // extern "C"

// ---- prepared batches -------------------------------------------------------------------------------

struct b200_batch {
    DevState* d = nullptr;
    size_t nobj = 0;
    uint64_t* d_off = nullptr;
    uint64_t* d_len = nullptr;
    uint32_t* d_seg_start = nullptr;
    uint32_t* d_want = nullptr;
    uint32_t* d_seg_raw = nullptr;
    uint32_t* d_got = nullptr;
    uint8_t* d_ok = nullptr;
    unsigned long long* d_nbad = nullptr;
    HostPlan plan;
    uint32_t small_objects = 0;
    uint64_t uniform_off0 = 0;  // equal blocks back to back (BatchArgs::uniform_*), found at creation
    uint32_t uniform_len = 0, uniform_objs = 0;
    uint32_t lane_objects = 0;  // many small objects, none above 32 KiB (BatchArgs::lane_objects)
    uint32_t* d_order = nullptr;  // ... sorted by the length of their streams (BatchArgs::lane_order)
    static constexpr int kRing = 256;  // per-run CUDA event triples: begin, row kernel end, all end
    cudaEvent_t ev[kRing][3] = {};
    uint64_t runs = 0, timed_from = 0;
    bool ran = false;
};

extern "C" {

int b200_batch_free(b200_batch* b) {
    if (!b) return 0;
    if (b->d) cudaSetDevice(b->d->dev);
    cudaFree(b->d_off);
    cudaFree(b->d_len);
    cudaFree(b->d_seg_start);
    cudaFree(b->d_want);
    cudaFree(b->d_seg_raw);
    cudaFree(b->d_order);
    cudaFree(b->d_got);
    cudaFree(b->d_ok);
    cudaFree(b->d_nbad);
    for (auto& t : b->ev)
        for (cudaEvent_t e : t)
            if (e) cudaEventDestroy(e);
    cudaGetLastError();
    delete b;
    return 0;
}

int b200_batch_create(b200_ctx* ctx, int dev, const uint64_t* off, const uint64_t* len, size_t nobj,
                      const uint32_t* want, b200_batch** out) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !out || (nobj && (!off || !len))) return fail(B200_ERR_INVALID, "b200_batch_create");
    if (nobj > kMaxBatchObjects) return fail(B200_ERR_TOO_LARGE, "b200_batch_create: more than 2^29 objects in one batch");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    b200_batch* b = new (std::nothrow) b200_batch;
    if (!b) return fail(B200_ERR_NOMEM, "b200_batch_create");
    b->d = d;
    b->nobj = nobj;
    // plan: the arena base is 16-byte aligned, so the cut of object i depends on off[i] only
    auto rows_of = [&](size_t i) { return cut_object(off[i], len[i]).rows; };
    b->plan = make_plan(nobj, rows_of, d->grid);
    std::vector<uint32_t> seg_start(3 * (nobj + 1));
    fill_task_prefix(nobj, rows_of, b->plan.p, seg_start.data());
    b->small_objects = (nobj >= 1024 && b->plan.max_long <= 16) ? 1 : 0;
    // equal blocks back to back (block-wise verification: pages, ssd_test's blocks)?  Then the fold kernel can take
    // them a lane per object (launch_batch decides, once it knows the arena's alignment).
    if (nobj >= 64 && len[0] >= 1024 && len[0] <= 32768 && len[0] % 512 == 0 && b->plan.nlong == 0 && b->plan.nshort == 0 &&
        b->plan.npartial == nobj) {
        bool uniform = true;
        for (size_t i = 0; i < nobj && uniform; i++) uniform = len[i] == len[0] && off[i] == off[0] + i * len[0];
        if (uniform) {
            b->uniform_off0 = off[0];
            b->uniform_len = (uint32_t)len[0];
            b->uniform_objs = (uint32_t)(nobj / 32 * 32);
        }
    }
    int rc = 0;
    auto dev_alloc = [&](void** p, size_t bytes) -> int {
        if (rc) return rc;
        cudaError_t e = cudaMalloc(p, bytes ? bytes : 16);
        if (e != cudaSuccess) rc = e == cudaErrorMemoryAllocation ? fail(B200_ERR_NOMEM, "cudaMalloc")
                                                                  : fail_cuda(e, "cudaMalloc");
        return rc;
    };
    dev_alloc((void**)&b->d_off, nobj * 8);
    dev_alloc((void**)&b->d_len, nobj * 8);
    dev_alloc((void**)&b->d_seg_start, 3 * (nobj + 1) * 4);
    if (want) dev_alloc((void**)&b->d_want, nobj * 4);
    // many small objects of any shape, none above 32 KiB: candidates for the fold kernel's lane-per-object form, which
    // leaves one word per object in seg_raw
    if (b->small_objects && !b->uniform_objs) {
        uint64_t longest = 0;
        for (size_t i = 0; i < nobj; i++) longest = std::max<uint64_t>(longest, len[i]);
        b->lane_objects = longest <= 32768 ? 1 : 0;
    }
    dev_alloc((void**)&b->d_seg_raw, std::max<size_t>(b->plan.ntasks, b->lane_objects ? nobj : 0) * 4);
    if (b->lane_objects) {
        // counting sort by stream length in 256-byte stages (the arena is 16-byte aligned, so an object's head follows
        // from its offset): the 32 lanes of a task then run the same number of stages
        constexpr uint32_t kBuckets = 32768 / 256 + 2;
        std::vector<uint32_t> start(kBuckets + 1, 0), order(nobj);
        auto stages_of = [&](size_t i) {
            const uint64_t h = (16 - (off[i] & 15)) & 15, head = h < len[i] ? h : len[i];
            return (uint32_t)((((len[i] - head) & ~(uint64_t)15) + 255) / 256);
        };
        for (size_t i = 0; i < nobj; i++) start[stages_of(i) + 1]++;
        for (uint32_t k = 0; k < kBuckets; k++) start[k + 1] += start[k];
        for (size_t i = 0; i < nobj; i++) order[start[stages_of(i)]++] = (uint32_t)i;
        dev_alloc((void**)&b->d_order, nobj * 4);
        if (!rc && cudaMemcpy(b->d_order, order.data(), nobj * 4, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = fail(B200_ERR_CUDA, "upload order");
    }
    dev_alloc((void**)&b->d_got, nobj * 4);
    dev_alloc((void**)&b->d_ok, nobj);
    dev_alloc((void**)&b->d_nbad, 8);
    auto cu = [&](cudaError_t e, const char* what) {
        if (!rc && e != cudaSuccess) rc = fail_cuda(e, what);
    };
    if (!rc && nobj) {
        cu(cudaMemcpy(b->d_off, off, nobj * 8, cudaMemcpyHostToDevice), "upload off");
        cu(cudaMemcpy(b->d_len, len, nobj * 8, cudaMemcpyHostToDevice), "upload len");
        if (want) cu(cudaMemcpy(b->d_want, want, nobj * 4, cudaMemcpyHostToDevice), "upload want");
    }
    if (!rc) cu(cudaMemcpy(b->d_seg_start, seg_start.data(), 3 * (nobj + 1) * 4, cudaMemcpyHostToDevice), "upload plan");
    for (auto& t : b->ev)
        for (cudaEvent_t& e : t)
            if (!rc) cu(cudaEventCreate(&e), "event");
    if (rc) {
        b200_batch_free(b);
        return rc;
    }
    *out = b;
    return 0;
}

int b200_batch_run_async(b200_batch* b, const void* base) {
    if (!b || (b->nobj && !base) || ((uintptr_t)base & 15))
        return fail(B200_ERR_INVALID, "b200_batch_run: arena base must be a 16-byte aligned device pointer");
    DevState* d = b->d;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(d->dev));
    BatchArgs a{};
    a.base = (const uint8_t*)base;
    a.off = b->d_off;
    a.len = b->d_len;
    a.seg_start = b->d_seg_start;
    a.want = b->d_want;
    a.seg_raw = b->d_seg_raw;
    a.got = b->d_got;
    a.ok = b->d_ok;
    a.nbad = b->d_nbad;
    if (int rc = apply_plan(d, b->plan, &a)) return rc;
    a.nobj = (uint32_t)b->nobj;
    a.small_objects = b->small_objects;
    a.uniform_off0 = b->uniform_off0;
    a.uniform_len = b->uniform_len;
    a.uniform_objs = b->uniform_objs;
    a.lane_objects = b->lane_objects;
    a.lane_order = b->d_order;
    a.init_crc = 0;
    cudaEvent_t* ev = b->ev[b->runs % b200_batch::kRing];
    CU(launch_batch(a, d->tabs, d->grid, d->stream, ev[0], ev[1], &d->side));
    CU(cudaEventRecord(ev[2], d->stream));
    b->runs++;
    b->ran = true;
    return 0;
}

int b200_batch_wait(b200_batch* b) {
    if (!b) return fail(B200_ERR_INVALID, "b200_batch_wait");
    CU(cudaSetDevice(b->d->dev));
    CU(cudaStreamSynchronize(b->d->stream));
    return 0;
}

int b200_batch_run(b200_batch* b, const void* base) {
    if (int rc = b200_batch_run_async(b, base)) return rc;
    return b200_batch_wait(b);
}

int b200_batch_results(b200_batch* b, uint32_t* got, uint8_t* ok, uint64_t* nbad) {
    if (!b || !b->ran) return fail(B200_ERR_INVALID, "b200_batch_results: batch has not run");
    CU(cudaSetDevice(b->d->dev));
    CU(cudaStreamSynchronize(b->d->stream));
    if (got && b->nobj) CU(cudaMemcpy(got, b->d_got, b->nobj * 4, cudaMemcpyDeviceToHost));
    if (ok && b->nobj) CU(cudaMemcpy(ok, b->d_ok, b->nobj, cudaMemcpyDeviceToHost));
    if (nbad) {
        unsigned long long v = 0;
        CU(cudaMemcpy(&v, b->d_nbad, 8, cudaMemcpyDeviceToHost));
        *nbad = v;
    }
    return 0;
}

int b200_batch_kernel_ms(b200_batch* b, float* total_ms, float* rows_kernel_ms) {
    if (!b || !b->ran) return fail(B200_ERR_INVALID, "b200_batch_kernel_ms: batch has not run");
    CU(cudaSetDevice(b->d->dev));
    CU(cudaStreamSynchronize(b->d->stream));
    uint64_t from = b->timed_from;
    if (b->runs - from > b200_batch::kRing) from = b->runs - b200_batch::kRing;
    double tot = 0, rows = 0;
    for (uint64_t r = from; r < b->runs; r++) {
        cudaEvent_t* ev = b->ev[r % b200_batch::kRing];
        float t = 0, k = 0;
        CU(cudaEventElapsedTime(&t, ev[0], ev[2]));
        CU(cudaEventElapsedTime(&k, ev[0], ev[1]));
        tot += t;
        rows += k;
    }
    uint64_t n = b->runs - from;
    if (total_ms) *total_ms = n ? (float)(tot / n) : 0.f;
    if (rows_kernel_ms) *rows_kernel_ms = n ? (float)(rows / n) : 0.f;
    b->timed_from = b->runs;
    return 0;
}

int b200_batch_launches(const b200_batch* b, int* launches_per_run) {
    if (!b || !launches_per_run) return fail(B200_ERR_INVALID, "b200_batch_launches");
    const HostPlan& h = b->plan;  // mirrors launch_batch
    const bool fold_long = fold_enabled(h.p.a) && h.nlong > 0;
    const bool fold_short = h.p.b && fold_enabled(h.p.b) && h.nshort > 0;
    // equal blocks on the fold kernel, a lane per object, finished there (taken when the arena is 16-byte aligned
    // at run time, which the count assumes)
    if (b->lane_objects && fold_enabled(kFoldMinRows)) {  // ragged small objects: one fold launch + one finalize
        *launches_per_run = 2;
        return 0;
    }
    const uint64_t lane_objs = b->small_objects && b->uniform_objs && fold_lane_block_ok(b->uniform_len) ? b->uniform_objs : 0;
    const bool rows = h.ntasks && (h.npartial > lane_objs || (!fold_long && h.nlong > 0) || (!fold_short && h.nshort > 0));
    const bool finalize = b->nobj > lane_objs;
    *launches_per_run = (fold_long ? 1 : 0) + (fold_short ? 1 : 0) + (lane_objs ? 1 : 0) + (rows ? 1 : 0) + (finalize ? 1 : 0);
    return 0;
}

int b200_crc32c_batch_device(b200_ctx* ctx, int dev, const void* base, const uint64_t* off, const uint64_t* len,
                             size_t nobj, const uint32_t* want, uint32_t* got, uint8_t* ok, uint64_t* nbad) {
    b200_batch* b = nullptr;
    if (int rc = b200_batch_create(ctx, dev, off, len, nobj, want, &b)) return rc;
    int rc = b200_batch_run(b, base);
    if (!rc) rc = b200_batch_results(b, got, ok, nbad);
    b200_batch_free(b);
    return rc;
}

int b200_crc32c_batch_host(b200_ctx* ctx, int dev, const void* const* bufs, const uint64_t* len, size_t nobj,
                           const uint32_t* want, uint32_t* got, uint8_t* ok, uint64_t* nbad) {
    DevState* d = get_dev(ctx, dev);
    if (!d || (nobj && (!bufs || !len))) return fail(B200_ERR_INVALID, "b200_crc32c_batch_host");
    // Objects are laid out back to back (16-byte aligned starts) in one of two device slabs.  Pinned
    // buffers (b200_alloc_pinned) are copied straight from where they lie — neighbouring buffers merge into
    // one cudaMemcpyAsync — pageable ones go through the pinned staging slab.  While slab k is verified on
    // the compute stream, slab k+1 is being filled by the copy engine.  An object larger than a slab is
    // streamed on its own.
    std::lock_guard<std::recursive_mutex> slk(d->stage_mu);
    {
        std::lock_guard<std::mutex> lk(d->mu);
        CU(cudaSetDevice(dev));
        if (int r = ensure_stage(d, kSlabBytes)) return r;
    }
    uint64_t bad_total = 0;
    int rc = 0;
    size_t i = 0;
    int slot = 0;
    auto collect = [&](int s) -> int {  // wait for slab s and hand its results to the caller
        HostSlab& hs = d->slab[s];
        if (!hs.busy) return 0;
        cudaError_t e = cudaEventSynchronize(hs.done);
        hs.busy = false;
        if (e != cudaSuccess) return fail_cuda(e, "cudaEventSynchronize");
        if (got) memcpy(got + hs.first, hs.h_got, hs.n * sizeof(uint32_t));
        if (ok) memcpy(ok + hs.first, hs.h_ok, hs.n);
        bad_total += *hs.h_nbad;
        return 0;
    };
    while (i < nobj && !rc) {
        if (len[i] > kSlabBytes) {  // a big object: stream it alone (it uses both device slabs)
            for (int s2 = 0; s2 < 2 && !rc; s2++) rc = collect(s2);
            if (rc) break;
            uint32_t c = 0;
            rc = b200_crc32c_stream_host(ctx, dev, 0, bufs[i], len[i], 0, &c);
            if (rc) break;
            if (got) got[i] = c;
            bool good = want ? c == want[i] : true;
            if (ok) ok[i] = good;
            bad_total += !good;
            i++;
            continue;
        }
        rc = collect(slot);  // device slab `slot` and its plan buffers are free once its last batch finished
        if (rc) break;
        HostSlab& hs = d->slab[slot];
        // choose the objects of this slab
        size_t first = i, used = 0, staged = 0;
        size_t cnt = 0;
        while (i < nobj && len[i] <= kSlabBytes && cnt < kMaxBatchObjects) {
            size_t start = (used + 15) & ~(size_t)15;
            if (start + len[i] > kSlabBytes) break;
            used = start + len[i];
            i++;
            cnt++;
        }
        std::lock_guard<std::mutex> lk(d->mu);
        cudaSetDevice(dev);
        if ((rc = hs.reserve(cnt))) break;
        // plan (the device slab is 256-byte aligned, so cuts depend on the offsets only)
        used = 0;
        for (size_t k = 0; k < cnt; k++) {
            size_t start = (used + 15) & ~(size_t)15;
            hs.h_off[k] = start;
            hs.h_len[k] = len[first + k];
            if (want) hs.h_want[k] = want[first + k];
            used = start + len[first + k];
        }
        auto rows_of = [&](size_t k) { return cut_object(hs.h_off[k], hs.h_len[k]).rows; };
        const HostPlan hp = make_plan(cnt, rows_of, d->grid);
        fill_task_prefix(cnt, rows_of, hp.p, hs.h_seg);
        if ((rc = hs.reserve_tasks(hp.ntasks))) break;
        BatchArgs a{};
        if ((rc = apply_plan(d, hp, &a))) break;
        cudaError_t e = cudaSuccess;
        auto cu = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
        cu(cudaMemcpyAsync(hs.d_off, hs.h_off, cnt * 8, cudaMemcpyHostToDevice, d->copy_stream));
        cu(cudaMemcpyAsync(hs.d_len, hs.h_len, cnt * 8, cudaMemcpyHostToDevice, d->copy_stream));
        cu(cudaMemcpyAsync(hs.d_seg, hs.h_seg, 3 * (cnt + 1) * 4, cudaMemcpyHostToDevice, d->copy_stream));
        if (want) cu(cudaMemcpyAsync(hs.d_want, hs.h_want, cnt * 4, cudaMemcpyHostToDevice, d->copy_stream));
        // data: merge host-contiguous pinned neighbours into single copies
        size_t k = 0;
        while (k < cnt && e == cudaSuccess) {
            const uint8_t* src = (const uint8_t*)bufs[first + k];
            size_t n = hs.h_len[k];
            if (n == 0) { k++; continue; }
            if (is_pinned(src)) {
                size_t run_begin = hs.h_off[k], run_bytes = n, j = k + 1;
                while (j < cnt && (const uint8_t*)bufs[first + j] == src + (hs.h_off[j] - run_begin) &&
                       is_pinned(bufs[first + j])) {
                    run_bytes = hs.h_off[j] + hs.h_len[j] - run_begin;
                    j++;
                }
                cu(cudaMemcpyAsync(d->d_stage[slot] + run_begin, src, run_bytes, cudaMemcpyHostToDevice, d->copy_stream));
                k = j;
            } else {
                // pageable: through this slab's half of pinned staging, kStageBytes at a time
                size_t done = 0;
                while (done < n && e == cudaSuccess) {
                    size_t piece = std::min(n - done, kStageBytes - staged % kStageBytes);
                    if (staged && staged % kStageBytes == 0) cu(cudaStreamSynchronize(d->copy_stream));
                    memcpy(d->h_stage[slot] + staged % kStageBytes, src + done, piece);
                    cu(cudaMemcpyAsync(d->d_stage[slot] + hs.h_off[k] + done, d->h_stage[slot] + staged % kStageBytes,
                                       piece, cudaMemcpyHostToDevice, d->copy_stream));
                    done += piece;
                    staged += piece;
                }
                k++;
            }
        }
        if (staged) cu(cudaStreamSynchronize(d->copy_stream));  // pinned staging of this slot is reusable next round
        cu(cudaEventRecord(d->ev_copied[slot], d->copy_stream));
        cu(cudaStreamWaitEvent(d->stream, d->ev_copied[slot], 0));
        if (e != cudaSuccess) { rc = fail_cuda(e, "slab upload"); break; }
        a.base = d->d_stage[slot];
        a.off = hs.d_off;
        a.len = hs.d_len;
        a.seg_start = hs.d_seg;
        a.want = want ? hs.d_want : nullptr;
        a.seg_raw = hs.d_seg_raw;
        a.got = hs.d_got;
        a.ok = hs.d_ok;
        a.nbad = hs.d_nbad;
        a.nobj = (uint32_t)cnt;
        cu(launch_batch(a, d->tabs, d->grid, d->stream, nullptr, nullptr, &d->side));
        cu(cudaMemcpyAsync(hs.h_got, hs.d_got, cnt * 4, cudaMemcpyDeviceToHost, d->stream));
        cu(cudaMemcpyAsync(hs.h_ok, hs.d_ok, cnt, cudaMemcpyDeviceToHost, d->stream));
        cu(cudaMemcpyAsync(hs.h_nbad, hs.d_nbad, 8, cudaMemcpyDeviceToHost, d->stream));
        cu(cudaEventRecord(hs.done, d->stream));
        if (e != cudaSuccess) { rc = fail_cuda(e, "slab verify"); break; }
        hs.busy = true;
        hs.first = first;
        hs.n = cnt;
        slot ^= 1;
    }
    for (int s = 0; s < 2; s++) {
        int r = collect(s);
        if (!rc) rc = r;
    }
    if (!rc && nbad) *nbad = bad_total;
    return rc;
}

}  // extern "C"

// ---- hash.Hash32 --------------------------------------------------------------------------------------

struct b200_hasher {
    b200_ctx* ctx = nullptr;
    DevState* d = nullptr;
    int dev = 0;
    // Two pinned staging buffers and their device twins.  A full buffer is handed to the GPU (H2D on the copy
    // stream, kernels on the compute stream, running CRC kept on the device) while the caller fills the other;
    // only Sum32 waits for the GPU.
    uint8_t* h_buf[2] = {nullptr, nullptr};
    uint8_t* d_buf[2] = {nullptr, nullptr};
    cudaEvent_t copied[2] = {nullptr, nullptr}, done[2] = {nullptr, nullptr};
    bool pending[2] = {false, false};
    uint32_t* d_state = nullptr;  // CRC of everything handed over so far
    uint32_t* h_state = nullptr;  // pinned
    int cur = 0;
    size_t fill = 0, cap = 0;
};

extern "C" {

int b200_hasher_free(b200_hasher* h) {
    if (!h) return 0;
    cudaSetDevice(h->dev);
    for (int b = 0; b < 2; b++) {
        if (h->pending[b]) cudaEventSynchronize(h->done[b]);
        cudaFreeHost(h->h_buf[b]);
        cudaFree(h->d_buf[b]);
        if (h->copied[b]) cudaEventDestroy(h->copied[b]);
        if (h->done[b]) cudaEventDestroy(h->done[b]);
    }
    cudaFree(h->d_state);
    cudaFreeHost(h->h_state);
    cudaGetLastError();
    delete h;
    return 0;
}

int b200_hasher_new(b200_ctx* ctx, int dev, b200_hasher** out) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !out) return fail(B200_ERR_INVALID, "b200_hasher_new");
    b200_hasher* h = new (std::nothrow) b200_hasher;
    if (!h) return fail(B200_ERR_NOMEM, "b200_hasher_new");
    h->ctx = ctx;
    h->d = d;
    h->dev = dev;
    h->cap = 8u << 20;  // four of the read loop's 2 MiB buffers (main.go:125)
    cudaError_t e = cudaSetDevice(dev);
    for (int b = 0; b < 2 && e == cudaSuccess; b++) {
        e = cudaHostAlloc(&h->h_buf[b], h->cap, cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaMalloc(&h->d_buf[b], h->cap);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->copied[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->done[b], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaMalloc(&h->d_state, sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaHostAlloc(&h->h_state, sizeof(uint32_t), cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaMemset(h->d_state, 0, sizeof(uint32_t));
    if (e != cudaSuccess) {
        b200_hasher_free(h);
        return fail_cuda(e, "b200_hasher_new");
    }
    *out = h;
    return 0;
}

// Hand the current staging buffer to the GPU (asynchronously) and switch to the other one.
static int hasher_flush(b200_hasher* h) {
    if (!h->fill) return 0;
    DevState* d = h->d;
    const int b = h->cur;
    {
        std::lock_guard<std::mutex> lk(d->mu);
        CU(cudaSetDevice(h->dev));
        CU(cudaMemcpyAsync(h->d_buf[b], h->h_buf[b], h->fill, cudaMemcpyHostToDevice, d->copy_stream));
        CU(cudaEventRecord(h->copied[b], d->copy_stream));
        CU(cudaStreamWaitEvent(d->stream, h->copied[b], 0));
        if (int rc = enqueue_single(d, h->d_buf[b], h->fill, 0, 0, h->d_state)) return rc;
        CU(cudaEventRecord(h->done[b], d->stream));
    }
    h->pending[b] = true;
    h->cur = b ^ 1;
    h->fill = 0;
    if (h->pending[h->cur]) {  // the buffer we are about to fill must have been consumed
        CU(cudaEventSynchronize(h->done[h->cur]));
        h->pending[h->cur] = false;
    }
    return 0;
}

int b200_hasher_write(b200_hasher* h, const void* p, size_t n) {
    if (!h || (n && !p)) return fail(B200_ERR_INVALID, "b200_hasher_write");
    const uint8_t* s = (const uint8_t*)p;
    while (n) {
        size_t take = std::min(n, h->cap - h->fill);
        memcpy(h->h_buf[h->cur] + h->fill, s, take);
        h->fill += take;
        s += take;
        n -= take;
        if (h->fill == h->cap)
            if (int rc = hasher_flush(h)) return rc;
    }
    return 0;
}

int b200_hasher_reserve(b200_hasher* h, void** ptr, size_t* avail) {
    if (!h || !ptr || !avail) return fail(B200_ERR_INVALID, "b200_hasher_reserve");
    if (h->fill == h->cap)
        if (int rc = hasher_flush(h)) return rc;
    *ptr = h->h_buf[h->cur] + h->fill;
    *avail = h->cap - h->fill;
    return 0;
}

int b200_hasher_commit(b200_hasher* h, size_t n) {
    if (!h || n > h->cap - h->fill) return fail(B200_ERR_INVALID, "b200_hasher_commit: more than was reserved");
    h->fill += n;
    if (h->fill == h->cap) return hasher_flush(h);
    return 0;
}

int b200_hasher_sum32(b200_hasher* h, uint32_t* out) {
    if (!h || !out) return fail(B200_ERR_INVALID, "b200_hasher_sum32");
    if (int rc = hasher_flush(h)) return rc;
    DevState* d = h->d;
    cudaEvent_t last = nullptr;
    {
        std::lock_guard<std::mutex> lk(d->mu);
        CU(cudaSetDevice(h->dev));
        CU(cudaMemcpyAsync(h->h_state, h->d_state, sizeof(uint32_t), cudaMemcpyDeviceToHost, d->stream));
        // wait for this hasher's own work only: other hashers keep the stream busy
        last = h->done[h->cur ^ 1];
        CU(cudaEventRecord(last, d->stream));
    }
    CU(cudaEventSynchronize(last));
    h->pending[0] = h->pending[1] = false;  // everything before `last` on the stream has finished
    *out = *h->h_state;
    return 0;
}

int b200_hasher_reset(b200_hasher* h) {
    if (!h) return fail(B200_ERR_INVALID, "b200_hasher_reset");
    DevState* d = h->d;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(h->dev));
    CU(cudaMemsetAsync(h->d_state, 0, sizeof(uint32_t), d->stream));  // ordered after the chunks already queued
    h->fill = 0;
    return 0;
}

}  // extern "C"

// ---- latency reduction ----------------------------------------------------------------------------------

namespace {

void fill_stats(const LatMoments& m, int64_t divisor, b200_lat_stats* stats) {
    memset(stats, 0, sizeof *stats);
    stats->count = m.count;
    if (m.count) {
        stats->min = m.min_ns / divisor;  // C division truncates toward zero, as Go's does
        stats->max = m.max_ns / divisor;
    }
    stats->sum = m.sum_lo;
    stats->sum_hi = m.sum_hi;
    stats->sumsq_lo = m.sumsq_lo;
    stats->sumsq_hi = m.sumsq_hi;
}

}  // namespace

extern "C" {

int b200_hist_fixed(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                    const double* edges, int nedges, int mode, uint64_t* counts, b200_lat_stats* stats) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !edges || !counts || (n && !d_samples) || ((uintptr_t)d_samples & 7) ||
        (mode != B200_HIST_NUMPY && mode != B200_HIST_OPENCENSUS) || nedges > B200_HIST_MAX_EDGES)
        return fail(B200_ERR_INVALID, "b200_hist_fixed");
    HistParams P;
    if (host_build_hist_params(edges, nedges, mode, divisor, &P))
        return fail(B200_ERR_INVALID, "b200_hist_fixed: edges must be ascending, divisor >= 1");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    CU(launch_hist((const long long*)d_samples, n, d->lat, stats != nullptr, d->stream, P));
    unsigned long long c[kMaxThresholds + 1];
    LatMoments m;
    CU(lat_read_results(d->lat, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    memcpy(c, d->lat.h_results->counts, sizeof c);
    m = d->lat.h_results->mom;
    uint64_t below = 0, above = 0;
    if (mode == B200_HIST_NUMPY) {
        for (int i = 0; i < nedges - 1; i++) counts[i] = c[i + 1];
        below = c[0];
        above = c[nedges];
    } else {
        for (int i = 0; i <= nedges; i++) counts[i] = c[i];
    }
    if (stats) {
        fill_stats(m, divisor, stats);
        stats->dropped_below = below;
        stats->dropped_above = above;
    }
    return 0;
}

int b200_hist_fixed_pair(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                         const double* edges_a, int nedges_a, int mode_a, uint64_t* counts_a,
                         const double* edges_b, int nedges_b, int mode_b, uint64_t* counts_b,
                         uint64_t* dropped_b, b200_lat_stats* stats) {
    DevState* d = get_dev(ctx, dev);
    auto mode_ok = [](int m) { return m == B200_HIST_NUMPY || m == B200_HIST_OPENCENSUS; };
    if (!d || !edges_a || !edges_b || !counts_a || !counts_b || (n && !d_samples) || ((uintptr_t)d_samples & 7) ||
        !mode_ok(mode_a) || !mode_ok(mode_b) || nedges_a > B200_HIST_MAX_EDGES || nedges_b > B200_HIST_MAX_EDGES)
        return fail(B200_ERR_INVALID, "b200_hist_fixed_pair");
    // One pass with the union of the two threshold sets; each histogram's bin is then a sum of union bins.
    long long ta[kMaxThresholds], tb[kMaxThresholds], u[2 * kMaxThresholds];
    const int ma = host_edges_to_thresholds(edges_a, nedges_a, mode_a, divisor, ta);
    const int mb = host_edges_to_thresholds(edges_b, nedges_b, mode_b, divisor, tb);
    if (ma < 0 || mb < 0) return fail(B200_ERR_INVALID, "b200_hist_fixed_pair: edges must be ascending, divisor >= 1");
    const int mu = (int)(std::unique(u, std::merge(ta, ta + ma, tb, tb + mb, u)) - u);
    if (mu > kMaxThresholds)
        return fail(B200_ERR_TOO_LARGE, "b200_hist_fixed_pair: more than 66 distinct thresholds between the two edge sets");
    HistParams P;
    if (host_build_hist_params_thr(u, mu, divisor, false, &P)) return fail(B200_ERR_INVALID, "b200_hist_fixed_pair");
    unsigned long long c[kMaxThresholds + 1];
    LatMoments m;
    {
        std::lock_guard<std::mutex> lk(d->mu);
        CU(cudaSetDevice(dev));
        CU(launch_hist((const long long*)d_samples, n, d->lat, stats != nullptr, d->stream, P));
        CU(lat_read_results(d->lat, d->stream));
        CU(cudaStreamSynchronize(d->stream));
        memcpy(c, d->lat.h_results->counts, sizeof c);
        m = d->lat.h_results->mom;
    }
    // union bin g holds the samples with exactly g union thresholds <= them, i.e. those in [u[g-1], u[g]); the
    // number of a histogram's own thresholds <= such a sample is the number of them <= u[g-1]
    auto split = [&](const long long* t, int mt, int mode, int nedges, uint64_t* counts, uint64_t* below, uint64_t* above) {
        uint64_t own[kMaxThresholds + 1] = {0};
        for (int g = 0; g <= mu; g++) {
            const int k = g == 0 ? 0 : (int)(std::upper_bound(t, t + mt, u[g - 1]) - t);
            own[k] += c[g];
        }
        *below = *above = 0;
        if (mode == B200_HIST_NUMPY) {
            for (int i = 0; i < nedges - 1; i++) counts[i] = own[i + 1];
            *below = own[0];
            *above = own[nedges];
        } else {
            for (int i = 0; i <= nedges; i++) counts[i] = own[i];
        }
    };
    uint64_t below_a, above_a, below_b, above_b;
    split(ta, ma, mode_a, nedges_a, counts_a, &below_a, &above_a);
    split(tb, mb, mode_b, nedges_b, counts_b, &below_b, &above_b);
    if (dropped_b) {
        dropped_b[0] = below_b;
        dropped_b[1] = above_b;
    }
    if (stats) {
        fill_stats(m, divisor, stats);
        stats->dropped_below = below_a;
        stats->dropped_above = above_a;
    }
    return 0;
}

int b200_minmax(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t* mn, int64_t* mx) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !mn || !mx || (n && !d_samples) || ((uintptr_t)d_samples & 7)) return fail(B200_ERR_INVALID, "b200_minmax");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    CU(launch_minmax((const long long*)d_samples, n, d->lat, d->stream));
    LatMoments m;
    CU(cudaMemcpyAsync(&m, d->lat.d_mom, sizeof m, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    *mn = m.min_ns;
    *mx = m.max_ns;
    return 0;
}

int b200_radix_count(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, uint64_t key_base, int shift,
                     int top, int nk, const uint64_t* prefix, uint64_t* counts) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !counts || (n && !d_samples) || ((uintptr_t)d_samples & 7) || nk < 1 || nk > B200_SELECT_MAX_K || shift < 0 ||
        top <= shift ||
        top > 64 || top - shift > kRadixBits || (top < 64 && !prefix))
        return fail(B200_ERR_INVALID, "b200_radix_count");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    if (top < 64)
        CU(cudaMemcpyAsync(d->lat.d_prefix, prefix, nk * sizeof(uint64_t), cudaMemcpyHostToDevice, d->stream));
    CU(launch_radix_pass((const long long*)d_samples, n, key_base, shift, top, nk, d->lat, d->stream));
    CU(cudaMemcpyAsync(d->lat.h_radix, d->lat.d_radix, (size_t)nk * kRadixBins * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                       d->stream));
    CU(cudaStreamSynchronize(d->stream));
    memcpy(counts, d->lat.h_radix, (size_t)nk * kRadixBins * sizeof(uint64_t));
    return 0;
}

int b200_radix_count_based(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, const uint64_t* key_bases,
                           int shift, int top, int nk, const uint64_t* prefix, uint64_t* counts) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !counts || !key_bases || !prefix || (n && !d_samples) || ((uintptr_t)d_samples & 7) || nk < 1 ||
        nk > B200_SELECT_MAX_K || shift < 0 || top <= shift || top >= 64 || top - shift > kRadixBits)
        return fail(B200_ERR_INVALID, "b200_radix_count_based");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    uint64_t both[2 * kMaxSelect] = {};
    for (int i = 0; i < nk; i++) {
        both[i] = prefix[i];
        both[kMaxSelect + i] = key_bases[i];
    }
    CU(cudaMemcpyAsync(d->lat.d_prefix, both, sizeof both, cudaMemcpyHostToDevice, d->stream));
    CU(launch_radix_pass((const long long*)d_samples, n, 0, shift, top, nk, d->lat, d->stream, true));
    CU(cudaMemcpyAsync(d->lat.h_radix, d->lat.d_radix, (size_t)nk * kRadixBins * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                       d->stream));
    CU(cudaStreamSynchronize(d->stream));
    memcpy(counts, d->lat.h_radix, (size_t)nk * kRadixBins * sizeof(uint64_t));
    return 0;
}

}  // extern "C"

namespace {

int try_sampled(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const uint64_t* k, int nk,
                int64_t* out, b200_lat_stats* stats = nullptr, bool no_squares = false);  // below, after the radix select it builds on

// Inputs up to this many samples are selected with all passes queued on the device (one host round trip);
// longer ones that the sampled select did not settle take the host-driven passes with bucket compaction
// (2^26 samples: 1.1 ms resident, 0.64 ms host-driven with compaction, 0.57 ms sampled).
size_t resident_select_max() {  // tuning / test override: B200_SELECT_RESIDENT_MAX=0 forces the host-driven passes
    static const size_t v = [] {
        const char* e = getenv("B200_SELECT_RESIDENT_MAX");
        return e && *e ? (size_t)strtoull(e, nullptr, 10) : (size_t)1 << 25;  // measured: tools/select_routes.py
    }();
    return v;
}

bool grow(long long** buf, size_t* cap, size_t want);  // below
bool select_trace_on();

// A stratified sample to take first, on the same stream and under the same lock: the select then runs over it
// (d->lat.d_sample, n = its length) with the bounds the sampling kernel leaves behind.
struct SampleFirst {
    const int64_t* src;
    size_t stride;
    unsigned long long outside;  // out: samples of the sample outside [0, 2^32)
};

// have_bounds: 0 none, 1 [lo, hi] given.  approx_max > 0 (one-kernel select only; the separate launches give exact
// ranks, which serve as well): out[i] is a bound on the rank's order statistic, from below — from above for the ranks
// in ceil_of — within approx_max ranks.
int select_resident(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, bool have_bounds,
                    int64_t lo, int64_t hi, const uint64_t* k, int nk, int64_t* out, uint64_t approx_max = 0,
                    const bool* ceil_of = nullptr, SampleFirst* first = nullptr) {
    DevState* d = get_dev(ctx, dev);
    if (!d) return fail(B200_ERR_INVALID, "b200_select_kth: unknown device");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    const bool fused = select_fused_grid() != 0;
    int bounds = have_bounds ? 1 : 0;
    if (first) {
        if (!grow(&d->lat.d_sample, &d->lat.sample_cap, n)) return 1;  // (the caller takes another route)
        CU(launch_sample((const long long*)first->src, first->stride, d->lat.d_sample, n, d->stream, d->lat.d_compact_n,
                         fused ? d->lat.d_mom : nullptr, &d->lat));
        d_samples = (const int64_t*)d->lat.d_sample;
        bounds = fused ? 2 : 0;
    }
    for (int g = 0; g < nk; g += kSelMax) {
        SelRanks r{};
        r.nk = std::min(kSelMax, nk - g);
        for (int i = 0; i < r.nk; i++) r.rank[i] = k[g + i];
        if (fused) {
            unsigned ceil_mask = 0;
            for (int i = 0; i < r.nk && ceil_of; i++) ceil_mask |= ceil_of[g + i] ? 1u << i : 0u;
            CU(launch_select_fused((const long long*)d_samples, n, r, bounds, lo, hi, approx_max, ceil_mask, d->lat,
                                   d->stream));
        } else {
            CU(launch_select_resident((const long long*)d_samples, n, r, bounds == 1, lo, hi, d->lat, d->stream));
        }
        CU(lat_read_results(d->lat, d->stream));  // the select's state (and, after a sampling kernel, its counter)
        CU(cudaStreamSynchronize(d->stream));
        const SelState& st = d->lat.h_results->sel;
        if (fused) *d->lat.sel_hist_clean = true;  // the kernel ran to its end: it leaves its histograms cleared
        if (first && g == 0) first->outside = d->lat.h_results->compact_n;
        if (st.bad || st.top != 0) {
            if (select_trace_on()) fprintf(stderr, "b200 select: resident select failed: bad=%d top=%d n=%zu\n", st.bad, st.top, n);
            return fail(B200_ERR_INVALID, "b200_select_kth: samples outside the given [lo, hi] range");
        }
        for (int i = 0; i < r.nk; i++) out[g + i] = st.result[i] / divisor;  // truncating division is monotone
        if (fused && select_trace_on() && st.nstamp > 1) {
            char msg[400];
            int at = snprintf(msg, sizeof msg, "one-kernel select of %d ranks among %zu, CTA 0 phase ends (us):", r.nk, n);
            for (int i = 1; i < st.nstamp && at < (int)sizeof msg - 12; i++)
                at += snprintf(msg + at, sizeof msg - at, " %.1f", (double)(st.stamp[i] - st.stamp[0]) * 1e-3);
            fprintf(stderr, "b200 select: %s\n", msg);
        }
    }
    return 0;
}

// After the first radix pass: copy the samples of the buckets still in play (distinct prefixes pf[0..np)) into
// the device's compaction buffer so that the remaining passes read `total` samples instead of n.  Returns 0
// and sets *out when done, 1 when the buckets are too large to be worth it (caller keeps the full data).
int compact_buckets(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, uint64_t key_base, int top,
                    const uint64_t* pf, int np, uint64_t total, const int64_t** out) {
    DevState* d = get_dev(ctx, dev);
    // worth it only when few samples survive (measured: at 25 % the filtering pass costs more than it saves):
    // one filtering pass + 2 passes over `total` samples instead of 2 passes over n
    if (!d || total == 0 || total > n / 8 || n < (1u << 20)) return 1;
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    if (d->lat.compact_cap < total) {
        cudaFree(d->lat.d_compact);
        d->lat.d_compact = nullptr;
        d->lat.compact_cap = 0;
        cudaError_t e = cudaMalloc(&d->lat.d_compact, total * sizeof(long long));
        if (e != cudaSuccess) {  // no room for the shortcut: not an error
            cudaGetLastError();
            return 1;
        }
        d->lat.compact_cap = total;
    }
    CU(cudaMemcpyAsync(d->lat.d_prefix, pf, np * sizeof(uint64_t), cudaMemcpyHostToDevice, d->stream));
    CU(launch_compact((const long long*)d_samples, n, key_base, top, np, d->lat, d->stream));
    unsigned long long got = 0;
    CU(cudaMemcpyAsync(&got, d->lat.d_compact_n, sizeof got, cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    if (got != total) return fail(B200_ERR_CUDA, "select: compaction count differs from the histogram");
    *out = (const int64_t*)d->lat.d_compact;
    return 0;
}

}  // namespace

extern "C" {

int b200_select_kth_range(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, int64_t lo_ns,
                          int64_t hi_ns, const uint64_t* k, int nk, int64_t* out) {
    if (!k || !out || nk < 1 || divisor < 1 || lo_ns > hi_ns) return fail(B200_ERR_INVALID, "b200_select_kth");
    if (n == 0) return fail(B200_ERR_EMPTY, "b200_select_kth: no samples (the reference panics, ssd_test/main.go:157)");
    for (int i = 0; i < nk; i++)
        if (k[i] >= n) return fail(B200_ERR_INVALID, "b200_select_kth: k >= n");
    DevState* dsel = get_dev(ctx, dev);
    if (!dsel) return fail(B200_ERR_INVALID, "b200_select_kth: unknown device");
    {   // the one-pass sampled select needs no bounds; when it misses, the radix passes below use them
        const int rc = try_sampled(ctx, dev, d_samples, n, divisor, k, nk, out);
        if (rc <= 0) return rc;
    }
    if (n <= resident_select_max()) return select_resident(ctx, dev, d_samples, n, divisor, true, lo_ns, hi_ns, k, nk, out);
    std::lock_guard<std::recursive_mutex> sel_lock(dsel->select_mu);
    const uint64_t kmin = (uint64_t)lo_ns ^ 0x8000000000000000ull, kmax = (uint64_t)hi_ns ^ 0x8000000000000000ull;
    const uint64_t range = kmax - kmin;
    int bits = 0;
    while (bits < 64 && (range >> bits)) bits++;
    std::vector<uint64_t> counts((size_t)kMaxSelect * kRadixBins);
    for (int g = 0; g < nk; g += kMaxSelect) {
        int m = std::min(kMaxSelect, nk - g);
        uint64_t prefix[kMaxSelect] = {0, 0, 0, 0}, rank[kMaxSelect];
        for (int i = 0; i < m; i++) rank[i] = k[g + i];
        int top = bits;
        bool first = true;
        const int64_t* cur = d_samples;  // what the next pass reads: all samples, or the compacted buckets
        size_t cur_n = n;
        if (top == 0) {  // all samples equal lo_ns: nothing to resolve
            for (int i = 0; i < m; i++) prefix[i] = 0;
        }
        while (top > 0) {
            int shift = std::max(top - kRadixBits, 0);
            int width = top - shift;
            // first pass: key' < 2^top, so every sample matches prefix 0 and one histogram serves all k
            int rc = b200_radix_count(ctx, dev, cur, cur_n, kmin, shift, top, first ? 1 : m, prefix, counts.data());
            if (rc) return rc;
            uint64_t bucket[kMaxSelect];
            for (int i = 0; i < m; i++) {
                const uint64_t* h = counts.data() + (first ? 0 : (size_t)i * kRadixBins);
                uint64_t cum = 0;
                int dgt = 0;
                for (; dgt < (1 << width); dgt++) {
                    if (rank[i] < cum + h[dgt]) break;
                    cum += h[dgt];
                }
                if (dgt == (1 << width))
                    return fail(B200_ERR_INVALID, "b200_select_kth: samples outside the given [lo, hi] range");
                rank[i] -= cum;
                prefix[i] = (prefix[i] << width) | (uint64_t)dgt;
                bucket[i] = h[dgt];
            }
            top = shift;
            if (first && top > 0) {
                // the wanted ranks now live in at most m buckets: keep just those samples for the next passes
                uint64_t pf[kMaxSelect], total = 0;
                int np = 0;
                for (int i = 0; i < m; i++) {
                    bool seen = false;
                    for (int j = 0; j < np; j++) seen |= pf[j] == prefix[i];
                    if (!seen) { pf[np++] = prefix[i]; total += bucket[i]; }
                }
                const int64_t* small = nullptr;
                rc = compact_buckets(ctx, dev, d_samples, n, kmin, top, pf, np, total, &small);
                if (rc < 0) return rc;
                if (rc == 0) { cur = small; cur_n = (size_t)total; }
            }
            first = false;
        }
        for (int i = 0; i < m; i++) {
            int64_t ns = (int64_t)((prefix[i] + kmin) ^ 0x8000000000000000ull);
            out[g + i] = ns / divisor;  // truncating division is monotone: k-th of v = trunc(k-th of ns / divisor)
        }
    }
    return 0;
}

}  // extern "C"

namespace {

thread_local int g_sampled_depth = 0;  // the selects a sampled select runs on its own buffers must not nest

size_t sampled_select_min() {  // inputs at least this long try the sampled shortcut first (tuning / test override)
    static const size_t v = [] {
        const char* e = getenv("B200_SELECT_SAMPLED_MIN");
        return e && *e ? (size_t)strtoull(e, nullptr, 10) : (size_t)1 << 24;
    }();
    return v;
}

bool select_trace_on() {  // B200_SELECT_TRACE=1
    static const bool on = [] {
        const char* e = getenv("B200_SELECT_TRACE");
        return e && *e && *e != '0';
    }();
    return on;
}

int sampled_trace(const char* what, int rc) {  // B200_SELECT_TRACE=1: say on stderr how a sampled select ended
    if (select_trace_on()) fprintf(stderr, "b200 select: sampled %s\n", what);
    return rc;
}

bool grow(long long** buf, size_t* cap, size_t want) {
    if (*cap >= want) return true;
    cudaFree(*buf);
    *buf = nullptr;
    *cap = 0;
    // a little headroom, so that a caller whose inputs creep up in length does not pay cudaFree + cudaMalloc
    // (about a millisecond each) on every call
    size_t room = want + want / 4;
    if (cudaMalloc(buf, room * sizeof(long long)) != cudaSuccess) {
        cudaGetLastError();
        room = want;
        if (cudaMalloc(buf, room * sizeof(long long)) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
    }
    *cap = room;
    return true;
}

// Exact order statistics in ONE pass over the samples instead of min/max + three radix passes:
//   1. a stratified sample (one input per stratum of `stride`) gives, for every wanted rank, a bracket
//      [lo, hi] of sample order statistics a few standard deviations either side of the rank's expected position;
//   2. one histogram pass over all samples with the bracket ends as thresholds counts the samples below each
//      bracket and keeps the samples inside the brackets (about 1 % of the input);
//   3. if every rank did fall into its bracket (checked with the exact counts), it is resolved among the kept
//      samples, by the ordinary radix select.
// The result is exact whenever it is given; a miss (returns 1) makes the caller take the radix passes.
int select_sampled(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const uint64_t* k,
                   int nk, int64_t* out, b200_lat_stats* stats, bool no_squares) {
    DevState* d = get_dev(ctx, dev);
    if (!d || nk > kSelMax) return 1;
    std::lock_guard<std::recursive_mutex> sel_lock(d->select_mu);
    struct Depth {
        Depth() { g_sampled_depth++; }
        ~Depth() { g_sampled_depth--; }
    } depth;
    const auto t0 = std::chrono::steady_clock::now();
    auto ms_since = [&](std::chrono::steady_clock::time_point t) {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t).count();
    };
    // the sample: ~2^22 inputs (B200_SELECT_SAMPLE_LOG2 for tuning runs)
    static const int sample_log2 = [] {
        const char* e = getenv("B200_SELECT_SAMPLE_LOG2");
        const int v = e && *e ? atoi(e) : 22;
        return v < 16 ? 16 : v > 26 ? 26 : v;
    }();
    const size_t stride = std::max<size_t>(16, (n + ((size_t)1 << sample_log2) - 1) >> sample_log2);
    const size_t s = n / stride;
    if (s < 4096) return sampled_trace("skip: input too short", 1);
    // sample rank of an input rank r is about r / stride, with a standard deviation below sqrt(s) / 2
    const uint64_t margin = (uint64_t)(2.5 * std::sqrt((double)s)) + 2;
    uint64_t want[2 * kSelMax];
    bool open_lo[kSelMax], open_hi[kSelMax], is_hi[2 * kSelMax] = {};
    int nwant = 0, lo_at[kSelMax], hi_at[kSelMax];
    for (int i = 0; i < nk; i++) {
        const uint64_t c = k[i] / stride;
        open_lo[i] = c < margin;
        open_hi[i] = c + margin + 1 >= s;
        lo_at[i] = hi_at[i] = -1;
        if (!open_lo[i]) { lo_at[i] = nwant; want[nwant++] = c - margin; }
        if (!open_hi[i]) { hi_at[i] = nwant; is_hi[nwant] = true; want[nwant++] = c + margin + 1; }
    }
    int64_t got[2 * kSelMax];
    unsigned long long outside = 0;
    if (nwant) {
        // bounds are enough here: a lower end at or below the sample's order statistic, an upper end at or above it,
        // good to a sixteenth of the `margin` either side (s <= 2^22 + change: always the resident select).  The
        // sampling kernel and the select run back to back on the stream: one host round trip for both.
        SampleFirst first{d_samples, stride, 0};
        const int rc = select_resident(ctx, dev, nullptr, s, 1, false, 0, 0, want, nwant, got,
                                       std::max<uint64_t>(1, margin / 16), is_hi, &first);
        if (rc) return rc < 0 ? rc : sampled_trace("skip: no memory for the sample", 1);
        outside = first.outside;
    } else {
        std::unique_lock<std::mutex> lk(d->mu);
        CU(cudaSetDevice(dev));
        if (!grow(&d->lat.d_sample, &d->lat.sample_cap, s)) return sampled_trace("skip: no memory for the sample", 1);
        CU(launch_sample((const long long*)d_samples, stride, d->lat.d_sample, s, d->stream, d->lat.d_compact_n));
        CU(cudaMemcpyAsync(&outside, d->lat.d_compact_n, sizeof outside, cudaMemcpyDeviceToHost, d->stream));
        CU(cudaStreamSynchronize(d->stream));
    }
    // the histogram pass takes negative samples and samples >= 2^32 ns one by one on its generic path: a
    // multi-second tail is fine, a data set that mostly lives there (not latencies) is faster by radix passes
    if (outside > s / 16) return sampled_trace("skip: too much of the input is outside the 32-bit fast path", 1);
    const double ms_sample = ms_since(t0);
    // brackets, ascending and disjoint (overlapping or touching ones are merged)
    struct Bracket { int64_t lo, hi; };
    const int64_t kMin = INT64_MIN, kMax = INT64_MAX - 1;  // hi + 1 must not overflow
    Bracket br[kSelMax];
    for (int i = 0; i < nk; i++) {
        br[i].lo = open_lo[i] ? kMin : got[lo_at[i]];
        br[i].hi = open_hi[i] ? kMax : std::min(got[hi_at[i]], kMax);
    }
    for (int i = 1; i < nk; i++)  // at most kSelMax entries: insertion sort by lo
        for (int j = i; j > 0 && br[j].lo < br[j - 1].lo; j--) std::swap(br[j], br[j - 1]);
    int nb = 0;
    for (int i = 0; i < nk; i++) {
        if (nb && br[i].lo <= br[nb - 1].hi + 1) br[nb - 1].hi = std::max(br[nb - 1].hi, br[i].hi);
        else br[nb++] = br[i];
    }
    long long thr[2 * kSelMax];
    for (int j = 0; j < nb; j++) {
        thr[2 * j] = br[j].lo;      // bins 2j+1 are the brackets, even bins the gaps
        thr[2 * j + 1] = br[j].hi + 1;
    }
    HistParams P;
    // a bracket that starts at INT64_MIN leaves bin 0 empty, which the builder allows (thresholds stay ascending)
    // the thresholds are raw nanoseconds; the divisor scales the moments (stats) and sets how far the kernel's fast
    // path reaches (a divisor with a power-of-two part takes samples beyond 2^32 ns in its stride)
    if (host_build_hist_params_thr(thr, 2 * nb, divisor, true, &P)) return sampled_trace("skip: thresholds", 1);
    // room for twice the expected number of samples inside the brackets, at least n/64; more than n/8 is not worth it
    const size_t expect = (size_t)nk * (2 * margin + 2) * stride;
    const size_t cap = std::min<size_t>(std::max<size_t>(2 * expect, n / 64), n / 8);
    unsigned long long c[kMaxThresholds + 1], kept = 0;
    LatMoments mom{};
    // The kept samples lie between the ends of the brackets: no min/max pass when those are finite — and then the
    // select over them is queued right behind the pass, with its ranks worked out on the device from the pass's
    // counts (bracket_ranks_kernel): one host round trip for both.  The host repeats the arithmetic on the counts it
    // reads back and only then believes the result.
    const bool bounded = br[0].lo != kMin && br[nb - 1].hi != kMax;
    const bool chained = bounded && select_fused_grid() != 0 && cap <= resident_select_max();
    SelState sel{};
    {
        std::unique_lock<std::mutex> lk(d->mu);
        CU(cudaSetDevice(dev));
        if (!grow(&d->lat.d_collect, &d->lat.collect_cap, cap)) return sampled_trace("skip: no memory for the brackets", 1);
        HistCollect col{d->lat.d_collect, d->lat.d_compact_n, cap};
        CU(launch_hist((const long long*)d_samples, n, d->lat, stats ? (no_squares ? 2 : 1) : 0, d->stream, P, &col));
        if (chained) {
            CU(launch_bracket_ranks(d->lat, nb, (const unsigned long long*)k, nk, cap, d->stream));
            CU(launch_select_fused((const long long*)d->lat.d_collect, cap, SelRanks{}, 1, br[0].lo, br[nb - 1].hi, 0, 0,
                                   d->lat, d->stream, true));
        }
        CU(lat_read_results(d->lat, d->stream));
        CU(cudaStreamSynchronize(d->stream));
        memcpy(c, d->lat.h_results->counts, sizeof c);
        mom = d->lat.h_results->mom;
        kept = d->lat.h_results->compact_n;
        if (chained) {
            sel = d->lat.h_results->sel;
            *d->lat.sel_hist_clean = true;  // the select ran to its end (or did not start): its histograms are clear
        }
    }
    const double ms_pass = ms_since(t0) - ms_sample;
    uint64_t inside = 0, total = 0;
    for (int j = 0; j <= 2 * nb; j++) total += c[j];
    for (int j = 0; j < nb; j++) inside += c[2 * j + 1];
    if (total != n || kept != inside) return fail(B200_ERR_CUDA, "select: bracket pass lost samples");
    if (inside > cap || inside == 0) return sampled_trace("miss: brackets too heavy", 1);
    // rank among the kept samples = (kept samples in earlier brackets) + (rank - samples below this bracket)
    uint64_t rank_kept[kSelMax];
    for (int i = 0; i < nk; i++) {
        uint64_t below = 0, kept_before = 0;
        int j = 0;
        for (;; j++) {
            if (j == nb) return sampled_trace("miss: rank beyond the brackets", 1);
            below += c[2 * j];
            if (k[i] < below) return sampled_trace("miss: rank between brackets", 1);
            if (k[i] < below + c[2 * j + 1]) break;
            below += c[2 * j + 1];
            kept_before += c[2 * j + 1];
        }
        rank_kept[i] = kept_before + (k[i] - below);
    }
    if (chained) {
        if (sel.bad || sel.top != 0) return sampled_trace("miss: the select over the kept samples", 1);
        for (int i = 0; i < nk; i++) out[i] = sel.result[i] / divisor;  // truncating division is monotone
    } else {
        int rc = inside <= resident_select_max()
                     ? select_resident(ctx, dev, (const int64_t*)d->lat.d_collect, (size_t)inside, divisor, bounded,
                                       bounded ? br[0].lo : 0, bounded ? br[nb - 1].hi : 0, rank_kept, nk, out)
                     : b200_select_kth(ctx, dev, (const int64_t*)d->lat.d_collect, (size_t)inside, divisor, rank_kept, nk, out);
        if (rc) return rc < 0 ? rc : 1;
    }
    if (stats) fill_stats(mom, divisor, stats);  // moments of the whole input, from the bracket pass
    char msg[160];
    snprintf(msg, sizeof msg, "ok (sample of %zu + brackets %.3f ms, pass %.3f ms, %llu kept + select %.3f ms)", s, ms_sample,
             ms_pass, (unsigned long long)inside, ms_since(t0) - ms_sample - ms_pass);
    return sampled_trace(msg, 0);
}

// 0 = out[] is filled, 1 = not attempted or missed (take the radix passes), < 0 = error
int try_sampled(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const uint64_t* k, int nk,
                int64_t* out, b200_lat_stats* stats, bool no_squares) {
    if (g_sampled_depth != 0 || n < sampled_select_min() || !d_samples || ((uintptr_t)d_samples & 7)) return 1;
    for (int i = 0; i < nk; i++)
        if (k[i] >= n) return 1;  // the caller reports it
    // up to kSelMax ranks share a pass; more take one pass per group (the moments come from the first)
    for (int g = 0; g < nk; g += kSelMax) {
        const int rc = select_sampled(ctx, dev, d_samples, n, divisor, k + g, std::min(kSelMax, nk - g), out + g,
                                      g == 0 ? stats : nullptr, no_squares);
        if (rc) return rc;  // a miss in any group: the caller redoes all of them by radix passes
    }
    return 0;
}

}  // namespace

extern "C" {

int b200_select_kth(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const uint64_t* k,
                    int nk, int64_t* out) {
    if (!k || !out || nk < 1 || divisor < 1) return fail(B200_ERR_INVALID, "b200_select_kth");
    if (n == 0) return fail(B200_ERR_EMPTY, "b200_select_kth: no samples (the reference panics, ssd_test/main.go:157)");
    {
        const int rc = try_sampled(ctx, dev, d_samples, n, divisor, k, nk, out);
        if (rc <= 0) return rc;  // done, or a real error; 1 = take the radix passes
    }
    if (n <= resident_select_max()) {
        for (int i = 0; i < nk; i++)
            if (k[i] >= n) return fail(B200_ERR_INVALID, "b200_select_kth: k >= n");
        if (!get_dev(ctx, dev) || !d_samples) return fail(B200_ERR_INVALID, "b200_select_kth");
        return select_resident(ctx, dev, d_samples, n, divisor, false, 0, 0, k, nk, out);
    }
    int64_t mn, mx;
    if (int rc = b200_minmax(ctx, dev, d_samples, n, &mn, &mx)) return rc;
    return b200_select_kth_range(ctx, dev, d_samples, n, divisor, mn, mx, k, nk, out);
}

int b200_select_sample(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, size_t stride,
                       const int64_t** d_sample, size_t* nsample) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !d_sample || !nsample || stride < 1 || (n && !d_samples) || ((uintptr_t)d_samples & 7))
        return fail(B200_ERR_INVALID, "b200_select_sample");
    std::lock_guard<std::recursive_mutex> sel_lock(d->select_mu);
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    const size_t s = n / stride;
    if (s && !grow(&d->lat.d_sample, &d->lat.sample_cap, s)) return fail(B200_ERR_NOMEM, "b200_select_sample");
    CU(launch_sample((const long long*)d_samples, stride, d->lat.d_sample, s, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    *d_sample = (const int64_t*)d->lat.d_sample;
    *nsample = s;
    return 0;
}

int b200_select_sample_bounds(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, size_t stride,
                              const uint64_t* ranks, const uint8_t* upper, int nranks, uint64_t slack, int64_t* out,
                              size_t* nsample) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !ranks || !out || !nsample || nranks < 1 || nranks > 2 * kSelMax || stride < 1 || !d_samples ||
        ((uintptr_t)d_samples & 7))
        return fail(B200_ERR_INVALID, "b200_select_sample_bounds");
    const size_t s = n / stride;
    for (int i = 0; i < nranks; i++)
        if (ranks[i] >= s) return fail(B200_ERR_INVALID, "b200_select_sample_bounds: rank >= n / stride");
    std::lock_guard<std::recursive_mutex> sel_lock(d->select_mu);
    bool up[2 * kSelMax] = {};
    for (int i = 0; i < nranks && upper; i++) up[i] = upper[i] != 0;
    SampleFirst first{d_samples, stride, 0};
    const int rc = select_resident(ctx, dev, nullptr, s, 1, false, 0, 0, ranks, nranks, out, slack, up, &first);
    if (rc > 0) return fail(B200_ERR_NOMEM, "b200_select_sample_bounds");
    if (rc == 0) *nsample = s;
    return rc;
}

namespace {
int bracket_pass(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const int64_t* thresholds,
                 int nthr, uint64_t cap, uint64_t* counts, const int64_t** d_kept, uint64_t* nkept, b200_lat_stats* stats,
                 bool no_squares);
}

int b200_bracket_pass(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                      const int64_t* thresholds, int nthr, uint64_t cap, uint64_t* counts, const int64_t** d_kept,
                      uint64_t* nkept, b200_lat_stats* stats) {
    return bracket_pass(ctx, dev, d_samples, n, divisor, thresholds, nthr, cap, counts, d_kept, nkept, stats, false);
}

int b200_bracket_pass_avg(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                          const int64_t* thresholds, int nthr, uint64_t cap, uint64_t* counts, const int64_t** d_kept,
                          uint64_t* nkept, b200_lat_stats* stats) {
    const int rc = bracket_pass(ctx, dev, d_samples, n, divisor, thresholds, nthr, cap, counts, d_kept, nkept, stats, true);
    if (rc == 0 && stats) stats->sumsq_lo = stats->sumsq_hi = 0;
    return rc;
}

}  // extern "C"

namespace {

int bracket_pass(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const int64_t* thresholds,
                 int nthr, uint64_t cap, uint64_t* counts, const int64_t** d_kept, uint64_t* nkept, b200_lat_stats* stats,
                 bool no_squares) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !thresholds || !counts || !d_kept || !nkept || nthr < 1 || nthr > B200_HIST_MAX_EDGES || divisor < 1 ||
        (n && !d_samples) || ((uintptr_t)d_samples & 7))
        return fail(B200_ERR_INVALID, "b200_bracket_pass");
    HistParams P;
    static_assert(sizeof(long long) == sizeof(int64_t), "threshold type");
    if (host_build_hist_params_thr((const long long*)thresholds, nthr, divisor, true, &P))
        return fail(B200_ERR_INVALID, "b200_bracket_pass: thresholds must be strictly ascending");
    std::lock_guard<std::recursive_mutex> sel_lock(d->select_mu);
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    if (cap && !grow(&d->lat.d_collect, &d->lat.collect_cap, (size_t)cap)) return fail(B200_ERR_NOMEM, "b200_bracket_pass");
    HistCollect col{d->lat.d_collect, d->lat.d_compact_n, cap};
    CU(launch_hist((const long long*)d_samples, n, d->lat, stats ? (no_squares ? 2 : 1) : 0, d->stream, P, &col));
    unsigned long long c[kMaxThresholds + 1], kept = 0;
    LatMoments m{};
    CU(lat_read_results(d->lat, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    memcpy(c, d->lat.h_results->counts, sizeof c);
    kept = d->lat.h_results->compact_n;
    m = d->lat.h_results->mom;
    for (int i = 0; i <= nthr; i++) counts[i] = c[i];
    *nkept = kept;
    *d_kept = (const int64_t*)d->lat.d_collect;
    if (stats) fill_stats(m, divisor, stats);
    return 0;
}

}  // namespace

extern "C" {

namespace {
int select_kth_with_stats(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const uint64_t* k,
                          int nk, int64_t* out, b200_lat_stats* stats, bool no_squares);
}

int b200_select_kth_stats(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                          const uint64_t* k, int nk, int64_t* out, b200_lat_stats* stats) {
    return select_kth_with_stats(ctx, dev, d_samples, n, divisor, k, nk, out, stats, false);
}

int b200_select_kth_avg(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor,
                        const uint64_t* k, int nk, int64_t* out, b200_lat_stats* stats) {
    const int rc = select_kth_with_stats(ctx, dev, d_samples, n, divisor, k, nk, out, stats, true);
    if (rc == 0) stats->sumsq_lo = stats->sumsq_hi = 0;  // not part of this call's contract on any route
    return rc;
}

}  // extern "C"

namespace {

int select_kth_with_stats(b200_ctx* ctx, int dev, const int64_t* d_samples, size_t n, int64_t divisor, const uint64_t* k,
                          int nk, int64_t* out, b200_lat_stats* stats, bool no_squares) {
    if (!k || !out || !stats || nk < 1 || divisor < 1 || !d_samples || ((uintptr_t)d_samples & 7))
        return fail(B200_ERR_INVALID, "b200_select_kth_stats");
    if (n == 0) return fail(B200_ERR_EMPTY, "b200_select_kth: no samples (the reference panics, ssd_test/main.go:157)");
    for (int i = 0; i < nk; i++)
        if (k[i] >= n) return fail(B200_ERR_INVALID, "b200_select_kth: k >= n");
    DevState* d = get_dev(ctx, dev);
    if (!d) return fail(B200_ERR_INVALID, "b200_select_kth: unknown device");
    {   // long inputs: moments and brackets in the same pass
        const int rc = try_sampled(ctx, dev, d_samples, n, divisor, k, nk, out, stats, no_squares);
        if (rc <= 0) return rc;
    }
    // otherwise one pass for the moments (a histogram with a single threshold), then the radix select between
    // the minimum and maximum that pass found
    LatMoments m;
    {
        HistParams P;
        const long long zero = 0;
        if (host_build_hist_params_thr(&zero, 1, divisor, false, &P)) return fail(B200_ERR_INVALID, "b200_select_kth_stats");
        std::lock_guard<std::mutex> lk(d->mu);
        CU(cudaSetDevice(dev));
        CU(launch_hist((const long long*)d_samples, n, d->lat, true, d->stream, P));
        CU(cudaMemcpyAsync(&m, d->lat.d_mom, sizeof m, cudaMemcpyDeviceToHost, d->stream));
        CU(cudaStreamSynchronize(d->stream));
    }
    fill_stats(m, divisor, stats);
    g_sampled_depth++;  // the sampled route had its chance above
    const int rc = b200_select_kth_range(ctx, dev, d_samples, n, divisor, m.min_ns, m.max_ns, k, nk, out);
    g_sampled_depth--;
    return rc;
}

}  // namespace

// ---- latency text codec + float64 histogram ----------------------------------------------------------------

extern "C" {

int b200_parse_latency_text(b200_ctx* ctx, int dev, const void* d_text, size_t nbytes, double* d_values,
                            uint8_t* d_status, size_t max_lines, uint64_t* nlines, uint64_t* first_bad_line,
                            uint64_t* nbad, uint64_t* ninexact) {
    DevState* d = get_dev(ctx, dev);
    if (!d || (nbytes && !d_text) || (max_lines && !d_values)) return fail(B200_ERR_INVALID, "b200_parse_latency_text");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    const size_t ntiles = text_tiles(nbytes);
    // scratch kept between calls (cudaMalloc + cudaFree of a few tens of MB cost more than the parse itself);
    // three words in front hold the summary
    const size_t need = 4 + text_scratch_words(nbytes);
    if (need > d->text_scratch_cap) {
        cudaFree(d->text_scratch);
        d->text_scratch = nullptr;
        d->text_scratch_cap = 0;
        const size_t room = need + need / 4;
        cudaError_t ea = cudaMalloc(&d->text_scratch, room * sizeof(unsigned long long));
        if (ea == cudaErrorMemoryAllocation) {
            cudaGetLastError();
            return fail(B200_ERR_NOMEM, "b200_parse_latency_text: scratch");
        }
        CU(ea);
        d->text_scratch_cap = room;
    }
    unsigned long long *summary = d->text_scratch, *tiles = d->text_scratch + 4;
    cudaError_t e = launch_text_parse((const unsigned char*)d_text, nbytes, d_values, d_status, max_lines, tiles, summary,
                                      d->stream);
    unsigned long long total = 0, sum3[3] = {~0ull, 0, 0};
    if (e == cudaSuccess) e = cudaMemcpyAsync(&total, tiles + ntiles, sizeof total, cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sum3, summary, sizeof sum3, cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(d->stream);
    if (e != cudaSuccess) return fail_cuda(e, "b200_parse_latency_text");
    if (nlines) *nlines = total;
    if (first_bad_line) *first_bad_line = sum3[0];
    if (nbad) *nbad = sum3[1];
    if (ninexact) *ninexact = sum3[2];
    if (total > max_lines) return fail(B200_ERR_TOO_LARGE, "b200_parse_latency_text: more lines than max_lines");
    return 0;
}

int b200_hist_fixed_f64(b200_ctx* ctx, int dev, const double* d_values, size_t n, const double* edges, int nedges,
                        int mode, uint64_t* counts, uint64_t* dropped_below, uint64_t* dropped_above,
                        uint64_t* nan_count, double* sum_hi, double* sum_lo) {
    DevState* d = get_dev(ctx, dev);
    if (!d || !edges || !counts || (n && !d_values) || (mode != B200_HIST_NUMPY && mode != B200_HIST_OPENCENSUS) ||
        nedges < (mode == B200_HIST_NUMPY ? 2 : 1) || nedges > B200_HIST_MAX_EDGES)
        return fail(B200_ERR_INVALID, "b200_hist_fixed_f64");
    for (int i = 0; i < nedges; i++)
        if (edges[i] != edges[i] || (i && !(edges[i] > edges[i - 1])))
            return fail(B200_ERR_INVALID, "b200_hist_fixed_f64: edges must be ascending");
    std::lock_guard<std::mutex> lk(d->mu);
    CU(cudaSetDevice(dev));
    const size_t nsum = 4 + 2 * (size_t)256 * hist_f64_grid();
    double* d_sums = nullptr;
    CU(cudaMalloc(&d_sums, nsum * sizeof(double)));
    std::vector<double> sums(nsum);
    unsigned long long c[kMaxThresholds + 1];
    cudaError_t e = launch_hist_f64(d_values, n, edges, nedges, mode, d->lat.d_counts, d_sums, d->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(c, d->lat.d_counts, sizeof c, cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sums.data(), d_sums, nsum * sizeof(double), cudaMemcpyDeviceToHost, d->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(d->stream);
    cudaFree(d_sums);
    if (e != cudaSuccess) return fail_cuda(e, "b200_hist_fixed_f64");
    uint64_t below = 0, above = 0;
    if (mode == B200_HIST_NUMPY) {
        for (int i = 0; i < nedges - 1; i++) counts[i] = c[i + 1];
        below = c[0];
        above = c[nedges];
    } else {
        for (int i = 0; i <= nedges; i++) counts[i] = c[i];
    }
    if (dropped_below) *dropped_below = below;
    if (dropped_above) *dropped_above = above;
    if (nan_count) {
        unsigned long long nn;
        memcpy(&nn, &sums[2], sizeof nn);
        *nan_count = nn;
    }
    // fold the per-thread double-double partial sums in a fixed order (two-sum), so the result is reproducible
    double hi = 0, lo = 0;
    for (size_t i = 4; i + 1 < nsum; i += 2) {
        double v = sums[i], s = hi + v, bb = s - hi;
        lo += (hi - (s - bb)) + (v - bb) + sums[i + 1];
        hi = s;
    }
    double s = hi + lo;
    if (sum_hi) *sum_hi = s;
    if (sum_lo) *sum_lo = lo - (s - hi);
    return 0;
}

}  // extern "C"

// ---- several devices driven by one process: per-device kernels, counts summed on the host ----

namespace {

// run f(i) for i in [0, n) on n host threads (device work of different GPUs overlaps); returns the first error
template <typename F>
int for_each_device(int n, F f) {
    std::vector<int> rc(n, 0);
    std::vector<std::thread> th;
    for (int i = 1; i < n; i++) th.emplace_back([&, i] { rc[i] = f(i); });
    rc[0] = f(0);
    for (auto& t : th) t.join();
    for (int r : rc)
        if (r) return r;
    return 0;
}

void add128(uint64_t& lo, uint64_t& hi, uint64_t alo, uint64_t ahi) {
    uint64_t s = lo + alo;
    hi += ahi + (s < lo);
    lo = s;
}

}  // namespace

extern "C" {

int b200_hist_fixed_multi(b200_ctx* ctx, const int* devs, const int64_t* const* d_samples, const size_t* n, int ndev,
                          int64_t divisor, const double* edges, int nedges, int mode, uint64_t* counts,
                          b200_lat_stats* stats) {
    if (!ctx || !devs || !d_samples || !n || ndev < 1 || !counts || nedges < 1 || nedges > B200_HIST_MAX_EDGES)
        return fail(B200_ERR_INVALID, "b200_hist_fixed_multi");
    const int nb = mode == B200_HIST_NUMPY ? nedges - 1 : nedges + 1;
    std::vector<std::vector<uint64_t>> c(ndev, std::vector<uint64_t>(B200_HIST_MAX_EDGES + 2));
    std::vector<b200_lat_stats> st(ndev);
    int rc = for_each_device(ndev, [&](int i) {
        return b200_hist_fixed(ctx, devs[i], d_samples[i], n[i], divisor, edges, nedges, mode, c[i].data(), &st[i]);
    });
    if (rc) return rc;
    for (int b = 0; b < nb; b++) {
        counts[b] = 0;
        for (int i = 0; i < ndev; i++) counts[b] += c[i][b];
    }
    if (stats) {
        b200_lat_stats t{};
        uint64_t sum_hi = 0;
        bool any = false;
        for (int i = 0; i < ndev; i++) {
            if (!st[i].count) continue;
            t.min = any ? std::min(t.min, st[i].min) : st[i].min;
            t.max = any ? std::max(t.max, st[i].max) : st[i].max;
            any = true;
            t.count += st[i].count;
            add128(t.sum, sum_hi, st[i].sum, (uint64_t)st[i].sum_hi);
            add128(t.sumsq_lo, t.sumsq_hi, st[i].sumsq_lo, st[i].sumsq_hi);
            t.dropped_below += st[i].dropped_below;
            t.dropped_above += st[i].dropped_above;
        }
        t.sum_hi = (int64_t)sum_hi;
        *stats = t;
    }
    return 0;
}

int b200_select_kth_multi(b200_ctx* ctx, const int* devs, const int64_t* const* d_samples, const size_t* n, int ndev,
                          int64_t divisor, const uint64_t* k, int nk, int64_t* out) {
    if (!ctx || !devs || !d_samples || !n || ndev < 1 || !k || !out || nk < 1 || divisor < 1)
        return fail(B200_ERR_INVALID, "b200_select_kth_multi");
    uint64_t total = 0;
    for (int i = 0; i < ndev; i++) total += n[i];
    if (total == 0) return fail(B200_ERR_EMPTY, "b200_select_kth_multi: no samples (the reference panics, ssd_test/main.go:157)");
    for (int i = 0; i < nk; i++)
        if (k[i] >= total) return fail(B200_ERR_INVALID, "b200_select_kth_multi: k >= n");
    std::vector<int64_t> mn(ndev), mx(ndev);
    int rc = for_each_device(ndev, [&](int i) { return b200_minmax(ctx, devs[i], d_samples[i], n[i], &mn[i], &mx[i]); });
    if (rc) return rc;
    int64_t gmn = INT64_MAX, gmx = INT64_MIN;
    for (int i = 0; i < ndev; i++)
        if (n[i]) { gmn = std::min(gmn, mn[i]); gmx = std::max(gmx, mx[i]); }
    const uint64_t kmin = (uint64_t)gmn ^ 0x8000000000000000ull, range = ((uint64_t)gmx ^ 0x8000000000000000ull) - kmin;
    int bits = 0;
    while (bits < 64 && (range >> bits)) bits++;
    std::vector<std::vector<uint64_t>> cnt(ndev, std::vector<uint64_t>((size_t)kMaxSelect * kRadixBins));
    for (int g = 0; g < nk; g += kMaxSelect) {
        const int m = std::min(kMaxSelect, nk - g);
        uint64_t prefix[kMaxSelect] = {0, 0, 0, 0}, rank[kMaxSelect];
        for (int i = 0; i < m; i++) rank[i] = k[g + i];
        int top = bits;
        bool first = true;
        while (top > 0) {
            const int shift = std::max(top - kRadixBits, 0), width = top - shift, nh = first ? 1 : m;
            rc = for_each_device(ndev, [&](int i) {
                return b200_radix_count(ctx, devs[i], d_samples[i], n[i], kmin, shift, top, nh, prefix, cnt[i].data());
            });
            if (rc) return rc;
            for (int i = 0; i < m; i++) {
                const size_t row = first ? 0 : (size_t)i * kRadixBins;
                uint64_t cum = 0;
                int dgt = 0;
                for (; dgt < (1 << width); dgt++) {
                    uint64_t h = 0;
                    for (int dv = 0; dv < ndev; dv++) h += cnt[dv][row + dgt];  // the host-side "all-reduce"
                    if (rank[i] < cum + h) break;
                    cum += h;
                }
                if (dgt == (1 << width)) return fail(B200_ERR_CUDA, "b200_select_kth_multi: inconsistent radix counts");
                rank[i] -= cum;
                prefix[i] = (prefix[i] << width) | (uint64_t)dgt;
            }
            top = shift;
            first = false;
        }
        for (int i = 0; i < m; i++) out[g + i] = (int64_t)((prefix[i] + kmin) ^ 0x8000000000000000ull) / divisor;
    }
    return 0;
}

}  // extern "C"
