// b200_crc32c_update_device_noalloc — CRC-32C of one device buffer for callers that must not allocate.
//
// Who this is for: a process that handed all of its device memory to its own pool at start-up (an inference
// server, a storage daemon with a fixed arena), so a later cudaMalloc either fails or eats into memory it has
// promised elsewhere; and read loops that cannot tolerate the device-wide synchronisation cudaMalloc and
// cudaFree imply.  The ordinary entry points grow their per-segment scratch with cudaMalloc on demand; this
// one keeps every byte it needs — the constant tables and the per-segment scratch of the row kernel — in the
// fixed array below, which is part of the loaded module.  Every GPU that loads libb200crc.so carries the array
// (about 270 KB) until the process exits, whether or not this entry point is ever called.
//
// Consequences the caller has to know:
//   * Size limit: B200_NOALLOC_MAX_BYTES (32 GiB) per call; larger inputs are rejected with
//     B200_ERR_TOO_LARGE.  Split them and fold the pieces with b200_crc32c_combine.
//   * One array per GPU, shared by every call in the process.  The library serialises calls with a mutex and
//     each call waits for its own result before returning, so calls made through this entry point never
//     overlap on the GPU.  That serialisation is what keeps them correct: two pieces of GPU work using the
//     array at the same time would race on it.
//   * The array keeps whatever the previous call left in it.  Nothing relies on old contents: the mismatch
//     counter is zeroed on every call and each per-segment slot is written before it is read.
//   * The first call on a given GPU also performs the one-time set-up for that GPU (it looks up the device
//     address of the array and copies the constant tables into it).  Make that first call before you begin
//     any stream capture.
//   * Work is issued on the legacy default stream of the device — a pageable-source cudaMemcpyAsync, the
//     kernels, then a blocking cudaMemcpy of the result — and the call returns after the result has been
//     copied back.  Legacy-stream work is not allowed while ANY thread of the process has a global-mode stream
//     capture open, so no call at all (not only the first) may overlap such a capture: it would fail and
//     invalidate the capture.
//   * The calling thread's current device is put back to what it was before the call.
#include <cstring>
#include <mutex>

#include "b200crc.h"
#include "crc_kernels.h"
#include "lat_kernels.h"

using namespace b200;

namespace {

constexpr size_t kOffAdvRow = 0;          // [1024] u32  A^512 slicing tables
constexpr size_t kOffAdvSeg = 4096;       // [1024] u32  per-call segment shift tables
constexpr size_t kOffColK = 8192;         // [128]  u32
constexpr size_t kOffXsq = 8704;          // [64]   u32
constexpr size_t kOffByteTab = 8960;      // [256]  u32
constexpr size_t kOffGot = 9984;          // u32
constexpr size_t kOffOk = 10000;          // u8
constexpr size_t kOffNbad = 10016;        // u64
constexpr size_t kOffAdv4 = 10240;        // [1024] u32  x^32 slicing tables (row kernel's segment epilogue)
constexpr size_t kOffSegRaw = 14336;      // [kMaxTasks] u32
constexpr uint32_t kMaxTasks = 65536;
constexpr size_t kScratchBytes = kOffSegRaw + (size_t)kMaxTasks * sizeof(uint32_t);
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for the runtime; it is NOT a device address.  The address the
// kernels use comes from cudaGetSymbolAddress in setup_device() below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

std::mutex g_mu;
unsigned char* g_dev_ptr[kMaxDevices] = {};
int g_grid[kMaxDevices] = {};

int setup_device(int dev) {  // caller holds g_mu and has made `dev` current
    if (g_dev_ptr[dev]) return B200_OK;
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    // the row kernel addresses its tables at absolute shared addresses: check once that this driver's dynamic
    // shared-memory window starts where that layout assumes (8 bytes of the array serve as the probe's output)
    e = smem_base_probe((unsigned long long*)(p + kOffNbad));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    uint32_t adv_row[1024], col_k[128], xsq[64], byte_tab[256], adv4[1024];
    host_build_adv_table(kRowBytes, adv_row);
    host_build_adv_table(4, adv4);
    host_build_col_k(col_k);
    host_build_xpow8_sq(xsq);
    host_build_byte_tab(byte_tab);
    e = cudaMemcpy(p + kOffAdvRow, adv_row, sizeof adv_row, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p + kOffColK, col_k, sizeof col_k, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p + kOffXsq, xsq, sizeof xsq, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p + kOffByteTab, byte_tab, sizeof byte_tab, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p + kOffAdv4, adv4, sizeof adv4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    g_grid[dev] = rows_kernel_grid(dev);
    g_dev_ptr[dev] = p;
    return B200_OK;
}

}  // namespace

extern "C" int b200_crc32c_update_device_noalloc(int dev, uint32_t crc, const void* dptr, size_t n, uint32_t* out) {
    if (!out || (n && !dptr) || dev < 0 || dev >= kMaxDevices) return B200_ERR_INVALID;
    if (n > B200_NOALLOC_MAX_BYTES) return B200_ERR_TOO_LARGE;
    std::lock_guard<std::mutex> lk(g_mu);
    struct RestoreDevice {  // leave the calling thread's current device as we found it
        int prev = -1;
        ~RestoreDevice() {
            if (prev >= 0) cudaSetDevice(prev);
            cudaGetLastError();
        }
    } restore;
    if (cudaGetDevice(&restore.prev) != cudaSuccess) restore.prev = -1;
    if (cudaSetDevice(dev) != cudaSuccess) {
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    if (int rc = setup_device(dev)) return rc;
    unsigned char* p = g_dev_ptr[dev];

    const uintptr_t addr = (uintptr_t)dptr, base = addr & ~(uintptr_t)15;
    BatchArgs a{};
    a.base = (const uint8_t*)base;
    a.single_off = addr - base;
    a.single_len = n;
    const ObjectCut cut = cut_object(addr, n);
    // same segment-length rule as the allocating path, then widened until the fixed scratch suffices
    uint64_t target = cut.rows / ((uint64_t)g_grid[dev] * kRowsWarps * 6);
    uint32_t rps = 16;
    while ((uint64_t)rps * 2 <= target && rps < 1024) rps *= 2;
    uint64_t ntasks = (cut.rows + rps - 1) / rps;
    while (ntasks > kMaxTasks) {
        rps *= 2;
        ntasks = (cut.rows + rps - 1) / rps;
    }
    uint32_t adv_seg[1024];
    host_build_adv_table((uint64_t)rps * kRowBytes, adv_seg);
    // pageable source: the runtime has taken its copy of adv_seg by the time this returns
    cudaError_t e = cudaMemcpyAsync(p + kOffAdvSeg, adv_seg, sizeof adv_seg, cudaMemcpyHostToDevice, 0);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    a.rows_per_seg = rps;
    a.ntasks = (uint32_t)ntasks;
    a.rows_per_seg2 = 0;  // one segment length
    a.adv_seg2 = nullptr;
    a.nfull_tasks = (uint32_t)(cut.rows / rps);
    a.nshort_tasks = 0;
    a.npartial_tasks = (uint32_t)(ntasks - cut.rows / rps);
    a.no_fold = 1;        // everything on the row kernel
    a.nobj = 1;
    a.seg_raw = (uint32_t*)(p + kOffSegRaw);
    a.got = (uint32_t*)(p + kOffGot);
    a.ok = p + kOffOk;
    a.nbad = (unsigned long long*)(p + kOffNbad);  // zeroed by launch_batch on every call
    a.adv_seg = (const uint32_t*)(p + kOffAdvSeg);
    a.init_crc = crc;
    DeviceTables t{};
    t.adv_row = (uint32_t*)(p + kOffAdvRow);
    t.col_k = (uint32_t*)(p + kOffColK);
    t.xpow8_sq = (uint32_t*)(p + kOffXsq);
    t.byte_tab = (uint32_t*)(p + kOffByteTab);
    t.adv4 = (uint32_t*)(p + kOffAdv4);
    e = launch_batch(a, t, g_grid[dev], 0, nullptr, nullptr);
    if (e == cudaSuccess) e = cudaMemcpy(out, p + kOffGot, sizeof(uint32_t), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    return B200_OK;
}
