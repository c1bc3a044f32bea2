// b200crc.hpp — C++17 host mirror of the Go surface the reference's hot path uses, over the C ABI in b200crc.h.
//
// The reference is Go (compiled code) and no Go toolchain exists in this image or on the GPU box, so this
// header is the compiled-language host side: same names, argument meaning and error behaviour as the Go
// identifiers it stands for, so that code reads like the reference's call sites.
//
//   crc32c::New()            crc32.New(crc32.MakeTable(crc32.Castagnoli))  -> hash.Hash32
//   crc32c::Update/Checksum  crc32.Update / crc32.Checksum, the statement storage.Reader.Read runs per chunk
//                            (reached from main.go:140)
//   crc32c::VerifyObjects    the per-object EOF check "storage: bad CRC on read" (main.go:141-143), batched
//   crc32c::Combine          CRC of A||B from the parts (GF(2) on the host)
//   latency::*               README.md:15-35 histogram/mean, metrics_exporter.go:22-34 distribution,
//                            benchmark-script/ssd_test/main.go:147-163 percentiles
// Errors: Go returns (value, error); here a failing C call throws b200::Error carrying the b200_status.
// hash.Hash.Write "never returns an error" in Go — a CUDA failure there throws as well (Go's binding panics).
// Header only; link with -lb200crc.
#ifndef B200CRC_HPP
#define B200CRC_HPP

#include <cstdint>
#include <cstring>
#include <istream>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "b200crc.h"

namespace b200 {

struct Error : std::runtime_error {
    int status;
    Error(int s, const std::string& what) : std::runtime_error(what), status(s) {}
};

inline void check(int rc) {
    if (rc != B200_OK) throw Error(rc, std::string(b200_strerror(rc)) + ": " + b200_last_error());
}

// Owns the library context for a set of devices (streams, tables, scratch).
class Context {
public:
    explicit Context(std::vector<int> devices = {0}) : devices_(std::move(devices)) {
        check(b200_init((int)devices_.size(), devices_.data(), &ctx_));
    }
    ~Context() { b200_shutdown(ctx_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    b200_ctx* get() const { return ctx_; }
    int device(int i = 0) const { return devices_.at(i); }
    int ndev() const { return (int)devices_.size(); }

    void* AllocDevice(size_t n, int dev) const { void* p = nullptr; check(b200_alloc_device(ctx_, dev, n, &p)); return p; }
    void FreeDevice(void* p, int dev) const { check(b200_free_device(ctx_, dev, p)); }
    // the replacement for ReadObject's `buf := make([]byte, 2*MB)` (main.go:125)
    void* AllocPinned(size_t n) const { void* p = nullptr; check(b200_alloc_pinned(ctx_, n, &p)); return p; }
    void FreePinned(void* p) const { check(b200_free_pinned(ctx_, p)); }
    void H2D(void* d, const void* h, size_t n, int dev) const { check(b200_memcpy_h2d(ctx_, dev, d, h, n)); }
    void D2H(void* h, const void* d, size_t n, int dev) const { check(b200_memcpy_d2h(ctx_, dev, h, d, n)); }

private:
    std::vector<int> devices_;
    b200_ctx* ctx_ = nullptr;
};

namespace crc32c {

constexpr uint32_t Castagnoli = 0x82f63b78;  // crc32.Castagnoli
constexpr int Size = 4;                      // crc32.Size

// storage.Reader: "storage: bad CRC on read: got %d, want %d"
struct BadCRC : std::runtime_error {
    size_t index;
    uint32_t got, want;
    BadCRC(size_t i, uint32_t g, uint32_t w)
        : std::runtime_error("storage: bad CRC on read: got " + std::to_string(g) + ", want " + std::to_string(w)),
          index(i), got(g), want(w) {}
};

inline uint32_t Update(const Context& c, uint32_t crc, const void* p, size_t n, int dev = -1) {
    uint32_t out = 0;
    check(b200_crc32c_update_host(c.get(), dev < 0 ? c.device() : dev, crc, p, n, &out));
    return out;
}
inline uint32_t Checksum(const Context& c, const void* p, size_t n) { return Update(c, 0, p, n); }
inline uint32_t UpdateDevice(const Context& c, uint32_t crc, const void* dptr, size_t n, int dev = -1) {
    uint32_t out = 0;
    check(b200_crc32c_update_device(c.get(), dev < 0 ? c.device() : dev, crc, dptr, n, &out));
    return out;
}
inline uint32_t Combine(uint32_t crc1, uint32_t crc2, uint64_t len2) { return b200_crc32c_combine(crc1, crc2, len2); }

// hash.Hash32
class Hash32 {
public:
    explicit Hash32(const Context& c, int dev = -1) { check(b200_hasher_new(c.get(), dev < 0 ? c.device() : dev, &h_)); }
    ~Hash32() { b200_hasher_free(h_); }
    Hash32(const Hash32&) = delete;
    Hash32& operator=(const Hash32&) = delete;

    size_t Write(const void* p, size_t n) { check(b200_hasher_write(h_, p, n)); return n; }  // n == len(p), no error
    uint32_t Sum32() { uint32_t v = 0; check(b200_hasher_sum32(h_, &v)); return v; }
    // appends the big-endian checksum to b; does not change the state
    std::vector<uint8_t> Sum(std::vector<uint8_t> b = {}) {
        uint32_t s = Sum32();
        b.push_back(uint8_t(s >> 24)); b.push_back(uint8_t(s >> 16)); b.push_back(uint8_t(s >> 8)); b.push_back(uint8_t(s));
        return b;
    }
    void Reset() { check(b200_hasher_reset(h_)); }
    int Size() const { return crc32c::Size; }
    int BlockSize() const { return 1; }
    // io.ReaderFrom: io.Copy(h, rc) in place of io.CopyBuffer(io.Discard, rc, buf) (main.go:140); the stream is
    // read straight into the hasher's pinned staging buffer
    int64_t ReadFrom(std::istream& r) {
        int64_t total = 0;
        for (;;) {
            void* p = nullptr;
            size_t avail = 0;
            check(b200_hasher_reserve(h_, &p, &avail));
            r.read(static_cast<char*>(p), (std::streamsize)avail);
            std::streamsize n = r.gcount();
            if (n > 0) { check(b200_hasher_commit(h_, (size_t)n)); total += n; }
            if (n < (std::streamsize)avail) return total;  // EOF
        }
    }

private:
    b200_hasher* h_ = nullptr;
};
inline std::unique_ptr<Hash32> New(const Context& c, int dev = -1) { return std::make_unique<Hash32>(c, dev); }

struct Buf {
    const void* data;
    size_t len;
};

struct VerifyResult {
    std::vector<size_t> bad;    // indices of mismatching objects; a mismatch never aborts the batch
    std::vector<uint32_t> got;
};

// VerifyObjects([]buf, []want) on host buffers
inline VerifyResult VerifyObjects(const Context& c, const std::vector<Buf>& bufs, const std::vector<uint32_t>& want,
                                  int dev = -1) {
    if (!want.empty() && want.size() != bufs.size()) throw Error(B200_ERR_INVALID, "len(bufs) != len(want)");
    std::vector<const void*> ptrs(bufs.size());
    std::vector<uint64_t> lens(bufs.size());
    for (size_t i = 0; i < bufs.size(); i++) { ptrs[i] = bufs[i].data; lens[i] = bufs[i].len; }
    VerifyResult r;
    r.got.resize(bufs.size());
    std::vector<uint8_t> ok(bufs.size());
    uint64_t nbad = 0;
    check(b200_crc32c_batch_host(c.get(), dev < 0 ? c.device() : dev, ptrs.data(), lens.data(), bufs.size(),
                                 want.empty() ? nullptr : want.data(), r.got.data(), ok.data(), &nbad));
    for (size_t i = 0; i < ok.size(); i++)
        if (!ok[i]) r.bad.push_back(i);
    return r;
}

// the same on a device-resident arena (BASELINE config 2)
inline VerifyResult VerifyObjectsDevice(const Context& c, const void* base, const std::vector<uint64_t>& off,
                                        const std::vector<uint64_t>& len, const std::vector<uint32_t>& want, int dev = -1) {
    VerifyResult r;
    r.got.resize(off.size());
    std::vector<uint8_t> ok(off.size());
    uint64_t nbad = 0;
    check(b200_crc32c_batch_device(c.get(), dev < 0 ? c.device() : dev, base, off.data(), len.data(), off.size(),
                                   want.empty() ? nullptr : want.data(), r.got.data(), ok.data(), &nbad));
    for (size_t i = 0; i < ok.size(); i++)
        if (!ok[i]) r.bad.push_back(i);
    return r;
}

// one object sharded over the context's devices in message order (BASELINE config 3)
inline uint32_t ShardedChecksum(const Context& c, const std::vector<const void*>& dptrs, const std::vector<uint64_t>& lens) {
    uint32_t out = 0;
    check(b200_crc32c_sharded(c.get(), dptrs.data(), lens.data(), (int)dptrs.size(), &out));
    return out;
}

}  // namespace crc32c

namespace latency {

constexpr int64_t kNsPerMs = 1000000, kNsPerUs = 1000;

inline std::vector<double> ReadmeEdges() {  // README.md:21-23: range(20, 100, 5)
    std::vector<double> e;
    for (int x = 20; x < 100; x += 5) e.push_back(x);
    return e;
}
inline std::vector<double> OcDefaultLatencyBounds() {  // ochttp.DefaultLatencyDistribution
    return {1, 2, 3, 4, 5, 6, 8, 10, 13, 16, 20, 25, 30, 40, 50, 65, 80, 100, 130, 160, 200, 250, 300, 400, 500, 650,
            800, 1000, 2000, 5000, 10000, 20000, 50000, 100000};
}

struct Distribution {
    std::vector<uint64_t> counts;
    b200_lat_stats stats{};
};

inline Distribution HistFixed(const Context& c, const int64_t* d_samples_ns, size_t n, int64_t divisor,
                              const std::vector<double>& edges, int mode, int dev = -1) {
    Distribution d;
    d.counts.resize(mode == B200_HIST_NUMPY ? edges.size() - 1 : edges.size() + 1);
    check(b200_hist_fixed(c.get(), dev < 0 ? c.device() : dev, d_samples_ns, n, divisor, edges.data(), (int)edges.size(),
                          mode, d.counts.data(), &d.stats));
    return d;
}

// Two histograms of the same samples from one pass (README bins + OpenCensus buckets of one run).
inline std::pair<Distribution, Distribution> HistFixedPair(const Context& c, const int64_t* d_samples_ns, size_t n,
                                                           int64_t divisor, const std::vector<double>& edges_a, int mode_a,
                                                           const std::vector<double>& edges_b, int mode_b, int dev = -1) {
    Distribution a, b;
    a.counts.resize(mode_a == B200_HIST_NUMPY ? edges_a.size() - 1 : edges_a.size() + 1);
    b.counts.resize(mode_b == B200_HIST_NUMPY ? edges_b.size() - 1 : edges_b.size() + 1);
    uint64_t dropped_b[2] = {0, 0};
    check(b200_hist_fixed_pair(c.get(), dev < 0 ? c.device() : dev, d_samples_ns, n, divisor, edges_a.data(),
                               (int)edges_a.size(), mode_a, a.counts.data(), edges_b.data(), (int)edges_b.size(), mode_b,
                               b.counts.data(), dropped_b, &a.stats));
    b.stats = a.stats;
    b.stats.dropped_below = dropped_b[0];
    b.stats.dropped_above = dropped_b[1];
    return {a, b};
}

struct Percentiles {  // the seven lines ssd_test prints (ssd_test/main.go:157-163), in sample units
    int64_t avg, p20, p50, p90, p99, min, max;
};

inline Percentiles ComputePercentiles(const Context& c, const int64_t* d_samples_ns, size_t n, int64_t divisor = kNsPerUs,
                                      int dev = -1) {
    if (n == 0) throw Error(B200_ERR_EMPTY, "percentiles of an empty sample set (Go panics at ssd_test/main.go:157)");
    const uint64_t size = n;
    uint64_t ks[4] = {size / 5, size / 2, (9 * size) / 10, (99 * size) / 100};
    int64_t out[4];
    b200_lat_stats st;
    check(b200_select_kth_avg(c.get(), dev < 0 ? c.device() : dev, d_samples_ns, n, divisor, ks, 4, out, &st));
    const int64_t sum = (int64_t)st.sum;  // Go's wrapping int64 sum
    return {sum / (int64_t)size, out[0], out[1], out[2], out[3], st.min, st.max};
}

}  // namespace latency

}  // namespace b200
#endif
