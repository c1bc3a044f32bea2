// Drives the C++ host mirror (include/b200crc.hpp) the way the reference's Go code drives hash/crc32 and its
// latency reduction, and checks every result against the CPU oracle (test infrastructure; linked here only).
#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <string>
#include <vector>

#include "b200crc.hpp"
#include "oracle.h"

#define REQUIRE(cond)                                                         \
    do {                                                                      \
        if (!(cond)) {                                                        \
            std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            return 1;                                                         \
        }                                                                     \
    } while (0)

int main() {
    using namespace b200;
    Context ctx({0});

    // crc32.Checksum / Update known answers
    REQUIRE(crc32c::Checksum(ctx, "123456789", 9) == 0xE3069283u);
    REQUIRE(crc32c::Checksum(ctx, "", 0) == 0u);
    std::vector<uint8_t> data(5u << 20);
    oracle_fill_payload(data.data(), 0, data.size() / 8, 0xC0FFEE);
    const uint32_t whole = oracle_crc32c_hw(0, data.data(), data.size());
    REQUIRE(crc32c::Checksum(ctx, data.data(), data.size()) == whole);
    uint32_t run = 0;
    for (size_t o = 0; o < data.size(); o += 777777) run = crc32c::Update(ctx, run, data.data() + o, std::min<size_t>(777777, data.size() - o));
    REQUIRE(run == whole);
    REQUIRE(crc32c::Combine(oracle_crc32c_hw(0, data.data(), 1000), oracle_crc32c_hw(0, data.data() + 1000, data.size() - 1000),
                            data.size() - 1000) == whole);

    // hash.Hash32
    auto h = crc32c::New(ctx);
    REQUIRE(h->Size() == 4 && h->BlockSize() == 1 && h->Sum32() == 0);
    REQUIRE(h->Write(data.data(), 10) == 10);
    h->Write(data.data() + 10, data.size() - 10);
    REQUIRE(h->Sum32() == whole && h->Sum32() == whole);
    auto sum = h->Sum({0xAB});
    REQUIRE(sum.size() == 5 && sum[0] == 0xAB && sum[1] == (whole >> 24) && sum[4] == (whole & 0xFF));
    h->Reset();
    std::istringstream in(std::string(reinterpret_cast<const char*>(data.data()), data.size()));
    REQUIRE(h->ReadFrom(in) == (int64_t)data.size());
    REQUIRE(h->Sum32() == whole);

    // VerifyObjects: flags exactly the corrupt object, never aborts
    const size_t nobj = 20, sz = 256 * 1024;
    std::vector<crc32c::Buf> bufs;
    std::vector<uint32_t> want;
    for (size_t i = 0; i < nobj; i++) {
        bufs.push_back({data.data() + i * sz, sz});
        want.push_back(oracle_crc32c_hw(0, data.data() + i * sz, sz));
    }
    auto ok = crc32c::VerifyObjects(ctx, bufs, want);
    REQUIRE(ok.bad.empty() && ok.got == want);
    want[7] ^= 1;
    auto bad = crc32c::VerifyObjects(ctx, bufs, want);
    REQUIRE(bad.bad.size() == 1 && bad.bad[0] == 7);

    // device-resident batch + sharded single object
    void* d = ctx.AllocDevice(data.size(), 0);
    ctx.H2D(d, data.data(), data.size(), 0);
    std::vector<uint64_t> off, len;
    for (size_t i = 0; i < nobj; i++) { off.push_back(i * sz + 3); len.push_back(sz - 5); }
    std::vector<uint32_t> w2;
    for (size_t i = 0; i < nobj; i++) w2.push_back(oracle_crc32c_hw(0, data.data() + off[i], len[i]));
    auto dv = crc32c::VerifyObjectsDevice(ctx, d, off, len, w2);
    REQUIRE(dv.bad.empty() && dv.got == w2);
    REQUIRE(crc32c::ShardedChecksum(ctx, {d}, {data.size()}) == whole);
    REQUIRE(crc32c::UpdateDevice(ctx, 0, d, data.size()) == whole);
    ctx.FreeDevice(d, 0);

    // latency: README histogram, OpenCensus buckets, ssd_test percentiles
    const size_t n = 300001;
    std::vector<int64_t> ns(n), ms(n), us(n);
    oracle_fill_latency_ns(ns.data(), 0, n, 0xC0FFEE);
    oracle_ns_to_ms(ns.data(), n, ms.data());
    oracle_ns_to_us(ns.data(), n, us.data());
    int64_t* dns = static_cast<int64_t*>(ctx.AllocDevice(n * 8, 0));
    ctx.H2D(dns, ns.data(), n * 8, 0);
    std::vector<double> msf(ms.begin(), ms.end());
    auto edges = latency::ReadmeEdges();
    std::vector<uint64_t> ref(edges.size() - 1);
    uint64_t dropped = oracle_hist_numpy(msf.data(), n, edges.data(), (int)edges.size(), ref.data());
    auto hist = latency::HistFixed(ctx, dns, n, latency::kNsPerMs, edges, B200_HIST_NUMPY);
    REQUIRE(hist.counts == ref && hist.stats.dropped_below + hist.stats.dropped_above == dropped);
    auto bounds = latency::OcDefaultLatencyBounds();
    std::vector<uint64_t> ocref(bounds.size() + 1);
    oracle_hist_opencensus(msf.data(), n, bounds.data(), (int)bounds.size(), ocref.data());
    REQUIRE(latency::HistFixed(ctx, dns, n, latency::kNsPerMs, bounds, B200_HIST_OPENCENSUS).counts == ocref);
    auto both = latency::HistFixedPair(ctx, dns, n, latency::kNsPerMs, edges, B200_HIST_NUMPY, bounds, B200_HIST_OPENCENSUS);
    REQUIRE(both.first.counts == hist.counts);
    REQUIRE(both.second.counts == ocref);
    REQUIRE(both.first.stats.dropped_below == hist.stats.dropped_below);
    REQUIRE(both.second.stats.sum == hist.stats.sum);
    oracle_percentiles op;
    REQUIRE(oracle_percentiles_i64(us.data(), n, &op) == 0);
    auto p = latency::ComputePercentiles(ctx, dns, n);
    REQUIRE(p.avg == op.avg && p.p20 == op.p20 && p.p50 == op.p50 && p.p90 == op.p90 && p.p99 == op.p99 &&
            p.min == op.min && p.max == op.max);
    bool threw = false;
    try { latency::ComputePercentiles(ctx, dns, 0); } catch (const Error& e) { threw = e.status == B200_ERR_EMPTY; }
    REQUIRE(threw);
    ctx.FreeDevice(dns, 0);
    std::puts("host mirror ok");
    return 0;
}
