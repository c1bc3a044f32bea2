// BASELINE config 1 as compiled code: W worker threads each re-read one 2 MiB in-memory object M times
// (main.go:119-156, 200-212), verifying CRC-32C at every EOF — once through the GPU-backed hash.Hash32
// (include/b200crc.hpp) and once with the CPU oracle (the C restatement of Go's SSE4.2 path) for comparison.
//   g++ -std=c++17 -O2 -pthread tools/readloop_bench.cpp -Iinclude -Ioracle -Lcustom-go-client-benchmark_b200 -lb200crc
//       -Loracle -loracle -Wl,-rpath,$PWD/custom-go-client-benchmark_b200 -Wl,-rpath,$PWD/oracle -o gpurun_out/readloop_bench
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "b200crc.hpp"
#include "oracle.h"

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
    const size_t kObj = 2u << 20, kRead = 2u << 20;
    const int reads = argc > 1 ? atoi(argv[1]) : 200;
    b200::Context ctx({0});
    for (int workers : {1, 4, 16, 48}) {
        std::vector<std::vector<uint8_t>> objs(workers, std::vector<uint8_t>(kObj));
        std::vector<uint32_t> want(workers);
        for (int w = 0; w < workers; w++) {
            oracle_fill_payload(objs[w].data(), (uint64_t)w * kObj / 8, kObj / 8, 0xC0FFEE);
            want[w] = oracle_crc32c_hw(0, objs[w].data(), kObj);
        }
        for (int mode = 0; mode < 2; mode++) {  // 0 = GPU hasher, 1 = CPU oracle
            std::atomic<int> bad{0};
            std::vector<std::unique_ptr<b200::crc32c::Hash32>> hs;
            if (mode == 0)
                for (int w = 0; w < workers; w++) hs.push_back(b200::crc32c::New(ctx));
            const double t0 = now();
            std::vector<std::thread> th;
            for (int w = 0; w < workers; w++)
                th.emplace_back([&, w] {
                    for (int i = 0; i < reads; i++) {
                        uint32_t got;
                        if (mode == 0) {
                            auto& h = *hs[w];
                            h.Reset();
                            for (size_t o = 0; o < kObj; o += kRead) h.Write(objs[w].data() + o, std::min(kRead, kObj - o));
                            got = h.Sum32();
                        } else {
                            got = 0;
                            for (size_t o = 0; o < kObj; o += kRead)
                                got = oracle_crc32c_hw(got, objs[w].data() + o, std::min(kRead, kObj - o));
                        }
                        if (got != want[w]) bad++;
                    }
                });
            for (auto& t : th) t.join();
            const double dt = now() - t0;
            std::printf("%2d workers x %d reads, %s: %8.2f GB/s  (%d bad)\n", workers, reads,
                        mode == 0 ? "GPU Hash32 (Write + Sum32 per read)" : "CPU oracle (SSE4.2)              ",
                        (double)workers * reads * kObj / dt / 1e9, bad.load());
        }
    }
    return 0;
}
