"""BASELINE config 1 through the GPU hasher: N workers x M reads of 2 MiB in-memory objects."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
pkg = importlib.import_module("custom-go-client-benchmark_b200")
c, rl = pkg.crc32c, pkg.readloop
ctx = c.Context([0])
MB = 1 << 20
for workers, reads in ((1, 200), (8, 100), (48, 40)):
    objs = []
    for w in range(workers):
        data = oracle.payload(2 * MB, first_word=w * 2 * MB // 8)
        objs.append(rl.InMemoryObject(data, oracle.crc32c(data)))
    rl.RunReadLoop(objs, 2, ctx)
    t0 = time.perf_counter()
    res = rl.RunReadLoop(objs, reads, ctx)
    dt = time.perf_counter() - t0
    assert not res.errors
    print(f"{workers:3d} workers x {reads:4d} reads: {res.bytes_read/dt/1e9:7.2f} GB/s, {dt/res.reads*1e6*workers:8.1f} us per read per worker, "
          f"median recorded latency {np.median(res.latencies_ns)/1e3:.0f} us")
# raw hasher: one 64 MiB Write + Sum32, and 2 MiB writes
h = c.New(ctx)
buf = oracle.payload(64 * MB)
for chunk in (64 * MB, 2 * MB, 64 * 1024):
    h.Reset(); t0 = time.perf_counter()
    for rep in range(4):
        for o in range(0, buf.size, chunk):
            h.Write(buf[o:o + chunk])
    s = h.Sum32(); dt = time.perf_counter() - t0
    print(f"Hash32.Write in {chunk>>10:6d} KiB pieces: {4*buf.size/dt/1e9:6.2f} GB/s")
