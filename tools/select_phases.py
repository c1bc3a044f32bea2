"""Where a seven-line percentiles call spends its time: the sampled select's own trace (B200_SELECT_TRACE=1), and the
resident radix select alone on inputs the size of the sample and of the kept samples."""
import ctypes as C, importlib, os, sys, time
import numpy as np
os.environ.setdefault("B200_SELECT_TRACE", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
n = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
ctx = crc32c.Context([0])
if os.environ.get("B200_L2_FETCH"):  # experiment: the device's L2 fetch granularity (bytes)
    from cuda import cudart
    print("set limit:", cudart.cudaDeviceSetLimit(cudart.cudaLimit.cudaLimitMaxL2FetchGranularity, int(os.environ["B200_L2_FETCH"])),
          cudart.cudaDeviceGetLimit(cudart.cudaLimit.cudaLimitMaxL2FetchGranularity))
d = ctx.alloc_device(n * 8)
ctx.fill_latency_ns(d, n, 0xC0FFEE)
for _ in range(2):
    lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
t0 = time.perf_counter()
for _ in range(reps):
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
print(f"percentiles({n}): {(time.perf_counter() - t0) / reps * 1e3:.3f} ms", p, flush=True)
if len(sys.argv) > 3:
    sys.exit(0)
for m, nk in ((n // 120, 8), (6_000_000, 4), (1 << 20, 8)):
    ks = [m * (i + 1) // (nk + 1) for i in range(nk)]
    lat.select_kth(d, m, 1, ks, ctx=ctx)
    t0 = time.perf_counter()
    for _ in range(20):
        lat.select_kth(d, m, 1, ks, ctx=ctx)
    print(f"resident select of {nk} ranks among {m}: {(time.perf_counter() - t0) / 20 * 1e3:.3f} ms", flush=True)
