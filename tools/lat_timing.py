"""Times the latency kernels on one GPU: python tools/lat_timing.py [nsamples]"""
import ctypes as C, importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
n = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000_000
ctx = crc32c.Context([0])
d = ctx.alloc_device(n * 8)
ctx.fill_latency_ns(d, n, 0xC0FFEE)
L, h = ctx._L, ctx._h


def timeit(f, reps=int(os.environ.get("LAT_REPS", "5"))):  # LAT_REPS=1 under ncu
    f()
    t0 = time.perf_counter()
    for _ in range(reps):
        f()
    return (time.perf_counter() - t0) / reps * 1e3


edges = lat.README_EDGES
counts = np.zeros(40, dtype=np.uint64)
st = pkg._lib.LatStats()
ms = timeit(lambda: L.b200_hist_fixed(h, 0, d, n, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, C.byref(st)))
print(f"hist numpy15 + moments : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s")
ms = timeit(lambda: L.b200_hist_fixed(h, 0, d, n, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, None))
print(f"hist numpy15 only      : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s")
oc = lat.OC_DEFAULT_LATENCY_BOUNDS
ms = timeit(lambda: L.b200_hist_fixed(h, 0, d, n, 1000000, oc.ctypes.data, oc.size, 1, counts.ctypes.data, C.byref(st)))
print(f"hist oc35 + moments    : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s")
ms = timeit(lambda: lat.hist_readme_and_opencensus(d, n, ctx=ctx))
print(f"numpy15 + oc35 + moments, one pass: {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s")
mn, mx = C.c_int64(), C.c_int64()
ms = timeit(lambda: L.b200_minmax(h, 0, d, n, C.byref(mn), C.byref(mx)))
print(f"minmax                 : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s")
ks = np.array([n // 5, n // 2, 9 * n // 10, 99 * n // 100], dtype=np.uint64)
out = np.zeros(4, dtype=np.int64)
ms = timeit(lambda: L.b200_select_kth(h, 0, d, n, 1000, ks.ctypes.data, 4, out.ctypes.data))
print(f"select 4 ranks         : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s   {out}")
ms = timeit(lambda: lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx))
print(f"percentiles (7 lines)  : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s   {lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)}")
# radix building blocks
kmin = (mn.value + (1 << 63)) & ((1 << 64) - 1)
bits = (mx.value - mn.value).bit_length()
cnt = np.zeros(4 * 4096, dtype=np.uint64)
pf = np.zeros(4, dtype=np.uint64)
ms = timeit(lambda: L.b200_radix_count(h, 0, d, n, kmin, bits - 12, bits, 1, pf.ctypes.data, cnt.ctypes.data))
print(f"radix pass 1 (1 histogram, all samples) : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s   bits={bits}")
pf[:] = [19, 23, 29, 38]
ms = timeit(lambda: L.b200_radix_count(h, 0, d, n, kmin, bits - 24, bits - 12, 4, pf.ctypes.data, cnt.ctypes.data))
print(f"radix pass 2 (4 prefixes, full data)    : {ms:8.3f} ms  {n*8/ms/1e6:8.1f} GB/s")
