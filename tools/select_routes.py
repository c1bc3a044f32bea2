"""Times b200_select_kth (4 ranks) by route and input length: python tools/select_routes.py
Run once per route with B200_SELECT_SAMPLED_MIN / B200_SELECT_RESIDENT_MAX set (they are read once per process)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
ctx = pkg.crc32c.Context([0]); lat = pkg.latency
nmax = 1 << 28
d = ctx.alloc_device(nmax * 8)
ctx.fill_latency_ns(d, nmax, 0xC0FFEE)
for lg in range(16, 29, 2):
    n = 1 << lg
    ks = [n // 5, n // 2, 9 * n // 10, 99 * n // 100]
    lat.select_kth(d, n, 1000, ks, ctx=ctx)
    t0 = time.perf_counter()
    for _ in range(5):
        out = lat.select_kth(d, n, 1000, ks, ctx=ctx)
    print(f"2^{lg}: {(time.perf_counter() - t0) / 5 * 1e3:8.3f} ms  {out}", flush=True)
