"""How the latency kernels behave when every sample is beyond the 32-bit fast path (reads of 5-100 s):
python tools/slowpath_timing.py [nsamples]"""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
ctx = pkg.crc32c.Context([0]); lat = pkg.latency
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
rng = np.random.default_rng(1)
for name, x in (("5-100 s", rng.integers(5 * 10**9, 10**11, n)), ("20-95 ms", rng.integers(2 * 10**7, 95 * 10**6, n)),
                ("half and half", np.where(rng.random(n) < 0.5, rng.integers(5 * 10**9, 10**11, n), rng.integers(2 * 10**7, 95 * 10**6, n)))):
    x = np.ascontiguousarray(x, dtype=np.int64)
    d = ctx.alloc_device(n * 8); ctx.h2d(d, x)
    ms = oracle_ms = None
    for f, label in ((lambda: lat.hist_opencensus(d, n, ctx=ctx), "oc35 + moments"), (lambda: lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx), "percentiles")):
        f(); t0 = time.perf_counter(); r = f(); dt = (time.perf_counter() - t0) * 1e3
        print(f"{name:14s} {label:16s} {dt:9.3f} ms  {n * 8 / dt / 1e6:8.1f} GB/s", flush=True)
    v = x // 10**6
    h = lat.hist_opencensus(d, n, ctx=ctx)
    ref = np.bincount(np.searchsorted(lat.OC_DEFAULT_LATENCY_BOUNDS, v.astype(np.float64), side="right"), minlength=35)
    assert list(h.counts) == list(ref) and h.sum == int(v.sum()) and h.min == int(v.min()), name
    ctx.free_device(d)
print("ok")
