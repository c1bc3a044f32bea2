"""One README histogram (with moments) and one bracket-style percentile call over n samples: a short program for ncu."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
ctx = crc32c.Context([0])
d = ctx.alloc_device(n * 8)
ctx.fill_latency_ns(d, n, 0xC0FFEE)
for _ in range(2):
    h = lat.hist_readme(d, n, ctx=ctx)
    o = lat.hist_opencensus(d, n, ctx=ctx)
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
print(h.counts.sum(), o.counts.sum(), p)
