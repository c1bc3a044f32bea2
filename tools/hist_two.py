"""README histogram with and without moments over n samples: a short program for ncu."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000_000
ctx = crc32c.Context([0])
d = ctx.alloc_device(n * 8)
ctx.fill_latency_ns(d, n, 0xC0FFEE)
L, h = ctx._L, ctx._h
edges = lat.README_EDGES
counts = np.zeros(40, dtype=np.uint64)
st = pkg._lib.LatStats()
for _ in range(2):
    L.b200_hist_fixed(h, 0, d, n, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, C.byref(st))
    L.b200_hist_fixed(h, 0, d, n, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, None)
print(counts[:15].sum())
