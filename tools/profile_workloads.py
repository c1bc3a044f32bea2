"""One short pass over every kernel of the library, for ncu (round-2 profiles): python tools/profile_workloads.py
  1 M objects of 4 KiB (row kernel + thread-per-object finalize), a ragged batch (both fold lengths + row kernel +
  finalize), README histogram with and without moments, OpenCensus distribution, the seven ssd_test lines (sample,
  bracket pass, resident select), a radix pass and a compaction (select forced onto the radix route), the text
  codec (2 M lines) and the float64 histogram.  Prints nothing that matters; the timings are ncu's."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
ctx = crc32c.Context([0])
L, h = ctx._L, ctx._h
total = 4 << 30
d = ctx.alloc_device(total)
ctx.fill_payload(d, total // 8, 0xC0FFEE)
rng = np.random.default_rng(0)
# small objects
n = 1 << 20
b = crc32c.DeviceBatch(np.arange(n, dtype=np.uint64) * 4096, np.full(n, 4096, dtype=np.uint64), None, ctx)
b.run(d); b.run(d); b.close()
# the same blocks 8 bytes off a 16-byte boundary: the row kernel's shape (the aligned ones above go to the fold kernel)
b = crc32c.DeviceBatch(np.arange(n - 1, dtype=np.uint64) * 4096 + 8, np.full(n - 1, 4096, dtype=np.uint64), None, ctx)
b.run(d); b.run(d); b.close()
# a million ragged objects under 8 KiB, unaligned: the fold kernel a lane per object + the thread-per-object finish
ln_t = rng.integers(0, 8192, size=n).astype(np.uint64)
off_t = np.arange(n, dtype=np.uint64) * np.uint64(4096 - 16) + rng.integers(0, 16, size=n).astype(np.uint64)
b = crc32c.DeviceBatch(off_t, np.minimum(ln_t, 4000), None, ctx)
b.run(d); b.run(d); b.close()
# ragged
ln = rng.integers(1, 4 << 20, size=2048).astype(np.uint64)
off = np.concatenate([[0], np.cumsum(ln[:-1] + rng.integers(0, 64, size=2047).astype(np.uint64))]).astype(np.uint64)
keep = off + ln <= np.uint64(total)
b = crc32c.DeviceBatch(off[keep], ln[keep], None, ctx)
b.run(d); b.run(d); b.close()
ctx.free_device(d)
# latency
ns = 200_000_000
ds = ctx.alloc_device(ns * 8)
ctx.fill_latency_ns(ds, ns, 0xC0FFEE)
edges = lat.README_EDGES
counts = np.zeros(40, dtype=np.uint64)
st = pkg._lib.LatStats()
for _ in range(2):
    L.b200_hist_fixed(h, 0, ds, ns, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, C.byref(st))
    L.b200_hist_fixed(h, 0, ds, ns, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, None)
    lat.hist_opencensus(ds, ns, ctx=ctx)
    lat.percentiles(ds, ns, lat.NS_PER_US, ctx=ctx)
mn, mx = C.c_int64(), C.c_int64()
L.b200_minmax(h, 0, ds, ns, C.byref(mn), C.byref(mx))
kmin = (mn.value + (1 << 63)) & ((1 << 64) - 1)
bits = (mx.value - mn.value).bit_length()
cnt = np.zeros(4 * 4096, dtype=np.uint64)
pf = np.zeros(4, dtype=np.uint64)
L.b200_radix_count(h, 0, ds, ns, kmin, bits - 12, bits, 1, pf.ctypes.data, cnt.ctypes.data)
pf[:] = [19, 23, 29, 38]
L.b200_radix_count(h, 0, ds, ns, kmin, bits - 24, bits - 12, 4, pf.ctypes.data, cnt.ctypes.data)
ctx.free_device(ds)
# text
nl = 2_000_000
vals = rng.integers(20_000_000, 95_000_000, size=100_000)
chunk = ("\n".join(lat.go_duration_line(int(v)) for v in vals) + "\n").encode()
data = np.frombuffer(chunk * (nl // 100_000), dtype=np.uint8)
dt = ctx.alloc_device(data.size)
ctx.h2d(dt, data)
for _ in range(2):
    p = lat.parse_latency_text(dt, data.size, nl, ctx=ctx)
    hf = lat.hist_fixed_f64(p.d_values, p.nlines, edges, ctx=ctx)
    ctx.free_device(p.d_values); ctx.free_device(p.d_status)
print("ok", int(hf.counts.sum()))
