"""Times the latency text codec on one GPU: python tools/text_timing.py [nlines] [reps]
20 M lines of `tr 'ms' ' '`-ed Go Duration strings (execute_pb.sh:4,8), ~238 MB."""
import ctypes as C, importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
nl = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
ctx = crc32c.Context([0])
ns = np.random.default_rng(1).integers(20_000_000, 95_000_000, size=200_000)
chunk = ("\n".join(lat.go_duration_line(int(v)) for v in ns) + "\n").encode()
data = np.frombuffer(chunk * (nl // 200_000), dtype=np.uint8)
dt = ctx.alloc_device(data.size)
ctx.h2d(dt, data)
dv = ctx.alloc_device(nl * 8); dst = ctx.alloc_device(nl)
out = [C.c_uint64() for _ in range(4)]
call = lambda: pkg._lib.check(ctx._L.b200_parse_latency_text(ctx._h, 0, dt, data.size, dv, dst, nl, *[C.byref(o) for o in out]))
call()
t0 = time.perf_counter()
for _ in range(reps):
    call()
ms = (time.perf_counter() - t0) / reps * 1e3
vals = np.zeros(1000, dtype=np.float64)
ctx.d2h(vals, dv)
want = np.array([float(l) for l in chunk.decode().split("\n")[:1000]])
assert (vals == want).all(), "parsed values differ from Python float()"
print(f"text parse: {out[0].value} lines, {data.size/1e6:.1f} MB in {ms:.3f} ms -> {data.size/ms/1e6:.1f} GB/s of text, "
      f"{out[0].value/ms/1e3:.1f} Mlines/s (bad={out[2].value})")
