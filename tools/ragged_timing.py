"""Times VerifyObjects on ragged device-resident batches and the text codec on one GPU."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
c, lat = pkg.crc32c, pkg.latency
ctx = c.Context([0])
total = 8 << 30
d = ctx.alloc_device(total + (8 << 20))
ctx.fill_payload(d, (total + (8 << 20)) // 8, 0xC0FFEE)
rng = np.random.default_rng(0)


def run(name, off, ln):
    b = c.DeviceBatch(off, ln, None, ctx)
    for _ in range(3): b.run(d)
    ctx.timer_start()
    K = 50
    for _ in range(K): b.run_async(d)
    ms = ctx.timer_stop() / K
    nbytes = int(ln.sum())
    print(f"{name:46s} {len(ln):8d} objs {nbytes/2**30:6.2f} GiB  {ms:8.3f} ms  {nbytes/ms/1e6:8.1f} GB/s  launches={b.launches()}")
    b.close()


ln = rng.integers(1, 4 << 20, size=4096).astype(np.uint64)
off = np.concatenate([[0], np.cumsum(ln[:-1] + rng.integers(0, 64, size=4095).astype(np.uint64))]).astype(np.uint64)
keep = off + ln <= np.uint64(total)   # stay inside the arena
off, ln = off[keep], ln[keep]
run("ragged: 4096 objs, len U[1, 4 MiB], unaligned", off, ln)
for size in (4096, 65536, 300_000, 1 << 20):
    n = min(total // size, 1 << 20)
    off = (np.arange(n, dtype=np.uint64) * np.uint64(size))
    run(f"uniform {size} B", off, np.full(n, size, dtype=np.uint64))
n = 1 << 20
ln = rng.integers(0, 8192, size=n).astype(np.uint64)
off = (np.arange(n, dtype=np.uint64) * np.uint64(8192) + rng.integers(0, 16, size=n).astype(np.uint64))
run("tiny ragged: 1M objs, len U[0, 8 KiB)", off, ln)

# text codec
nl = 20_000_000
ns = np.random.default_rng(1).integers(20_000_000, 95_000_000, size=200_000)
chunk = ("\n".join(lat.go_duration_line(int(v)) for v in ns) + "\n").encode()
reps = nl // 200_000
data = np.frombuffer(chunk * reps, dtype=np.uint8)
dt = ctx.alloc_device(data.size)
ctx.h2d(dt, data)
import ctypes as C
dv = ctx.alloc_device(nl * 8); dst = ctx.alloc_device(nl)
out = [C.c_uint64() for _ in range(4)]
call = lambda: pkg._lib.check(ctx._L.b200_parse_latency_text(ctx._h, 0, dt, data.size, dv, dst, nl, *[C.byref(o) for o in out]))
call()
t0 = time.perf_counter()
for _ in range(5):
    call()
ms = (time.perf_counter() - t0) / 5 * 1e3
print(f"text parse: {out[0].value} lines, {data.size/1e6:.1f} MB in {ms:.2f} ms -> {data.size/ms/1e6:.1f} GB/s, {out[0].value/ms/1e3:.1f} Mlines/s (bad={out[2].value})")
