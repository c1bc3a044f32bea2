import importlib, time, numpy as np, sys, os
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module('custom-go-client-benchmark_b200')
c = pkg.crc32c
ctx = c.Context([0])
nobj, size = 4096, 2 << 20
d = ctx.alloc_device(nobj * size)
ctx.fill_payload(d, nobj * size // 8, 0xC0FFEE)
off = np.arange(nobj, dtype=np.uint64) * size
ln = np.full(nobj, size, dtype=np.uint64)
for rps in (None,):
    b = c.DeviceBatch(off, ln, None, ctx)
    for i in range(3): b.run(d)
    ts = []
    for i in range(10):
        b.run(d); ts.append(b.kernel_ms())
    tot = np.median([t[0] for t in ts]); rows = np.median([t[1] for t in ts])
    print(f"batch 4096x2MiB: total {tot:.3f} ms rows {rows:.3f} ms -> {nobj*size/rows/1e6:.1f} GB/s (rows), {nobj*size/tot/1e6:.1f} GB/s (total)")
for i in range(5): b.run(d)
ctx.timer_start()
K = 300
for i in range(K): b.run_async(d)
ms = ctx.timer_stop() / K
print(f"sustained {K} back-to-back steps: {ms:.4f} ms/step -> {nobj*size/ms/1e6:.1f} GB/s  variant={os.environ.get('B200_FOLD_VARIANT','0')} nofold={os.environ.get('B200_NO_FOLD','')}")
t0 = time.time(); w = c.UpdateDevice(0, d, nobj * size, ctx); t1 = time.time()
print("single 8GiB object", hex(w), f"{(t1-t0)*1e3:.2f} ms wall")
os.system("nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv")
