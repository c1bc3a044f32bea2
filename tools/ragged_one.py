"""One ragged VerifyObjects batch, for ncu launch lists: python tools/ragged_one.py"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
c = pkg.crc32c
ctx = c.Context([0])
total = 8 << 30
d = ctx.alloc_device(total)
ctx.fill_payload(d, total // 8, 0xC0FFEE)
rng = np.random.default_rng(0)
ln = rng.integers(1, 4 << 20, size=4096).astype(np.uint64)
off = np.concatenate([[0], np.cumsum(ln[:-1] + rng.integers(0, 64, size=4095).astype(np.uint64))]).astype(np.uint64)
keep = off + ln <= np.uint64(total)
off, ln = off[keep], ln[keep]
b = c.DeviceBatch(off, ln, None, ctx)
for _ in range(4):
    b.run(d)
print("ms", b.kernel_ms(), "launches", b.launches())
