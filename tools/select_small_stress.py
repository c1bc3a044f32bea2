"""The resident (one-kernel) select against numpy.partition on inputs of 1 .. 8 M samples: random sizes, ranks,
alignments and distributions.  python tools/select_small_stress.py [iterations]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 100
ctx = crc32c.Context([0])
rng = np.random.default_rng(7)
cap = 8_400_000
d = ctx.alloc_device(cap * 8 + 16)
bad = 0
for it in range(iters):
    n = int(rng.choice([1, 2, 3, 100, 4097, 65536, 1_000_000, 4_166_666, 8_000_001])) if it % 3 else int(rng.integers(1, cap))
    kind = it % 4
    if kind == 0:
        ns = rng.integers(20_000_000, 95_000_000, size=n)
    elif kind == 1:
        ns = rng.integers(-2**62, 2**62, size=n)
    elif kind == 2:
        ns = np.full(n, 57_000_000) + rng.integers(0, 3, size=n)
    else:
        ns = (rng.exponential(5e7, size=n)).astype(np.int64)
        ns[rng.integers(0, n, size=max(1, n // 200))] = 9_000_000_000
    ns = ns.astype(np.int64)
    off = int(rng.integers(0, 2))
    ctx.h2d(d + 8 * off, ns)
    nk = int(rng.integers(1, 9))
    ks = sorted(int(x) for x in rng.integers(0, n, size=nk))
    want = [int(x) for x in np.partition(ns, ks)[ks]]
    try:
        got = lat.select_kth(d + 8 * off, n, 1, ks, ctx=ctx)
    except Exception as e:
        got = repr(e)[:120]
    if got != want:
        bad += 1
        print(f"MISMATCH it={it} n={n} kind={kind} off={off} ks={ks} got={got} want={want}", flush=True)
print("mismatches:", bad, "of", iters)
