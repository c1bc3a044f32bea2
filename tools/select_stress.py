"""Randomised check of b200_select_kth / b200_select_kth_stats on inputs long enough for the sampled route
(>= 2^24 samples) against numpy.partition: python tools/select_stress.py [cases] [seed]"""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
ctx = pkg.crc32c.Context([0]); lat = pkg.latency
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 12
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
nmax = 1 << 25
d = ctx.alloc_device(nmax * 8 + 8)
kinds = ["lognormal", "uniform64", "exp", "bimodal", "mostly_constant", "sorted", "reversed", "small_range", "negative"]
bad = 0
for c in range(cases):
    n = int(rng.integers(1 << 24, nmax))
    kind = kinds[c % len(kinds)]
    if kind == "lognormal": x = (rng.lognormal(17, 0.6, n)).astype(np.int64)
    elif kind == "uniform64": x = rng.integers(-(1 << 62), 1 << 62, n)
    elif kind == "exp": x = rng.exponential(5e7, n).astype(np.int64)
    elif kind == "bimodal": x = np.where(rng.random(n) < 0.3, rng.normal(2e7, 1e6, n), rng.normal(9e9, 3e9, n)).astype(np.int64)
    elif kind == "mostly_constant": x = np.where(rng.random(n) < 0.9, 55_000_000, rng.integers(0, 10**8, n)).astype(np.int64)
    elif kind == "sorted": x = np.sort(rng.integers(0, 10**9, n))
    elif kind == "reversed": x = np.sort(rng.integers(0, 10**9, n))[::-1].copy()
    elif kind == "small_range": x = rng.integers(1000, 1064, n)
    else: x = -rng.exponential(3e6, n).astype(np.int64)
    x = np.ascontiguousarray(x, dtype=np.int64)
    start = int(rng.integers(0, 2))
    ctx.h2d(d + 8 * start, x)
    div = int(rng.choice([1, 1000, 1_000_000, 7]))
    nk = int(rng.integers(1, 5))
    ks = sorted({int(k) for k in rng.integers(0, n, nk)} | ({0} if c % 3 == 0 else set()) | ({n - 1} if c % 4 == 0 else set()))[:4]
    t0 = time.perf_counter()
    got = lat.select_kth(d + 8 * start, n, div, ks, ctx=ctx)
    ms = (time.perf_counter() - t0) * 1e3
    part = np.partition(x, ks)
    want = [int(abs(int(part[k])) // div * (1 if part[k] >= 0 else -1)) for k in ks]
    p = lat.percentiles(d + 8 * start, n, div, ctx=ctx)
    k4 = [n // 5, n // 2, 9 * n // 10, 99 * n // 100]
    part = np.partition(x, k4)
    tr = lambda v: int(abs(int(v)) // div * (1 if v >= 0 else -1))
    q = (np.abs(x) // div) * np.sign(x)
    s = int(q.sum())
    want_p = (abs(s) // n * (1 if s >= 0 else -1), *[tr(part[k]) for k in k4], int(q.min()), int(q.max()))
    ok = got == want and (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == want_p
    bad += not ok
    print(f"{c:2d} {kind:16s} n={n} div={div} ks={ks} {ms:7.3f} ms {'ok' if ok else 'MISMATCH ' + str((got, want, p, want_p))}", flush=True)
print("mismatches:", bad)
sys.exit(1 if bad else 0)
