"""GPU parity tests for the latency reduction path (histogram, distribution moments, exact percentiles)."""
import importlib
import json
import os

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency


@pytest.fixture(scope="module")
def ctx():
    c = crc32c.Context([0])
    yield c
    c.close()


def upload(ctx, a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    d = ctx.alloc_device(max(a.nbytes, 8))
    if a.size:
        ctx.h2d(d, a)
    return d


@pytest.fixture(scope="module")
def samples(ctx):
    n = 1_000_003
    d = ctx.alloc_device(n * 8)
    ctx.fill_latency_ns(d, n, oracle.SEED)
    yield oracle.latency_ns(n), d, n
    ctx.free_device(d)


def test_device_generator_matches_oracle(ctx, samples):
    ns, d, n = samples
    back = np.empty(n, dtype=np.int64)
    ctx.d2h(back, d)
    assert (back == ns).all()


def test_readme_histogram(ctx, samples):
    ns, d, n = samples
    ms = oracle.ns_to_ms(ns)
    ref, dropped = oracle.hist_numpy(ms.astype(np.float64))
    h = lat.hist_readme(d, n, ctx=ctx)
    assert list(h.counts) == list(ref)
    assert h.dropped_below + h.dropped_above == dropped
    assert h.dropped_below == int((ms < 20).sum()) and h.dropped_above == int((ms > 95).sum())


def test_opencensus_distribution(ctx, samples):
    ns, d, n = samples
    ms = oracle.ns_to_ms(ns)
    ref = oracle.hist_opencensus(ms.astype(np.float64))
    h = lat.hist_opencensus(d, n, ctx=ctx)
    assert list(h.counts) == list(ref)
    s = oracle.stats_i64(ms)
    assert (h.count, h.min, h.max) == (s.count, s.min, s.max)
    assert h.sum == int(ms.sum()) and h.sum_wrapped == int(ms.sum())
    assert h.sumsq == int((ms.astype(object) ** 2).sum())
    # float views.  Ours is the exactly rounded sum/n; the reference's means are order dependent float
    # recurrences (OpenCensus Welford, README left-to-right sum) that drift by O(n) roundings, so the
    # tolerance is 1e-12 relative (observed ~1.6e-15 * 100 at n = 1e6).
    assert h.mean == float(int(ms.sum()) / ms.size) or abs(h.mean - int(ms.sum()) / ms.size) <= np.spacing(h.mean)
    assert abs(h.mean - s.mean_welford) <= 1e-12 * abs(h.mean)
    assert abs(h.mean - s.mean_seq) <= 1e-12 * abs(h.mean)
    assert abs(h.sum_of_squared_dev - s.sum_sq_dev) <= 1e-9 * s.sum_sq_dev


def test_two_histograms_from_one_pass(ctx, samples):
    ns, d, n = samples
    ms = oracle.ns_to_ms(ns).astype(np.float64)
    readme, oc = lat.hist_readme_and_opencensus(d, n, ctx=ctx)
    ref, dropped = oracle.hist_numpy(ms)
    assert list(readme.counts) == list(ref) and readme.dropped_below + readme.dropped_above == dropped
    assert list(oc.counts) == list(oracle.hist_opencensus(ms)) and oc.dropped_below == oc.dropped_above == 0
    one = lat.hist_opencensus(d, n, ctx=ctx)
    assert (oc.count, oc.min, oc.max, oc.sum, oc.sumsq) == (one.count, one.min, one.max, one.sum, one.sumsq)
    assert (readme.sum, readme.sumsq) == (one.sum, one.sumsq)
    # any two edge sets: fractional edges that collapse onto one threshold, shared edges, both modes, negatives
    rng = np.random.default_rng(9)
    v = rng.integers(-40_000_000, 150_000_000, size=70_001).astype(np.int64)
    dv = upload(ctx, v)
    for div in (1_000_000, 1000, 7):
        q = ((np.abs(v) // div) * np.sign(v)).astype(np.float64)
        ea = np.array([-3.5, -3.25, 0, 0.25, 0.5, 19.9, 20, 95, 1e4], dtype=np.float64)
        eb = np.array([-10, 0, 20, 20.5, 77, 1e4, 1e5], dtype=np.float64)
        for ma, mb in ((pkg._lib.HIST_NUMPY, pkg._lib.HIST_OPENCENSUS), (pkg._lib.HIST_OPENCENSUS, pkg._lib.HIST_NUMPY),
                       (pkg._lib.HIST_NUMPY, pkg._lib.HIST_NUMPY)):
            a, b = lat.hist_fixed_pair(dv, v.size, div, ea, ma, eb, mb, ctx=ctx)
            for h, e, m in ((a, ea, ma), (b, eb, mb)):
                want = lat.hist_fixed(dv, v.size, div, e, m, ctx=ctx)
                assert list(h.counts) == list(want.counts), (div, m)
                assert (h.dropped_below, h.dropped_above, h.sum, h.min, h.max) == (
                    want.dropped_below, want.dropped_above, want.sum, want.min, want.max)
                if m == pkg._lib.HIST_NUMPY:
                    ref, dropped = oracle.hist_numpy(q, e)
                    assert list(h.counts) == list(ref) and h.dropped_below + h.dropped_above == dropped
                else:
                    assert list(h.counts) == list(oracle.hist_opencensus(q, e))
    # too many distinct thresholds between the two sets
    big_a, big_b = np.arange(0, 64, dtype=np.float64), np.arange(0.0, 64.0) + 100
    with pytest.raises(pkg.B200Error) as e:
        lat.hist_fixed_pair(dv, v.size, 1, big_a, pkg._lib.HIST_OPENCENSUS, big_b, pkg._lib.HIST_OPENCENSUS, ctx=ctx)
    assert e.value.status == pkg._lib.ERR_TOO_LARGE
    ctx.free_device(dv)


def test_histogram_edge_values(ctx):
    ms_vals = [19, 20, 24, 25, 90, 94, 95, 96, 100, 0, -5]
    ns = np.array([v * 1_000_000 for v in ms_vals] + [19_999_999, 20_000_000, 24_999_999, 25_000_000,
                                                      95_999_999, 96_000_000, -1, -999_999, -1_000_000], dtype=np.int64)
    d = upload(ctx, ns)
    ms = oracle.ns_to_ms(ns)
    ref, dropped = oracle.hist_numpy(ms.astype(np.float64))
    h = lat.hist_readme(d, ns.size, ctx=ctx)
    assert list(h.counts) == list(ref) and h.dropped_below + h.dropped_above == dropped
    oc = lat.hist_opencensus(d, ns.size, ctx=ctx)
    assert list(oc.counts) == list(oracle.hist_opencensus(ms.astype(np.float64)))
    s = oracle.stats_i64(ms)
    assert (oc.count, oc.min, oc.max, oc.sum) == (s.count, s.min, s.max, int(ms.sum()))
    ctx.free_device(d)


def test_fractional_and_negative_edges(ctx):
    rng = np.random.default_rng(5)
    ns = rng.integers(-50_000_000, 200_000_000, size=50_001).astype(np.int64)
    d = upload(ctx, ns)
    for div in (1_000_000, 1000, 1, 7):
        v = (np.abs(ns) // div) * np.sign(ns)  # Go truncation
        edges = np.array([-30.5, -10, -0.5, 0, 0.25, 3, 19.999, 20, 20.5, 77, 1e5], dtype=np.float64)
        h = lat.hist_fixed(d, ns.size, div, edges, pkg._lib.HIST_NUMPY, ctx=ctx)
        ref, dropped = oracle.hist_numpy(v.astype(np.float64), edges)
        assert list(h.counts) == list(ref), div
        assert h.dropped_below + h.dropped_above == dropped
        oc = lat.hist_fixed(d, ns.size, div, edges, pkg._lib.HIST_OPENCENSUS, ctx=ctx)
        assert list(oc.counts) == list(oracle.hist_opencensus(v.astype(np.float64), edges))
        assert oc.sum == int(v.sum()) and oc.min == v.min() and oc.max == v.max()
    ctx.free_device(d)


def test_fast_path_cells_and_divisors(ctx):
    """The 32-bit fast path: every log-linear cell boundary below 2^32 (and the pair cells below 64), thresholds
    that split a cell, cells that hold two thresholds (generic path), and each way the kernel divides."""
    rng = np.random.default_rng(11)
    vals = list(range(0, 200))
    for e in range(6, 33):
        for mant in (0, 1, 13, 31):
            b = (1 << e) + (mant << (e - 5))
            vals += [b - 1, b, b + 1]
    vals += [(1 << 32) - 1, 1 << 32, (1 << 32) + 1, 1 << 40, -1, -(1 << 33)]
    ns = np.concatenate([np.array(vals, dtype=np.int64),
                         rng.integers(0, 1 << 32, size=200_000, dtype=np.int64),
                         rng.integers(0, 4096, size=100_000, dtype=np.int64),
                         rng.integers(-(1 << 34), 1 << 36, size=5_000, dtype=np.int64)])
    rng.shuffle(ns)
    d = upload(ctx, ns)
    absn = np.abs(ns).astype(object)
    for div in (1, 2, 16, 17, 24, 1000, 1_000_000, 999_983, (1 << 31) + 1, (1 << 32) + 5):
        v = np.array([int(a) // div * (1 if x >= 0 else -1) for a, x in zip(absn, ns)], dtype=object)  # Go truncation
        vf = v.astype(np.float64)
        top = float((1 << 36) // div + 2)
        for edges in (np.array([0, 1, 2, 3, 4, 5, 6.5, 7, 33, 34, 63, 64, 65, 100, 1000, 4000, top]) ,
                      np.array([0.5, 2, 1e3, 1e6, 2.0 ** 31, 2.0 ** 32 - 1, 2.0 ** 32, 2.0 ** 33, 2.0 ** 37])):
            edges = np.unique(edges.astype(np.float64))
            oc = lat.hist_fixed(d, ns.size, div, edges, pkg._lib.HIST_OPENCENSUS, ctx=ctx)
            assert list(oc.counts) == list(oracle.hist_opencensus(vf, edges)), (div, edges)
            assert oc.sum == int(v.sum()) and oc.sumsq == int((v * v).sum()), div
            assert (oc.min, oc.max) == (int(v.min()), int(v.max())), div
            h = lat.hist_fixed(d, ns.size, div, edges, pkg._lib.HIST_NUMPY, ctx=ctx)
            ref, dropped = oracle.hist_numpy(vf, edges)
            assert list(h.counts) == list(ref) and h.dropped_below + h.dropped_above == dropped, (div, edges)
    ctx.free_device(d)


def test_histogram_with_the_maximum_number_of_edges(ctx, samples):
    ns, d, n = samples
    ms = oracle.ns_to_ms(ns)
    edges = np.linspace(1.0, 190.0, pkg._lib.HIST_MAX_EDGES if hasattr(pkg._lib, "HIST_MAX_EDGES") else 64)
    for mode, ref in ((pkg._lib.HIST_OPENCENSUS, oracle.hist_opencensus(ms.astype(np.float64), edges)),
                      (pkg._lib.HIST_NUMPY, oracle.hist_numpy(ms.astype(np.float64), edges)[0])):
        h = lat.hist_fixed(d, n, lat.NS_PER_MS, edges, mode, ctx=ctx)
        assert list(h.counts) == list(ref)
        assert h.sum == int(ms.sum())


def test_histogram_unaligned_and_tiny(ctx, samples):
    ns, d, n = samples
    for start, cnt in [(1, 1001), (3, 2), (5, 1), (0, 0), (7, 100000)]:
        ms = oracle.ns_to_ms(ns[start:start + cnt])
        ref, dropped = oracle.hist_numpy(ms.astype(np.float64))
        h = lat.hist_readme(d + 8 * start, cnt, ctx=ctx)
        assert list(h.counts) == list(ref) and h.count == cnt


def test_percentiles_match_reference_rule(ctx, samples):
    ns, d, n = samples
    us = oracle.ns_to_us(ns)
    o = oracle.percentiles_i64(us)
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (o.avg, o.p20, o.p50, o.p90, o.p99, o.min, o.max)


@pytest.mark.parametrize("n", [1, 2, 5, 10, 99, 100, 101, 4097])
def test_percentile_index_rule_small_n(ctx, n):
    rng = np.random.default_rng(n)
    ns = rng.integers(0, 10**10, size=n).astype(np.int64)
    d = upload(ctx, ns)
    o = oracle.percentiles_i64(oracle.ns_to_us(ns))
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (o.avg, o.p20, o.p50, o.p90, o.p99, o.min, o.max)
    ctx.free_device(d)


def test_select_every_rank_with_duplicates_and_negatives(ctx):
    rng = np.random.default_rng(2)
    ns = np.concatenate([rng.integers(-1000, 1000, size=300), np.full(200, 42), np.full(50, -7),
                         rng.integers(-2**62, 2**62, size=20)]).astype(np.int64)
    rng.shuffle(ns)
    d = upload(ctx, ns)
    ks = list(range(ns.size))
    got = lat.select_kth(d, ns.size, 1, ks, ctx=ctx)
    assert got == [int(x) for x in np.sort(ns)]
    got3 = lat.select_kth(d, ns.size, 3, [0, 100, 569], ctx=ctx)
    v = np.sort((np.abs(ns.astype(object)) // 3) * np.sign(ns))
    assert got3 == [int(v[0]), int(v[100]), int(v[569])]
    ctx.free_device(d)


def test_select_with_bucket_compaction_unaligned(ctx):
    """n >= 2^20 makes the select copy the surviving buckets after its first pass; unaligned start, odd count."""
    n = 3_000_002
    d = ctx.alloc_device(n * 8)
    ctx.fill_latency_ns(d, n, oracle.SEED, first=77)
    ns = oracle.latency_ns(n, first=77)
    for start, cnt in [(1, n - 1), (0, n), (3, 2_500_000)]:
        srt = np.sort(oracle.ns_to_us(ns[start:start + cnt]))
        ks = [0, 1, cnt // 5, cnt // 2, 9 * cnt // 10, 99 * cnt // 100, cnt - 2, cnt - 1]
        got = lat.select_kth(d + 8 * start, cnt, 1000, ks, ctx=ctx)
        assert got == [int(srt[k]) for k in ks], (start, cnt)
    # spread-out values: each wanted rank's bucket holds ~1/4096 of the samples, so the compaction path runs
    rng = np.random.default_rng(9)
    wide = rng.integers(-2**39, 2**39, size=(1 << 21) + 3).astype(np.int64)
    dw = upload(ctx, wide)
    for start, cnt in [(0, wide.size), (1, wide.size - 2)]:
        srt = np.sort(wide[start:start + cnt])
        ks = [0, 7, cnt // 5, cnt // 2, cnt // 2 + 1, 9 * cnt // 10, 99 * cnt // 100, cnt - 1]
        assert lat.select_kth(dw + 8 * start, cnt, 1, ks, ctx=ctx) == [int(srt[k]) for k in ks]
    ctx.free_device(dw)
    # a spike: most samples in one bucket (compaction is skipped, the full-data path must agree)
    spike = np.full(2_000_000, 57_000_123, dtype=np.int64)
    spike[::1000] = np.arange(2000) * 5_000_000
    ds = upload(ctx, spike)
    srt = np.sort(spike)
    assert lat.select_kth(ds, spike.size, 1, [0, 999, 1_000_000, spike.size - 1], ctx=ctx) == [
        int(srt[0]), int(srt[999]), int(srt[1_000_000]), int(srt[-1])]
    ctx.free_device(ds)
    ctx.free_device(d)


SAMPLED_SELECT_CODE = r"""
import importlib, sys
import numpy as np
import oracle
pkg = importlib.import_module('custom-go-client-benchmark_b200')
crc32c, lat = pkg.crc32c, pkg.latency
ctx = crc32c.Context([0])

def check(ns, ranks, div, start=0):
    ns = np.ascontiguousarray(ns, dtype=np.int64)
    d = ctx.alloc_device(ns.nbytes)
    ctx.h2d(d, ns)
    part = np.sort(ns[start:])
    got = lat.select_kth(d + 8 * start, part.size, div, ranks, ctx=ctx)
    want = [int(abs(int(part[k])) // div * (1 if part[k] >= 0 else -1)) for k in ranks]
    assert list(got) == want, (list(got), want)
    ctx.free_device(d)
    print('case done', flush=True)
    sys.stderr.write('--\n'); sys.stderr.flush()

n = 3_000_001
base = oracle.latency_ns(n)
q = lambda m, ps: [int(m * p) for p in ps]
check(base, q(n, [0.2, 0.5, 0.9, 0.99]), 1_000_000)            # the README / ssd_test ranks
check(base, q(n - 1, [0.2, 0.5, 0.9, 0.99]), 1, start=1)        # 16-byte misaligned start
check(base, [0, n // 2, n - 1], 1000)                           # brackets open at both ends
check(base, [n // 2, n // 2 + 1, n // 2 + 2, n // 2 + 700], 1)  # overlapping brackets merge
rng = np.random.default_rng(3)
wide = rng.integers(-(1 << 40), 1 << 40, size=n)                # negative, and beyond the 32-bit fast path
check(wide, q(n, [0.001, 0.25, 0.75, 0.999]), 7)
check(np.sort(base), q(n, [0.1, 0.5, 0.95]), 1)                 # sorted input
period = (np.arange(n) % 48) * 1_000_000 + rng.integers(0, 1000, size=n)   # 48 workers, round robin
check(period, q(n, [0.2, 0.5, 0.9, 0.99]), 1)
coarse = (base // 1_000_000) * 1_000_000                        # ~100 distinct values: brackets get too heavy
check(coarse, q(n, [0.2, 0.5, 0.9, 0.99]), 1_000_000)
check(np.full(n, 12345, dtype=np.int64), [0, n // 2, n - 1], 1)  # all equal

def seven(ns):  # what ssd_test prints: ranks and moments from ONE call (b200_select_kth_stats)
    d = ctx.alloc_device(ns.nbytes)
    ctx.h2d(d, ns)
    o = oracle.percentiles_i64(oracle.ns_to_us(ns))
    p = lat.percentiles(d, ns.size, lat.NS_PER_US, ctx=ctx)
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (o.avg, o.p20, o.p50, o.p90, o.p99, o.min, o.max)
    ctx.free_device(d)
    sys.stderr.write('--\n'); sys.stderr.flush()

seven(base)                                                     # moments from the bracket pass
seven(np.ascontiguousarray(wide))                               # negative samples: truncation toward zero
seven(np.ascontiguousarray(coarse))                             # a miss: moments from an extra pass
stalls = np.where(rng.random(n) < 0.03, rng.integers(5 * 10**9, 10**11, size=n), base)   # p99 among 5-100 s reads:
check(stalls, q(n, [0.5, 0.9, 0.99, 0.999]), 1000)              # brackets beyond the 32-bit fast path
check(base, q(n, [0.01, 0.05, 0.1, 0.25, 0.5, 0.75, 0.9, 0.99, 0.999]), 1000)   # nine ranks: a pass of 8 and a pass of 1
print('ok')
"""


def test_sampled_select_matches_sort():
    """The one-pass sampled select (taken for >= 2^24 samples by default) forced onto 3 * 10^6 samples: exact on every
    kind of input, whether its brackets catch the ranks (trace says ok) or it hands over to the radix passes."""
    import subprocess, sys
    env = dict(os.environ, B200_SELECT_SAMPLED_MIN="100000", B200_SELECT_TRACE="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", SAMPLED_SELECT_CODE], env=env, cwd=root, text=True,
                       capture_output=True, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr
    cases = r.stderr.split("--\n")[:-1]
    assert len(cases) == 14, r.stderr
    assert cases[13].count("sampled ok") == 2, cases[13]
    for i in (0, 1, 2, 3, 5, 6, 9, 12, 13):  # the brackets must catch the ranks on all the spread-out inputs
        assert "sampled ok" in cases[i], (i, cases[i])
    for i in (7, 8, 11):             # heavy duplicates: a miss, then the radix passes
        assert "sampled miss" in cases[i], (i, cases[i])
    for i in (4, 10):                # mostly negative / beyond 2^32: not latency-shaped, left to the radix passes
        assert "sampled skip" in cases[i], (i, cases[i])


def test_host_driven_select_still_matches():
    """Inputs up to 2^25 samples take the device-resident select and longer ones the sampled select, so the
    host-driven passes with bucket compaction only run after a sampled miss on a very long input.  Force them
    (and switch the other two off) and run the select tests of this file again."""
    import subprocess, sys
    env = dict(os.environ, B200_SELECT_RESIDENT_MAX="0", B200_SELECT_SAMPLED_MIN=str(1 << 62))
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "-k",
                        "(select or percentile or golden or multi_device) and not host_driven and not sampled",
                        os.path.join(root, "tests", "test_latency_gpu.py")],
                       env=env, cwd=root, text=True, capture_output=True, timeout=900)
    assert r.returncode == 0 and " passed" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("nk", [1, 2, 3, 5, 8, 9, 17])
def test_select_any_number_of_ranks(ctx, samples, nk):
    """The device-resident select resolves up to 8 ranks together (one histogram while they share a prefix, one
    each afterwards) and groups beyond that; a single rank keeps a single histogram throughout."""
    ns, d, n = samples
    m = 300_000
    srt = np.sort(oracle.ns_to_us(ns[:m]))
    rng = np.random.default_rng(nk)
    for _ in range(3):
        ks = [int(x) for x in rng.integers(0, m, size=nk)]
        assert lat.select_kth(d, m, 1000, ks, ctx=ctx) == [int(srt[k]) for k in ks]


def test_multi_rank_sampled_select_building_blocks(ctx):
    """b200_select_sample + b200_bracket_pass under the multi-rank protocol (latency.sampled_select), here with one
    rank: exact ranks and exact moments from one pass; tests/test_dist_cpu.py runs the protocol across two ranks."""
    n = 5_000_001
    ns = oracle.latency_ns(n)
    d = upload(ctx, ns)
    red = lat._Reducer()
    for start, div in ((0, 1000), (1, 1)):
        part = ns[start:]
        m = part.size
        ks = [m // 5, m // 2, 9 * m // 10, 99 * m // 100]
        r = lat._sampled_select_on_gpu(red, ctx._L, ctx._h, 0, d + 8 * start, m, m, div, ks, True)
        assert r is not None
        vals, st, _ = r
        v = part // div
        srt = np.sort(v)
        assert vals == [int(srt[k]) for k in ks]
        assert (st.count, st.min, st.max, st.sum) == (m, int(v.min()), int(v.max()), int(v.sum()))
        # by fractions, the total left to the first collective, the moments riding along (latency.percentiles' form)
        r = lat._sampled_select_on_gpu(red, ctx._L, ctx._h, 0, d + 8 * start, m, None, div, None, "avg",
                                       fractions=[(1, 5), (1, 2), (9, 10), (99, 100)],
                                       extra_fn=lambda s: [s.count, s.min, s.max, s.sum, s.sumsq_lo])
        assert r is not None and r[0] == vals
        assert r[2] == [[m, int(v.min()) & (2**64 - 1), int(v.max()), int(v.sum()) & (2**64 - 1), 0]]
    # the building blocks directly: strata and odd-bin bookkeeping
    import ctypes as C
    p, cnt = C.c_void_p(), C.c_size_t()
    assert ctx._L.b200_select_sample(ctx._h, 0, d, n, 1000, C.byref(p), C.byref(cnt)) == 0 and cnt.value == n // 1000
    sample = np.empty(cnt.value, dtype=np.int64)
    ctx.d2h(sample, p.value)
    strata = ns[:cnt.value * 1000].reshape(-1, 1000)
    assert all(sample[i] in strata[i] for i in range(0, cnt.value, 97))
    thr = np.array([30_000_000, 30_100_000, 60_000_000, 60_000_001], dtype=np.int64)
    counts = np.zeros(5, dtype=np.uint64)
    kp, nkept = C.c_void_p(), C.c_uint64()
    assert ctx._L.b200_bracket_pass(ctx._h, 0, d, n, 1, thr.ctypes.data, 4, 1 << 20, counts.ctypes.data, C.byref(kp),
                                    C.byref(nkept), None) == 0
    c = np.searchsorted(thr, ns, side="right")
    assert list(counts) == list(np.bincount(c, minlength=5))
    kept = np.empty(nkept.value, dtype=np.int64)
    ctx.d2h(kept, kp.value)
    assert nkept.value == int((c % 2 == 1).sum()) and (np.sort(kept) == np.sort(ns[c % 2 == 1])).all()
    ctx.free_device(d)


def test_select_all_equal_and_empty(ctx):
    ns = np.full(1000, 123456789, dtype=np.int64)
    d = upload(ctx, ns)
    assert lat.select_kth(d, 1000, 1000, [0, 500, 999], ctx=ctx) == [123456] * 3
    with pytest.raises(pkg.B200Error) as e:
        lat.select_kth(d, 0, 1, [0], ctx=ctx)
    assert e.value.status == pkg._lib.ERR_EMPTY
    with pytest.raises(pkg.B200Error):
        lat.select_kth(d, 1000, 1, [1000], ctx=ctx)
    with pytest.raises(ZeroDivisionError):
        lat.percentiles(d, 0, ctx=ctx)
    # the one-call form (ranks + moments) reports the same conditions
    import ctypes as C
    L, h = ctx._L, ctx._h
    st = pkg._lib.LatStats()
    out = np.zeros(1, dtype=np.int64)

    def stats_call(n, k, ptr=None, divisor=1000):
        karr = np.array([k], dtype=np.uint64)
        return L.b200_select_kth_stats(h, 0, d if ptr is None else ptr, n, divisor, karr.ctypes.data, 1,
                                       out.ctypes.data, C.byref(st))
    assert stats_call(1000, 999) == 0 and out[0] == 123456
    assert (st.count, st.min, st.max, st.sum) == (1000, 123456, 123456, 123456 * 1000)
    assert stats_call(0, 0) == pkg._lib.ERR_EMPTY
    assert stats_call(1000, 1000) == pkg._lib.ERR_INVALID      # k >= n
    assert stats_call(999, 0, ptr=d + 4) == pkg._lib.ERR_INVALID  # samples must be 8-byte aligned
    assert stats_call(1000, 0, divisor=0) == pkg._lib.ERR_INVALID
    ctx.free_device(d)


def test_multi_device_form_with_one_gpu_standing_in_for_three(ctx, samples):
    """b200_hist_fixed_multi / b200_select_kth_multi: shards on 'several devices' (device 0 three times)."""
    import ctypes as C
    ns, d, n = samples
    cuts = [0, 200_001, 200_001, n]           # one empty shard
    ptrs = (C.c_void_p * 3)(*[d + 8 * cuts[i] for i in range(3)])
    sizes = (C.c_size_t * 3)(*[cuts[i + 1] - cuts[i] for i in range(3)])
    devs = (C.c_int * 3)(0, 0, 0)
    L = ctx._L
    edges = lat.README_EDGES
    counts = np.zeros(15, dtype=np.uint64)
    st = pkg._lib.LatStats()
    pkg._lib.check(L.b200_hist_fixed_multi(ctx._h, devs, ptrs, sizes, 3, 1_000_000, edges.ctypes.data, edges.size, 0,
                                           counts.ctypes.data, C.byref(st)))
    ms = oracle.ns_to_ms(ns)
    ref, dropped = oracle.hist_numpy(ms.astype(np.float64))
    assert list(counts) == list(ref) and st.dropped_below + st.dropped_above == dropped
    assert (st.count, st.min, st.max, st.sum) == (n, ms.min(), ms.max(), int(ms.sum()))
    assert (st.sumsq_hi << 64) | st.sumsq_lo == int((ms.astype(object) ** 2).sum())
    ks = np.array([0, n // 5, n // 2, 9 * n // 10, 99 * n // 100, n - 1], dtype=np.uint64)
    out = np.zeros(ks.size, dtype=np.int64)
    pkg._lib.check(L.b200_select_kth_multi(ctx._h, devs, ptrs, sizes, 3, 1000, ks.ctypes.data, ks.size, out.ctypes.data))
    srt = np.sort(oracle.ns_to_us(ns))
    assert list(out) == [int(srt[int(k)]) for k in ks]


def test_golden_fixture(ctx):
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))["latency"]
    n = g["n"]
    d = ctx.alloc_device(n * 8)
    ctx.fill_latency_ns(d, n, g["seed"])
    h = lat.hist_readme(d, n, ctx=ctx)
    assert [int(x) for x in h.counts] == g["numpy15_counts"]
    assert h.dropped_below + h.dropped_above == g["numpy15_dropped"]
    assert [int(x) for x in lat.hist_opencensus(d, n, ctx=ctx).counts] == g["oc35_counts"]
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
    assert [p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max] == g["us_percentiles"]
    ctx.free_device(d)


def test_large_sample_set_against_the_oracle(ctx):
    """1.25e8 samples (one GPU's share of BASELINE config 5) on the library's DEFAULT route thresholds — far above
    the 2^24 samples from which b200_select_kth takes the one-pass sampled route — against numpy.partition and
    numpy's histogram rule on the oracle-generated samples."""
    n = 125_000_000
    ns = oracle.latency_ns(n)
    d = ctx.alloc_device(n * 8)
    ctx.fill_latency_ns(d, n, oracle.SEED)
    h = lat.hist_readme(d, n, ctx=ctx)
    ms = ns // 1_000_000
    cnt = np.bincount(ms)
    vals = np.nonzero(cnt)[0]
    ref, _ = np.histogram(vals.astype(np.float64), lat.README_EDGES, weights=cnt[vals].astype(np.float64))
    assert [int(x) for x in h.counts] == [int(x) for x in ref]
    assert (h.dropped_below, h.dropped_above) == (int(cnt[:20].sum()), int(cnt[96:].sum()))
    assert (h.count, h.min, h.max, h.sum) == (n, int(ms.min()), int(ms.max()), int(ms.sum()))
    oc = lat.hist_opencensus(d, n, ctx=ctx)
    assert int(oc.counts.sum()) == n
    ks = [n // 5, n // 2, 9 * n // 10, 99 * n // 100]
    us = ns // 1000
    del ns, ms
    us_sum, us_min, us_max = int(us.sum()), int(us.min()), int(us.max())
    us.partition(ks)
    want = [int(us[k]) for k in ks]
    assert lat.select_kth(d, n, lat.NS_PER_US, ks, ctx=ctx) == want
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (us_sum // n, *want, us_min, us_max)
    ctx.free_device(d)


def test_default_routes_with_p99_beyond_32_bits(ctx):
    """2^25 + 12345 samples on default thresholds, 3 % of them multi-second stalls (5 s .. 9 s): p99 lies beyond
    2^32 ns, where the histogram pass leaves its 32-bit fast path.  Checked against numpy.partition."""
    n = (1 << 25) + 12345
    ns = oracle.latency_ns(n, seed=4242)
    idx = np.arange(0, n, 33)
    ns[idx] = 5_000_000_000 + (idx.astype(np.int64) * 7919) % 4_000_000_000
    d = ctx.alloc_device(n * 8)
    ctx.h2d(d, ns)
    ks = [0, n // 5, n // 2, 9 * n // 10, 99 * n // 100, n - 1]
    srt = np.partition(ns, ks)
    assert int(srt[ks[4]]) > 1 << 32
    assert lat.select_kth(d, n, 1, ks, ctx=ctx) == [int(srt[k]) for k in ks]
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
    us = ns // 1000
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (
        int(us.sum()) // n, *[int(srt[k]) // 1000 for k in ks[1:5]], int(us.min()), int(us.max()))
    ctx.free_device(d)


@pytest.mark.parametrize("burst,heavy", [(6000, False), (0, True)])
def test_bracket_pass_when_a_tile_parks_more_than_its_queue_holds(ctx, burst, heavy):
    """The collecting pass parks kept samples in a per-warp queue of 128: a run of consecutive samples equal to the
    median fills a tile's queue several times over (the tile is then redone one sample at a time), and an input that
    is mostly one value does so everywhere (the brackets come out too heavy and the radix passes take over, but the
    pass must still have counted every sample).  Checked against numpy.partition."""
    n = (1 << 24) + 4321
    ns = oracle.latency_ns(n, seed=99)
    med = int(np.partition(ns, n // 2)[n // 2])
    if burst:
        for at in (12345, n // 3, n - burst - 7):
            ns[at:at + burst] = med
    if heavy:
        ns[np.arange(n) % 5 != 0] = med
    d = ctx.alloc_device(n * 8)
    ctx.h2d(d, ns)
    ks = [n // 5, n // 2, 9 * n // 10, 99 * n // 100]
    srt = np.partition(ns, ks)
    assert lat.select_kth(d, n, 1, ks, ctx=ctx) == [int(srt[k]) for k in ks]
    p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
    us = ns // 1000
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (
        int(us.sum()) // n, *[int(srt[k]) // 1000 for k in ks], int(us.min()), int(us.max()))
    ctx.free_device(d)


def test_one_kernel_select_against_numpy_many_shapes(ctx):
    """The resident (one-kernel, cooperative) select on 80 random inputs of 1 .. 2.1 M samples: latency-like,
    full-range int64, three distinct values, exponential with multi-second outliers; 1..8 ranks; both alignments.
    (A rank re-read after its update made about 3 % of such calls fail before the fix this guards.)"""
    rng = np.random.default_rng(7)
    cap = 2_100_000
    d = ctx.alloc_device(cap * 8 + 16)
    for it in range(80):
        n = int(rng.choice([1, 2, 3, 100, 4097, 65536, 1_000_000, 2_000_001])) if it % 3 else int(rng.integers(1, cap))
        kind = it % 4
        if kind == 0:
            ns = rng.integers(20_000_000, 95_000_000, size=n)
        elif kind == 1:
            ns = rng.integers(-2**62, 2**62, size=n)
        elif kind == 2:
            ns = np.full(n, 57_000_000) + rng.integers(0, 3, size=n)
        else:
            ns = rng.exponential(5e7, size=n).astype(np.int64)
            ns[rng.integers(0, n, size=max(1, n // 200))] = 9_000_000_000
        ns = ns.astype(np.int64)
        off = int(rng.integers(0, 2))
        ctx.h2d(d + 8 * off, ns)
        ks = sorted(int(x) for x in rng.integers(0, n, size=int(rng.integers(1, 9))))
        want = [int(x) for x in np.partition(ns, ks)[ks]]
        assert lat.select_kth(d + 8 * off, n, 1, ks, ctx=ctx) == want, (it, n, kind, off, ks)
    ctx.free_device(d)


def test_select_kth_avg_is_select_kth_stats_without_the_squares(ctx):
    """b200_select_kth_avg (what latency.percentiles calls): same ranks, count, sum, min, max as b200_select_kth_stats
    on the sampled route (n >= 2^24) and on the short route; its sums of squares are 0 by contract."""
    import ctypes as C
    L, h = ctx._L, ctx._h
    for n in ((1 << 24) + 77, 300_001):
        ns = oracle.latency_ns(n, seed=n)
        d = ctx.alloc_device(n * 8)
        ctx.h2d(d, ns)
        karr = np.array([n // 5, n // 2, 9 * n // 10, 99 * n // 100], dtype=np.uint64)
        res = []
        for fn in (L.b200_select_kth_stats, L.b200_select_kth_avg):
            out, st = np.zeros(4, dtype=np.int64), pkg._lib.LatStats()
            pkg._lib.check(fn(h, 0, d, n, 1000, karr.ctypes.data, 4, out.ctypes.data, C.byref(st)))
            res.append((list(out), st.count, st.sum, st.sum_hi, st.min, st.max, st.sumsq_lo, st.sumsq_hi))
        assert res[0][:6] == res[1][:6]
        assert res[1][6:] == (0, 0) and res[0][6] != 0
        us = ns // 1000
        srt = np.partition(ns, [int(k) for k in karr])
        assert res[1][0] == [int(srt[int(k)]) // 1000 for k in karr]
        assert (res[1][1], res[1][2], res[1][4], res[1][5]) == (n, int(us.sum()), int(us.min()), int(us.max()))
        ctx.free_device(d)


def test_radix_count_with_a_key_base_per_histogram(ctx):
    """b200_radix_count_based against numpy: three histograms over the same samples, each with its own key base
    (the lower end of its bracket), a common 20-bit key range; first pass (prefix 0) and a second pass."""
    n = 300_007
    ns = oracle.latency_ns(n, seed=5)
    d = ctx.alloc_device(n * 8)
    ctx.h2d(d, ns)
    keys = ns.view(np.uint64) ^ np.uint64(1 << 63)
    bases = [int(np.uint64(b) ^ np.uint64(1 << 63)) for b in (20_000_000, 47_000_000, 71_500_000)]
    top = 20
    for shift, prefixes in ((8, [0, 0, 0]), (0, [5, 300, 4095])):
        barr, parr = np.array(bases, dtype=np.uint64), np.array(prefixes, dtype=np.uint64)
        counts = np.zeros(3 * lat.RADIX_BINS, dtype=np.uint64)
        t = top if shift else 8
        pkg._lib.check(ctx._L.b200_radix_count_based(ctx._h, 0, d, n, barr.ctypes.data, shift, t, 3, parr.ctypes.data,
                                                     counts.ctypes.data))
        for i in range(3):
            k = keys - np.uint64(bases[i])
            sel = k[(k >> np.uint64(t)) == np.uint64(prefixes[i])]
            dg = ((sel >> np.uint64(shift)) & np.uint64((1 << (t - shift)) - 1)).astype(np.int64)
            want = np.bincount(dg, minlength=lat.RADIX_BINS).astype(np.uint64)
            assert (counts[i * lat.RADIX_BINS:(i + 1) * lat.RADIX_BINS] == want).all(), (shift, i)
    ctx.free_device(d)


def test_sample_and_bounds_in_one_call(ctx):
    """b200_select_sample_bounds: the same stratified sample as b200_select_sample; with slack 0 its order statistics
    exactly, with slack > 0 bounds from below / above that are off by at most `slack` sample ranks or 4096 ns."""
    import ctypes as C
    n, stride = 6_000_003, 16
    ns = oracle.latency_ns(n, seed=17)
    d = upload(ctx, ns)
    L, h = ctx._L, ctx._h
    p, cnt = C.c_void_p(), C.c_size_t()
    assert L.b200_select_sample(h, 0, d, n, stride, C.byref(p), C.byref(cnt)) == 0
    s = cnt.value
    sample = np.empty(s, dtype=np.int64)
    ctx.d2h(sample, p.value)
    srt = np.sort(sample)
    ranks = np.array([0, 7, s // 5, s // 2, s // 2 + 1, 9 * s // 10, 99 * s // 100, s - 1, 3, 4], dtype=np.uint64)
    upper = np.array([0, 1, 0, 1, 0, 1, 0, 1, 1, 0], dtype=np.uint8)
    for slack in (0, 40):
        out = np.zeros(ranks.size, dtype=np.int64)
        got_s = C.c_size_t()
        pkg._lib.check(L.b200_select_sample_bounds(h, 0, d, n, stride, ranks.ctypes.data, upper.ctypes.data, ranks.size,
                                                   slack, out.ctypes.data, C.byref(got_s)))
        assert got_s.value == s
        for r, u, v in zip(ranks, upper, out):
            exact = int(srt[int(r)])
            if slack == 0:
                assert int(v) == exact
            elif u:
                between = int(np.searchsorted(srt, v, "right")) - int(r) - 1
                assert v >= exact and (v - exact < 4096 or between <= slack), (r, v, exact, between)
            else:
                between = int(r) - int(np.searchsorted(srt, v, "left"))
                assert v <= exact and (exact - v < 4096 or between <= slack), (r, v, exact, between)
    bad = np.array([s], dtype=np.uint64)
    assert L.b200_select_sample_bounds(h, 0, d, n, stride, bad.ctypes.data, None, 1, 0, out.ctypes.data,
                                       C.byref(got_s)) == pkg._lib.ERR_INVALID
    ctx.free_device(d)
