"""Pins the CPU oracle (oracle/oracle.c) before anything is compared against it.

The reference has no tests or golden vectors for this path, so the pins are the public
CRC-32C known answers (RFC 3720 B.4, the "123456789" check value) and numpy.histogram, which is what the
README script's plt.hist calls (README.md:32).
"""
import json
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import oracle

KATS = [
    (b"", 0x00000000),
    (b"a", 0xC1D04330),
    (b"123456789", 0xE3069283),
    (bytes(32), 0x8A9136AA),
    (b"\xff" * 32, 0x62A8AB43),
    (bytes(range(32)), 0x46DD794E),
    (bytes(range(31, -1, -1)), 0x113FDB5C),
    # RFC 3720 B.4: an iSCSI SCSI Read (10) command PDU
    (bytes([0x01, 0xC0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0x14, 0, 0, 0, 0, 0, 0x04, 0, 0, 0, 0, 0x14, 0, 0,
            0, 0x18, 0x28, 0, 0, 0, 0, 0, 0, 0, 0x02, 0, 0, 0, 0, 0, 0, 0]), 0xD9963A56),
]


def py_bitwise(data: bytes, crc: int = 0) -> int:
    """Independent pure-Python statement of CRC-32C (reflected 0x82F63B78)."""
    s = crc ^ 0xFFFFFFFF
    for b in data:
        s ^= b
        for _ in range(8):
            s = (s >> 1) ^ (0x82F63B78 if s & 1 else 0)
    return s ^ 0xFFFFFFFF


@pytest.mark.parametrize("impl", ["bitwise", "slice8", "hw"])
@pytest.mark.parametrize("data,want", KATS)
def test_known_answers(impl, data, want):
    assert oracle.crc32c(data, impl=impl) == want
    assert py_bitwise(data) == want


def test_hw_path_is_real_sse42_here():
    # informational: the CPU baseline is only "Go-like" when the SSE4.2 path is active
    assert oracle.lib().oracle_crc32c_hw_available() in (0, 1)


@settings(max_examples=60, deadline=None)
@given(st.binary(min_size=0, max_size=20000), st.integers(0, 2**32 - 1))
def test_impls_agree(data, crc):
    a = oracle.crc32c(data, crc, "bitwise")
    assert oracle.crc32c(data, crc, "slice8") == a
    assert oracle.crc32c(data, crc, "hw") == a


def test_hw_three_lane_path_large_unaligned():
    buf = oracle.payload(5 * 12288 + 777)
    for start in (0, 1, 3, 7, 9):
        d = buf[start:]
        assert oracle.crc32c(d, impl="hw") == oracle.crc32c(d, impl="slice8")


@settings(max_examples=40, deadline=None)
@given(st.binary(min_size=0, max_size=5000), st.lists(st.integers(0, 5000), max_size=6))
def test_update_chunk_invariance(data, cuts):
    # Update(Update(0,a),b) == Update(0,a||b): the property storage.Reader.Read relies on (main.go:140)
    cuts = sorted(min(c, len(data)) for c in cuts)
    crc, prev = 0, 0
    for c in cuts + [len(data)]:
        crc = oracle.crc32c(data[prev:c], crc)
        prev = c
    assert crc == oracle.crc32c(data)


def test_chunked_37():
    d = bytes(oracle.payload(100000))
    crc = 0
    for i in range(0, len(d), 37):
        crc = oracle.crc32c(d[i:i + 37], crc)
    assert crc == oracle.crc32c(d)


@settings(max_examples=40, deadline=None)
@given(st.binary(max_size=3000), st.binary(max_size=3000))
def test_combine_identity(a, b):
    assert oracle.combine(oracle.crc32c(a), oracle.crc32c(b), len(b)) == oracle.crc32c(a + b)


def test_combine_huge_length_matches_streaming():
    # 2^36-byte scale shift constants: combine with a long run of zero bytes, checked by associativity
    c = oracle.crc32c(b"hello")
    z1 = oracle.combine(c, oracle.crc32c(bytes(1 << 20)), 1 << 20)
    assert z1 == oracle.crc32c(b"hello" + bytes(1 << 20))
    # x^(8(a+b)) == x^(8a) x^(8b)
    a, b = (1 << 36) - 12345, 987654321
    assert oracle.gf2_mul(oracle.xpow8(a), oracle.xpow8(b)) == oracle.xpow8(a + b)


def test_xpow8_small():
    assert oracle.xpow8(0) == 0x80000000
    assert oracle.xpow8(1) == 0x00800000
    assert oracle.xpow8(4) == 0x82F63B78  # x^32 mod P is the polynomial itself


def test_verify_objects_flags_exactly_the_corrupt_one():
    n, sz = 32, 4096
    arena = oracle.payload(n * sz).copy()
    off = np.arange(n, dtype=np.uint64) * sz
    ln = np.full(n, sz, dtype=np.uint64)
    want, ok, bad = oracle.verify_objects(arena, off, ln)
    assert bad == 0 and ok.all()
    arena[5 * sz + 100] ^= 0x10
    got, ok, bad = oracle.verify_objects(arena, off, ln, want, threads=4)
    assert bad == 1 and list(np.nonzero(~ok)[0]) == [5]
    assert got[5] != want[5] and (np.delete(got, 5) == np.delete(want, 5)).all()


# ---- histogram / stats / percentiles -------------------------------------------------------------

def test_readme_edges_shape():
    assert list(oracle.README_EDGES) == [float(x) for x in range(20, 100, 5)]
    assert oracle.README_EDGES.size == 16  # 15 bins, 20..95 (README.md:21-23), not 20..100


def test_hist_numpy_edge_cases():
    x = [19.999, 20, 24.999, 25, 90, 94.999, 95, 95.0001, 100]
    counts, dropped = oracle.hist_numpy(x)
    exp = [2, 1] + [0] * 12 + [3]
    assert list(counts) == exp and dropped == 3
    assert list(np.histogram(x, list(range(20, 100, 5)))[0]) == exp


@settings(max_examples=30, deadline=None)
@given(st.lists(st.floats(min_value=-10, max_value=200, allow_nan=False), max_size=400))
def test_hist_numpy_matches_numpy(x):
    counts, dropped = oracle.hist_numpy(x)
    ref = np.histogram(np.array(x, dtype=np.float64), list(range(20, 100, 5)))[0]
    assert list(counts) == list(ref)
    assert dropped == len(x) - int(ref.sum())


def test_hist_numpy_on_generated_ms():
    ns = oracle.latency_ns(200000)
    ms = oracle.ns_to_ms(ns).astype(np.float64)
    counts, dropped = oracle.hist_numpy(ms)
    ref = np.histogram(ms, list(range(20, 100, 5)))[0]
    assert list(counts) == list(ref) and dropped == ms.size - int(ref.sum())
    assert dropped > 0 and counts.sum() > 0.9 * ms.size


def test_hist_opencensus_rule():
    x = [0, 0.5, 1, 1.5, 99.9, 100, 100000, 1e9, -3]
    c = oracle.hist_opencensus(x)
    assert c.size == 35 and c.sum() == len(x)
    assert c[0] == 3          # 0, 0.5, -3  (< 1)
    assert c[1] == 2          # 1, 1.5      (< 2)
    assert c[17] == 1         # 99.9 < 100
    assert c[18] == 1         # 100 < 130
    assert c[34] == 2         # 100000 and 1e9 overflow
    ref = np.searchsorted(oracle.OC_BOUNDS, np.array(x), side="right")
    assert list(np.bincount(ref, minlength=35)) == list(c)


def test_generator_is_integer_exact_and_in_range():
    ns = oracle.latency_ns(100000)
    assert ns.min() >= 0 and ns.max() < 10_000_000_000
    assert ((ns >= 20_000_000) & (ns < 95_000_000)).mean() > 0.97
    # counter based: any slice regenerates identically
    assert (oracle.latency_ns(1000, first=5000) == ns[5000:6000]).all()
    assert np.unique(ns).size < ns.size  # exact duplicates exist
    assert (oracle.ns_to_ms(ns) == ns // 1_000_000).all()
    assert (oracle.ns_to_us(ns) == ns // 1000).all()


def test_payload_counter_based():
    a = oracle.payload(4096)
    b = oracle.payload(1024, first_word=128)
    assert (a[1024:2048] == b).all()
    assert int(a[:8].view(np.uint64)[0]) == oracle.lib().oracle_splitmix64(oracle.SEED, 0)


@pytest.mark.parametrize("n", [1, 2, 5, 10, 99, 100, 101, 1000])
def test_percentile_index_rule(n):
    rng = np.random.default_rng(n)
    v = rng.integers(0, 10**7, size=n, dtype=np.int64)
    p = oracle.percentiles_i64(v)
    a = np.sort(v)
    assert p.p20 == a[n // 5] and p.p50 == a[n // 2] and p.p90 == a[(9 * n) // 10]
    assert p.p99 == a[(99 * n) // 100] and p.min == a[0] and p.max == a[-1]
    assert p.avg == int(a.sum()) // n
    ks = [0, n // 2, n - 1]
    assert list(oracle.select_kth_i64(v, ks)) == [a[k] for k in ks]


def test_percentiles_empty_raises():
    with pytest.raises(ZeroDivisionError):
        oracle.percentiles_i64(np.zeros(0, dtype=np.int64))


def test_stats():
    v = oracle.ns_to_ms(oracle.latency_ns(50000))
    s = oracle.stats_i64(v)
    assert s.count == v.size and s.min == v.min() and s.max == v.max() and s.sum == int(v.sum())
    assert abs(s.mean_welford - v.mean()) < 1e-9 * abs(v.mean())
    assert abs(s.mean_seq - v.mean()) < 1e-9 * abs(v.mean())
    assert abs(s.sum_sq_dev - ((v - v.mean()) ** 2).sum()) < 1e-6 * s.sum_sq_dev


def test_golden_fixture():
    path = os.path.join(os.path.dirname(__file__), "golden", "golden.json")
    g = json.load(open(path))
    for case in g["crc"]:
        d = oracle.payload(case["nbytes"], seed=case["seed"], first_word=case["first_word"])[case["skip"]:]
        assert oracle.crc32c(d) == case["crc32c"], case
        assert oracle.crc32c(d, impl="slice8") == case["crc32c"]
    lat = g["latency"]
    ns = oracle.latency_ns(lat["n"], seed=lat["seed"])
    ms = oracle.ns_to_ms(ns)
    counts, dropped = oracle.hist_numpy(ms.astype(np.float64))
    assert list(counts) == lat["numpy15_counts"] and dropped == lat["numpy15_dropped"]
    assert list(oracle.hist_opencensus(ms.astype(np.float64))) == lat["oc35_counts"]
    p = oracle.percentiles_i64(oracle.ns_to_us(ns))
    assert [p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max] == lat["us_percentiles"]
