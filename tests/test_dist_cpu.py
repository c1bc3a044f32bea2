"""world_size-2 gloo tests (CPU) of the host-side multi-rank logic: shard CRC folding, the count all-reduce
behind the histogram, and the distributed radix-select loop.  The per-rank GPU primitives are replaced here by
numpy stand-ins with the same contract (b200_minmax / b200_radix_count / b200_hist_fixed partials); the GPU
primitives themselves are covered by the -m gpu tests."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
lat = pkg.latency
SIGN, M64 = 1 << 63, (1 << 64) - 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _np_count_fn(samples):
    keys = (samples.astype(np.int64).view(np.uint64) ^ np.uint64(SIGN))

    def count_fn(key_base, shift, top, prefixes):
        # key_base: one for all histograms (b200_radix_count) or a list, one each (b200_radix_count_based)
        bases = key_base if isinstance(key_base, (list, tuple)) else [key_base] * len(prefixes)
        out = np.zeros((len(prefixes), lat.RADIX_BINS), dtype=np.uint64)
        width = top - shift
        for i, p in enumerate(prefixes):
            k = (keys - np.uint64(bases[i])).astype(np.uint64)
            sel = k if top >= 64 else k[(k >> np.uint64(top)) == np.uint64(p)]
            dg = ((sel >> np.uint64(shift)) & np.uint64((1 << width) - 1)).astype(np.int64)
            out[i] = np.bincount(dg, minlength=lat.RADIX_BINS).astype(np.uint64)
        return out.reshape(-1)
    return count_fn


def _np_sampled_fns(samples, divisor, seed):
    """numpy stand-ins with the contracts of b200_select_sample / b200_select_kth on the sample / b200_bracket_pass
    and the radix building blocks over the kept samples."""
    st = {}

    def sample_fn(stride):
        rng = np.random.default_rng(seed)
        s = samples.size // stride
        st["sample"] = np.sort(samples[np.arange(s) * stride + rng.integers(0, stride, size=s)]) if s else samples[:0]
        return s

    def sample_select_fn(ranks, upper=None, slack=0):  # (exact order statistics are bounds with no slack)
        return [int(st["sample"][r]) for r in ranks]

    def bracket_fn(thr, cap):
        c = np.searchsorted(np.array(thr, dtype=np.int64), samples, side="right")  # thresholds <= sample
        counts = np.bincount(c, minlength=len(thr) + 1).astype(np.uint64)
        kept = samples[c % 2 == 1]
        st["kept"] = kept[:cap]
        return counts, kept.size, None

    return (sample_fn, sample_select_fn, bracket_fn, lambda *a: _np_count_fn(st["kept"])(*a),
            lambda: (int(st["kept"].min()), int(st["kept"].max())) if st["kept"].size else ((1 << 63) - 1, -(1 << 63)))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # --- 1. sharded object: rank r hashes bytes [r*L/N, (r+1)*L/N); rank 0 folds (crc, len) pairs ------
        L = 3_000_001
        data = oracle.payload(L)
        lo, hi = L * rank // world, L * (rank + 1) // world
        mine = oracle.crc32c(data[lo:hi])            # stand-in for b200_crc32c_update_device on this rank
        gathered = [None] * world
        dist.all_gather_object(gathered, (mine, hi - lo))
        acc = 0
        for c, ln in gathered:
            acc = pkg.crc32c.Combine(acc, c, ln)     # the product's host-side GF(2) fold
        assert acc == oracle.crc32c(data)

        # --- 2. histogram merge ---------------------------------------------------------------------------
        n = 100_000
        ns_all = oracle.latency_ns(n)
        ns = ns_all[n * rank // world: n * (rank + 1) // world]
        ms = oracle.ns_to_ms(ns)
        counts, dropped = oracle.hist_numpy(ms.astype(np.float64))
        below, above = int((ms < 20).sum()), int((ms > 95).sum())
        red = lat._Reducer()
        assert red.active
        d = lat.merge_distribution(red, counts, ms.size, int(ms.min()), int(ms.max()), int(ms.sum()),
                                   int((ms.astype(object) ** 2).sum()), below, above)
        ms_all = oracle.ns_to_ms(ns_all)
        ref, ref_dropped = oracle.hist_numpy(ms_all.astype(np.float64))
        assert list(d.counts) == list(ref) and d.dropped_below + d.dropped_above == ref_dropped
        assert (d.count, d.min, d.max, d.sum) == (n, ms_all.min(), ms_all.max(), int(ms_all.sum()))
        assert d.sumsq == int((ms_all.astype(object) ** 2).sum())
        s = oracle.stats_i64(ms_all)
        assert abs(d.mean - s.mean_welford) <= 1e-12 * d.mean
        assert abs(d.sum_of_squared_dev - s.sum_sq_dev) <= 1e-9 * s.sum_sq_dev

        # --- 3. distributed radix select ------------------------------------------------------------------
        for divisor in (1000, 1):
            ks = [0, n // 5, n // 2, 9 * n // 10, 99 * n // 100, n - 1]
            got = lat.radix_select(red, ns.size, lambda: (int(ns.min()), int(ns.max())), _np_count_fn(ns), divisor, ks)
            srt = np.sort(ns_all // divisor)
            assert got == [int(srt[k]) for k in ks], (divisor, got)
        # negatives + duplicates + an empty rank
        rng = np.random.default_rng(1)
        all_v = np.concatenate([rng.integers(-5000, 5000, 900), np.full(300, -17), rng.integers(-2**62, 2**62, 10)]
                               ).astype(np.int64)
        part = all_v if rank == 0 else all_v[:0]
        big = (1 << 63) - 1
        mm = (lambda: (int(part.min()), int(part.max()))) if part.size else (lambda: (big, -big - 1))
        got = lat.radix_select(red, part.size, mm, _np_count_fn(part), 1, [0, 600, all_v.size - 1])
        srt = np.sort(all_v)
        assert got == [int(srt[0]), int(srt[600]), int(srt[-1])]
        # --- 4. sampled select: one bracket pass per rank, a handful of integers per collective --------------
        n = 4_000_000                                 # (below ~10^6 the brackets of four ranks are wider than the buffers)
        ns_all = oracle.latency_ns(n)
        for split in (n // 2, n // 3):                # equal and unequal shares
            ns = ns_all[:split] if rank == 0 else ns_all[split:]
            ks = [n // 5, n // 2, 9 * n // 10, 99 * n // 100]
            r = lat.sampled_select(red, ns.size, n, 1000, ks, *_np_sampled_fns(ns, 1000, rank))
            assert r is not None, "the brackets should catch the ranks on spread-out data"
            srt = np.sort(ns_all // 1000)
            assert r[0] == [int(srt[k]) for k in ks], r[0]
        # the same by fractions (the total comes out of the first collective), with words of the caller's riding
        # along in the second: what latency.percentiles does across ranks — four collectives in all
        ns = ns_all[:n // 3] if rank == 0 else ns_all[n // 3:]
        before = lat._Reducer.tally["collectives"]
        r = lat.sampled_select(red, ns.size, None, 1000, None, *_np_sampled_fns(ns, 1000, rank),
                               fractions=[(1, 5), (1, 2), (9, 10), (99, 100)], extra_fn=lambda st: [rank + 1, ns.size],
                               min_total=1 << 20)
        used = lat._Reducer.tally["collectives"] - before
        assert r is not None and used in (4, 5), (r is None, used)  # (a wide p99 bracket takes a third radix pass)
        srt = np.sort(ns_all // 1000)
        assert r[0] == [int(srt[n * a // b]) for a, b in ((1, 5), (1, 2), (9, 10), (99, 100))]
        assert r[2] == [[1, n // 3], [2, n - n // 3]]
        # too few samples for the route: every rank says so after the first collective
        assert lat.sampled_select(red, ns.size, None, 1000, None, *_np_sampled_fns(ns, 1000, rank),
                                  fractions=[(1, 2)], min_total=n + 1) is None
        # extreme ranks open the brackets at the ends; ranks whose data differ widen them: exact or a clean miss
        shifted = ns_all[:n // 2] if rank == 0 else ns_all[n // 2:] + 3_000_000
        both = np.concatenate([ns_all[:n // 2], ns_all[n // 2:] + 3_000_000])
        for ks in ([0, n // 2, n - 1], [n // 3]):
            r = lat.sampled_select(red, shifted.size, n, 1, ks, *_np_sampled_fns(shifted, 1, rank))
            if r is not None:
                srt = np.sort(both)
                assert r[0] == [int(srt[k]) for k in ks]
        # heavy duplicates: brackets too heavy for the buffers -> None on every rank (the caller falls back)
        coarse = (ns_all // 5_000_000) * 5_000_000
        part = coarse[:n // 2] if rank == 0 else coarse[n // 2:]
        assert lat.sampled_select(red, part.size, n, 1, [n // 2], *_np_sampled_fns(part, 1, rank)) is None
        # a rank with almost nothing cannot bracket: None everywhere
        part = ns_all[:100] if rank == 0 else ns_all[100:]
        assert lat.sampled_select(red, part.size, n, 1, [n // 2], *_np_sampled_fns(part, 1, rank)) is None
        q.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_radix_select_single_process_matches_sort():
    red = lat._Reducer()
    assert not red.active
    rng = np.random.default_rng(0)
    v = rng.integers(0, 10**10, size=5000).astype(np.int64)
    ks = [0, 1, 2500, 4999]
    got = lat.radix_select(red, v.size, lambda: (int(v.min()), int(v.max())), _np_count_fn(v), 1000, ks)
    srt = np.sort(v // 1000)
    assert got == [int(srt[k]) for k in ks]
    with pytest.raises(ZeroDivisionError):
        lat.radix_select(red, 0, lambda: (0, 0), _np_count_fn(v[:0]), 1, [0])
    with pytest.raises(ValueError):
        lat.radix_select(red, v.size, lambda: (int(v.min()), int(v.max())), _np_count_fn(v), 1, [5000])


def test_moment_words_round_trip_and_merge():
    """The 64-bit words a rank contributes to the one-collective merges: negative 128-bit sums, sums of squares
    above 2^64, extremes across ranks, a rank without samples (its extremes must not count)."""
    a = lat._moment_words([3, 0, 7], 10, -5, 40, -(1 << 70) - 3, (1 << 100) + 9, 1, 2)
    b = lat._moment_words([1, 1, 1], 3, -9, 12, (1 << 70) + 1, (1 << 64) - 1, 0, 5)
    empty = lat._moment_words([0, 0, 0], 0, (1 << 63) - 1, -(1 << 63), 0, 0, 0, 0)
    assert all(0 <= w < 1 << 64 for w in [int(x) & ((1 << 64) - 1) for x in a + b + empty])
    rows = [[int(x) & ((1 << 64) - 1) for x in r] for r in (a, empty, b)]  # what _Reducer.gather hands back
    counts, cnt, mn, mx, total, sumsq, below, above = lat._merge_moment_words(rows, 3)
    assert list(counts) == [4, 1, 8] and (cnt, below, above) == (13, 1, 7)
    assert (mn, mx) == (-9, 40)
    assert total == -2 and sumsq == (1 << 100) + 9 + (1 << 64) - 1
    # nobody holds a sample
    _, cnt, mn, mx, _, _, _, _ = lat._merge_moment_words([rows[1], rows[1]], 3)
    assert cnt == 0 and mn == (1 << 63) - 1 and mx == -(1 << 63)
    # not distributed: gather is the identity on one row
    assert lat._Reducer().gather([1, -1, (1 << 64) - 1]) == [[1, (1 << 64) - 1, (1 << 64) - 1]]


def test_bracket_radix_select_single_process():
    """Order statistics resolved inside their own brackets (keys relative to the bracket's lower end, one key base per
    wanted rank): against a sort, with brackets of different widths, touching key ranges and a one-value bracket."""
    rng = np.random.default_rng(3)
    v = np.concatenate([rng.integers(1000, 1400, 3000), rng.integers(5000, 5001, 50), rng.integers(70_000, 90_000, 2000),
                        rng.integers(-500, -100, 700)]).astype(np.int64)
    brackets = [(-500, -101), (1000, 1399), (5000, 5000), (70_000, 89_999)]
    srt = np.sort(v)
    red = lat._Reducer()
    located, want = [], []
    for j, (lo, hi) in enumerate(brackets):
        inside = srt[(srt >= lo) & (srt <= hi)]
        for r in (0, inside.size // 2, inside.size - 1):
            located.append((j, r))
            want.append(int(inside[r]))
    for divisor in (1, 7):
        got = lat.bracket_radix_select(red, _np_count_fn(v), brackets, located, divisor)
        assert got == [abs(w) // divisor * (1 if w >= 0 else -1) for w in want]
