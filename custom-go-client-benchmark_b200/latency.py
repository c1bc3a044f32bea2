"""Host-side mirror of the reference's latency reduction, over the C ABI.

  hist_readme      README.md:15-35   numpy.histogram(x, range(20,100,5)) of per-read milliseconds
  hist_opencensus  metrics_exporter.go:22-34  ochttp.DefaultLatencyDistribution buckets + distribution data
  percentiles      benchmark-script/ssd_test/main.go:147-163  avg/P20/P50/P90/p99/min/max by floor index
Samples live on the device as int64 nanoseconds (time.Duration); `divisor` restates Duration.Milliseconds()
(main.go:146) or .Microseconds() (ssd_test/main.go:79).  With world_size > 1 (one process per GPU) the
bin / digit counts are summed across ranks with torch.distributed (NCCL on GPUs, gloo in CPU tests).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from fractions import Fraction
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import HIST_NUMPY, HIST_OPENCENSUS, RADIX_BINS, SELECT_MAX_K, LatStats, check
from .crc32c import Context, default_context

README_EDGES = np.array(list(range(20, 100, 5)), dtype=np.float64)  # README.md:21-23 (15 bins, 20..95)
OC_DEFAULT_LATENCY_BOUNDS = np.array(
    [1, 2, 3, 4, 5, 6, 8, 10, 13, 16, 20, 25, 30, 40, 50, 65, 80, 100, 130, 160, 200, 250, 300, 400, 500, 650, 800,
     1000, 2000, 5000, 10000, 20000, 50000, 100000], dtype=np.float64)  # ochttp.DefaultLatencyDistribution
NS_PER_MS, NS_PER_US = 1_000_000, 1_000
_SIGN = 1 << 63
_M64 = (1 << 64) - 1


@dataclass
class Distribution:
    """Shape of OpenCensus view.DistributionData (what metrics_exporter.go:22-34 exports)."""
    count: int
    min: int
    max: int
    sum: int                 # exact integer sum of the scaled samples
    sum_wrapped: int         # the same reduced to int64 the way Go's `sum += x` wraps
    sumsq: int               # exact integer sum of squares
    counts: np.ndarray
    dropped_below: int = 0
    dropped_above: int = 0

    @property
    def mean(self) -> float:
        return float(Fraction(self.sum, self.count)) if self.count else 0.0

    @property
    def sum_of_squared_dev(self) -> float:
        if not self.count:
            return 0.0
        return float(Fraction(self.sumsq) - Fraction(self.sum * self.sum, self.count))

    def to_opencensus(self, bounds=None) -> dict:
        """The fields of OpenCensus view.DistributionData that metrics_exporter.go:22-34 would export
        (SURVEY 8f.2), from the exact integer partials.  Count == 0 keeps OpenCensus' zero values."""
        return {"Count": self.count, "Min": float(self.min) if self.count else 0.0,
                "Max": float(self.max) if self.count else 0.0, "Mean": self.mean,
                "SumOfSquaredDev": self.sum_of_squared_dev, "CountPerBucket": [int(c) for c in self.counts],
                "bounds": None if bounds is None else [float(b) for b in bounds]}

    @property
    def avg_go_int(self) -> int:
        """sum/int64(size) with Go's truncating integer division (ssd_test/main.go:157)."""
        s = self.sum_wrapped
        q = abs(s) // self.count
        return q if s >= 0 else -q


def _signed(lo: int, hi: int) -> int:
    v = (hi << 64) | lo if hi >= 0 else ((hi + (1 << 64)) << 64) | lo
    if hi < 0:
        v -= 1 << 128
    return v


class _Reducer:
    """Sums small integer arrays over ranks.  No-op when not distributed."""

    # how many collectives the reductions of this process issued and the wall-clock seconds spent in them
    # (bench.py reports both beside the multi-rank timings)
    tally = {"collectives": 0, "seconds": 0.0}
    _buffers: dict = {}

    def __init__(self, group=None, device=None):
        self.group, self.device = group, device
        try:
            import torch.distributed as dist
            self.dist = dist if dist.is_available() and dist.is_initialized() else None
        except Exception:
            self.dist = None

    @property
    def active(self):
        return self.dist is not None and self.dist.get_world_size(self.group) > 1

    def _run(self, arr: np.ndarray, op: str) -> np.ndarray:
        if not self.active:
            return arr
        import time

        import torch
        t0 = time.perf_counter()
        a = np.ascontiguousarray(arr).astype(np.int64, copy=False)
        if self.device is None:
            t = torch.from_numpy(a.copy())
            self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op), group=self.group)
            out = t.numpy()
        else:
            # pinned staging and a device buffer kept per (device, length): see gather()
            key = (str(self.device), a.size, "reduce")
            buf = _Reducer._buffers.get(key)
            if buf is None:
                buf = (torch.empty(a.size, dtype=torch.int64).pin_memory(),
                       torch.empty(a.size, dtype=torch.int64, device=self.device))
                _Reducer._buffers[key] = buf
            pin, dev = buf
            pin.numpy()[:] = a.reshape(-1)
            dev.copy_(pin, non_blocking=True)
            self.dist.all_reduce(dev, op=getattr(self.dist.ReduceOp, op), group=self.group)
            pin.copy_(dev, non_blocking=True)
            torch.cuda.current_stream(dev.device).synchronize()
            out = pin.numpy().copy().reshape(a.shape)
        _Reducer.tally["collectives"] += 1
        _Reducer.tally["seconds"] += time.perf_counter() - t0
        return out

    def gather(self, values) -> list:
        """Every rank's list of Python ints (each within 64 bits, signed or unsigned; all lists equally long), as
        a list of lists of ints in [0, 2^64) — ONE collective (all_gather), from which sums, minima and maxima of
        any width are then formed locally.  Not distributed: [values]."""
        vals = [int(v) & _M64 for v in values]
        if not self.active:
            return [vals]
        import time

        import torch
        t0 = time.perf_counter()
        world = self.dist.get_world_size(self.group)
        arr = np.array(vals, dtype=np.uint64).view(np.int64)
        if self.device is None:
            mine = torch.from_numpy(arr.copy())
            out = torch.empty(world * mine.numel(), dtype=torch.int64)
            self.dist.all_gather_into_tensor(out, mine, group=self.group)
            rows = out.numpy().view(np.uint64).reshape(world, -1)
        else:
            # pinned staging and device buffers are kept per (device, length): a collective of a few hundred bytes
            # is all overhead, and allocating four tensors was a good part of it
            key = (str(self.device), arr.size, world)
            buf = _Reducer._buffers.get(key)
            if buf is None:
                buf = (torch.empty(arr.size, dtype=torch.int64).pin_memory(),
                       torch.empty(arr.size, dtype=torch.int64, device=self.device),
                       torch.empty(world * arr.size, dtype=torch.int64, device=self.device),
                       torch.empty(world * arr.size, dtype=torch.int64).pin_memory())
                _Reducer._buffers[key] = buf
            pin_in, dev_in, dev_out, pin_out = buf
            pin_in.numpy()[:] = arr
            dev_in.copy_(pin_in, non_blocking=True)
            self.dist.all_gather_into_tensor(dev_out, dev_in, group=self.group)
            pin_out.copy_(dev_out, non_blocking=True)
            torch.cuda.current_stream(dev_out.device).synchronize()
            rows = pin_out.numpy().view(np.uint64).reshape(world, -1)
        _Reducer.tally["collectives"] += 1
        _Reducer.tally["seconds"] += time.perf_counter() - t0
        return [[int(x) for x in r] for r in rows]

    def sum_u64(self, arr: np.ndarray) -> np.ndarray:
        # split into 32-bit halves so the int64 all-reduce cannot overflow for < 2^31 ranks
        a = arr.astype(np.uint64)
        halves = np.concatenate([(a & np.uint64(0xFFFFFFFF)).astype(np.int64), (a >> np.uint64(32)).astype(np.int64)])
        r = self._run(halves, "SUM").astype(object)  # one collective for both halves
        return r[:a.size] + r[a.size:] * (1 << 32)  # python ints

    def sum_counts(self, arr: np.ndarray) -> np.ndarray:
        """Sum of sample counts over ranks as int64: one collective and no Python-int arithmetic.  For counts
        only (their total is the number of samples, far below 2^63); 64-bit limbs go through sum_u64."""
        return self._run(np.ascontiguousarray(arr).astype(np.int64), "SUM")

    def min_i64(self, v: int) -> int:
        return int(self._run(np.array([v], dtype=np.int64), "MIN")[0])

    def max_i64(self, v: int) -> int:
        return int(self._run(np.array([v], dtype=np.int64), "MAX")[0])

    def minmax_i64(self, mn: int, mx: int):
        """(min over ranks of mn, max over ranks of mx) in one collective: MIN of [mn, ~mx] (~ reverses the order
        of int64 without overflowing at INT64_MIN)."""
        r = self._run(np.array([mn, ~mx], dtype=np.int64), "MIN")
        return int(r[0]), ~int(r[1])


def _moment_words(counts, cnt, mn, mx, total, sumsq, below, above) -> list:
    """One rank's exact partials as 64-bit words for _Reducer.gather: counts..., cnt, below, above, the signed
    128-bit sum (two words, two's complement), the 128-bit sum of squares (two words), min, max."""
    t = int(total) & ((1 << 128) - 1)
    q = int(sumsq)
    return ([int(c) for c in counts] + [int(cnt), int(below), int(above), t & _M64, t >> 64, q & _M64, q >> 64,
                                        int(mn), int(mx)])


def _merge_moment_words(rows, nb: int):
    """The ranks' _moment_words -> (counts, cnt, mn, mx, total, sumsq, below, above) of all of them."""
    s64 = lambda w: w - (1 << 64) if w >= _SIGN else w
    counts = np.array([sum(r[i] for r in rows) for i in range(nb)], dtype=np.uint64)
    cnt, below, above = (sum(r[nb + i] for r in rows) for i in range(3))
    total = 0
    for r in rows:
        t = r[nb + 3] | (r[nb + 4] << 64)
        total += t - (1 << 128) if t >> 127 else t
    sumsq = sum(r[nb + 5] | (r[nb + 6] << 64) for r in rows)
    have = [r for r in rows if r[nb]]  # ranks that hold samples
    mn = min((s64(r[nb + 7]) for r in have), default=(1 << 63) - 1)
    mx = max((s64(r[nb + 8]) for r in have), default=-(1 << 63))
    return counts, cnt, mn, mx, total, sumsq, below, above


def merge_distribution(red: _Reducer, counts: np.ndarray, cnt: int, mn: int, mx: int, total: int, sumsq: int,
                       below: int, above: int) -> Distribution:
    """Combine one rank's exact integer partials with the other ranks' (sum / min / max all-reduces)."""
    nb = counts.size
    if red.active:
        rows = red.gather(_moment_words(counts, cnt, mn, mx, total, sumsq, below, above))  # ONE collective
        counts, cnt, mn, mx, total, sumsq, below, above = _merge_moment_words(rows, nb)
    wrapped = total & _M64
    wrapped = wrapped - (1 << 64) if wrapped >= _SIGN else wrapped
    return Distribution(cnt, mn, mx, total, wrapped, sumsq, counts, below, above)


def hist_fixed(d_samples: int, n: int, divisor: int, edges, mode: int, ctx: Optional[Context] = None,
               dev: Optional[int] = None, group=None, reduce_device=None) -> Distribution:
    ctx = ctx or default_context()
    edges = np.ascontiguousarray(edges, dtype=np.float64)
    nb = edges.size - 1 if mode == HIST_NUMPY else edges.size + 1
    counts = np.zeros(nb, dtype=np.uint64)
    st = LatStats()
    check(ctx._L.b200_hist_fixed(ctx._h, ctx._dev(dev), d_samples, n, divisor, edges.ctypes.data, edges.size, mode,
                                 counts.ctypes.data, C.byref(st)))
    total = _signed(st.sum, st.sum_hi)
    sumsq = (st.sumsq_hi << 64) | st.sumsq_lo
    return merge_distribution(_Reducer(group, reduce_device), counts, st.count, st.min, st.max, total, sumsq,
                              st.dropped_below, st.dropped_above)


def hist_fixed_pair(d_samples: int, n: int, divisor: int, edges_a, mode_a: int, edges_b, mode_b: int,
                    ctx: Optional[Context] = None, dev: Optional[int] = None, group=None, reduce_device=None):
    """Two histograms of the same samples (same divisor) from one pass over them: (Distribution a, Distribution b),
    each as hist_fixed would return it."""
    ctx = ctx or default_context()
    ea, eb = np.ascontiguousarray(edges_a, dtype=np.float64), np.ascontiguousarray(edges_b, dtype=np.float64)
    ca = np.zeros(ea.size - 1 if mode_a == HIST_NUMPY else ea.size + 1, dtype=np.uint64)
    cb = np.zeros(eb.size - 1 if mode_b == HIST_NUMPY else eb.size + 1, dtype=np.uint64)
    dropped_b = np.zeros(2, dtype=np.uint64)
    st = LatStats()
    check(ctx._L.b200_hist_fixed_pair(ctx._h, ctx._dev(dev), d_samples, n, divisor, ea.ctypes.data, ea.size, mode_a,
                                      ca.ctypes.data, eb.ctypes.data, eb.size, mode_b, cb.ctypes.data,
                                      dropped_b.ctypes.data, C.byref(st)))
    total, sumsq = _signed(st.sum, st.sum_hi), (st.sumsq_hi << 64) | st.sumsq_lo
    red = _Reducer(group, reduce_device)
    return (merge_distribution(red, ca, st.count, st.min, st.max, total, sumsq, st.dropped_below, st.dropped_above),
            merge_distribution(red, cb, st.count, st.min, st.max, total, sumsq, int(dropped_b[0]), int(dropped_b[1])))


def hist_readme_and_opencensus(d_samples_ns: int, n: int, **kw):
    """README.md:21-23,32 and metrics_exporter.go:28 on per-read milliseconds, from one pass.  (Integer milliseconds,
    as main.go:146 records them: see hist_readme.)"""
    return hist_fixed_pair(d_samples_ns, n, NS_PER_MS, README_EDGES, HIST_NUMPY, OC_DEFAULT_LATENCY_BOUNDS,
                           HIST_OPENCENSUS, **kw)


def hist_readme(d_samples_ns: int, n: int, **kw) -> Distribution:
    """README.md:21-23,32 on per-read milliseconds as main.go:146 records them: trunc(ns / 10^6), INTEGER
    milliseconds.  That is the script's answer when its input lines are whole milliseconds; on the fractional
    Duration strings of execute_pb.sh the script bins float(line) instead — a sample of 95.3 ms is dropped there and
    lands in the closed last bin [90, 95] here — and that form is parse_latency_text + hist_fixed_f64."""
    return hist_fixed(d_samples_ns, n, NS_PER_MS, README_EDGES, HIST_NUMPY, **kw)


def hist_opencensus(d_samples_ns: int, n: int, **kw) -> Distribution:
    """metrics_exporter.go:28 distribution view on per-read milliseconds."""
    return hist_fixed(d_samples_ns, n, NS_PER_MS, OC_DEFAULT_LATENCY_BOUNDS, HIST_OPENCENSUS, **kw)


def radix_select(red: _Reducer, n_local: int, minmax_fn, count_fn, divisor: int, ks: Sequence[int],
                 n_global: Optional[int] = None, global_bounds: Optional[tuple] = None):
    """MSD radix select over the union of all ranks' samples.

    minmax_fn() -> (min, max) of this rank's raw samples; count_fn(key_base, shift, top, prefixes) -> uint64
    array [len(prefixes)][4096] of this rank's digit counts (b200_minmax / b200_radix_count on the GPU).
    After every local count the histograms are summed over ranks, so every rank walks the same digits.
    """
    ks = [int(k) for k in ks]
    ntot = int(red.sum_u64(np.array([n_local], dtype=np.uint64))[0]) if (n_global is None and red.active) else (
        n_local if n_global is None else n_global)
    if ntot == 0:
        raise ZeroDivisionError("percentiles of an empty sample set (Go panics at ssd_test/main.go:157)")
    if any(k >= ntot or k < 0 for k in ks):
        raise ValueError("k >= n")
    if global_bounds is not None:  # bounds every rank already agrees on: no collective needed
        gmn, gmx = int(global_bounds[0]), int(global_bounds[1])
    else:
        mn, mx = minmax_fn()
        gmn, gmx = red.minmax_i64(mn, mx)
    kmin, kmax = (gmn + _SIGN) & _M64, (gmx + _SIGN) & _M64
    bits = (kmax - kmin).bit_length()
    res = []
    for g in range(0, len(ks), SELECT_MAX_K):
        grp = ks[g:g + SELECT_MAX_K]
        m = len(grp)
        prefix, rank = [0] * m, list(grp)
        top, first = bits, True
        while top > 0:
            shift = max(top - 12, 0)
            width = top - shift
            # first pass: key' < 2^top, so every sample matches prefix 0 and one histogram serves all k
            pfx = [0] if first else prefix
            tot = red.sum_counts(np.asarray(count_fn(kmin, shift, top, pfx), dtype=np.uint64).reshape(-1))
            for i in range(m):
                row = tot[0:RADIX_BINS] if first else tot[i * RADIX_BINS:(i + 1) * RADIX_BINS]
                cs = np.cumsum(row[:1 << width])
                dgt = int(np.searchsorted(cs, rank[i], side="right"))  # first digit whose running count exceeds the rank
                if dgt == 1 << width:
                    raise RuntimeError("inconsistent radix counts")
                rank[i] -= int(cs[dgt - 1]) if dgt else 0
                prefix[i] = (prefix[i] << width) | dgt
            top, first = shift, False
        for i in range(m):
            ns = ((prefix[i] + kmin) & _M64) - _SIGN  # undo the sign flip
            q = abs(ns) // divisor                    # Go's truncating division
            res.append(q if ns >= 0 else -q)
    return res


SAMPLED_SELECT_MIN = 1 << 24      # total samples from which the multi-rank select tries one sampled pass first
_I64_MIN, _I64_MAX = -(1 << 63), (1 << 63) - 1


def merge_brackets(brackets):
    """[(lo, hi)] closed intervals -> ascending, disjoint, non-touching intervals (b200_select_kth does the same)."""
    out = []
    for lo, hi in sorted((int(a), int(b)) for a, b in brackets):
        if out and lo <= out[-1][1] + 1:
            out[-1][1] = max(out[-1][1], hi)
        else:
            out.append([lo, hi])
    return [(a, b) for a, b in out]


def locate_in_brackets(counts, k: int):
    """counts[c] = samples with exactly c thresholds (lo0, hi0+1, lo1, hi1+1, ...) below or at them, so odd c are
    the brackets.  Returns the 0-based rank of global rank k among the kept (odd-bin) samples, or None when k fell
    between brackets (a miss)."""
    below = kept_before = 0
    for j in range((len(counts) - 1) // 2):
        below += int(counts[2 * j])
        if k < below:
            return None
        inside = int(counts[2 * j + 1])
        if k < below + inside:
            return kept_before + (k - below)
        below += inside
        kept_before += inside
    return None


def locate_bracket(counts, k: int):
    """As locate_in_brackets, but (bracket index, 0-based rank of global rank k among that bracket's samples)."""
    below = 0
    for j in range((len(counts) - 1) // 2):
        below += int(counts[2 * j])
        if k < below:
            return None
        inside = int(counts[2 * j + 1])
        if k < below + inside:
            return j, k - below
        below += inside
    return None


def bracket_radix_select(red: _Reducer, count_fn, brackets, located, divisor: int):
    """The order statistics among the kept samples, each resolved INSIDE its own bracket: keys are taken relative to
    the bracket's lower end, so they are as many bits wide as the widest bracket (~18 for 10^9 latencies) and two
    12-bit passes settle them — against three over the span of all brackets together.  located[i] = (bracket
    index, rank inside that bracket).  A pass is one local count (count_fn with a LIST of key bases, one per wanted
    rank; up to SELECT_MAX_K per call) and ONE all-reduce of the digit counts.  (A key range is the width of the
    widest bracket rounded up to a power of two: what it takes in beyond a bracket's upper end is larger than every
    sample of the bracket and cannot change a rank inside it; what lies below the lower end wraps to a huge key.)"""
    m = len(located)
    kbase = [(int(brackets[j][0]) + _SIGN) & _M64 for j, _ in located]
    top = max((int(brackets[j][1]) - int(brackets[j][0])).bit_length() for j, _ in located)
    prefix, rank = [0] * m, [int(r) for _, r in located]
    while top > 0:
        shift = max(top - 12, 0)
        width = top - shift
        parts = [np.asarray(count_fn(kbase[g:g + SELECT_MAX_K], shift, top, prefix[g:g + SELECT_MAX_K]),
                            dtype=np.uint64).reshape(-1) for g in range(0, m, SELECT_MAX_K)]
        tot = red.sum_counts(np.concatenate(parts))
        for i in range(m):
            cs = np.cumsum(tot[i * RADIX_BINS: i * RADIX_BINS + (1 << width)])
            dgt = int(np.searchsorted(cs, rank[i], side="right"))
            if dgt == 1 << width:
                raise RuntimeError("inconsistent radix counts")
            rank[i] -= int(cs[dgt - 1]) if dgt else 0
            prefix[i] = (prefix[i] << width) | dgt
        top = shift
    res = []
    for i in range(m):
        ns = ((prefix[i] + kbase[i]) & _M64) - _SIGN
        q = abs(ns) // divisor  # Go's truncating division
        res.append(q if ns >= 0 else -q)
    return res


def sampled_select(red: _Reducer, n_local: int, n_global: Optional[int], divisor: int, ks: Sequence[int], sample_fn,
                   sample_select_fn, bracket_fn, kept_count_fn, kept_minmax_fn, fractions=None, extra_fn=None,
                   min_total: int = 0):
    """One pass over every rank's samples instead of min/max + three radix passes (the multi-rank form of what
    b200_select_kth does on one device).  FOUR collectives of a few dozen integers (the last two a few
    thousand): the brackets, the bracket-pass counts, two radix passes inside the brackets.

      sample_fn(stride) -> s                 this rank's stratified sample (one input per `stride`), its size
      sample_select_fn(ranks, upper, slack) -> values   order statistics of that sample (local), or bounds on them
                                             to within `slack` sample ranks: from below, from above where upper[i]
      bracket_fn(thresholds, cap) -> (counts, nkept, stats)   the histogram pass that keeps the odd bins
      kept_count_fn / kept_minmax_fn         radix-select building blocks over the kept samples
      fractions                              [(num, den)]: the wanted ranks are n_global * num // den; n_global may
                                             then be None (it comes out of the first collective) and ks is ignored
      extra_fn(stats) -> [ints]              words gathered from every rank along with the counts (the caller's
                                             moments): returned as one list per rank
      min_total                              None is returned (by every rank alike) when the ranks hold fewer

    Every rank brackets each wanted quantile with ITS sample, a few standard deviations either side; the union of
    the ranks' brackets (min of the lower ends, max of the upper ends) holds the global order statistic whenever
    every local bracket holds the local quantile, whatever the ranks' distributions are - it is just wide when
    they differ, and then (or on a rare tail event) the exact counts of the pass say so and None is returned:
    the caller takes the radix passes.  Returns (values, stats, extra rows) otherwise."""
    world = red.dist.get_world_size(red.group) if red.active else 1
    nk = len(fractions) if fractions is not None else len(ks)
    # (without the total, ranks are taken to hold about equal shares: it only sizes the sample)
    stride = max(16, -(-(n_global if n_global is not None else n_local * world) // (1 << 22)))
    s = sample_fn(stride)
    ok = s >= 1024
    margin = int(2.5 * np.sqrt(max(s, 1))) + 2
    want, upper, where = [], [], []
    for i in range(nk):
        c = fractions[i][0] * s // fractions[i][1] if fractions is not None else int(ks[i]) * s // n_global
        lo_at = hi_at = None
        if ok and c >= margin:
            lo_at = len(want); want.append(c - margin); upper.append(False)
        if ok and c + margin + 1 < s:
            hi_at = len(want); want.append(c + margin + 1); upper.append(True)
        where.append((lo_at, hi_at))
    # (bounds good to a sixteenth of the margin are enough for bracket ends, and a radix round cheaper)
    got = sample_select_fn(want, upper, max(1, margin // 16)) if want else []
    s64 = lambda w: w - (1 << 64) if w >= _SIGN else w
    vec = [int(n_local), 1 if ok else 0]
    vec += [_I64_MIN if a is None else int(got[a]) for a, _ in where]
    vec += [_I64_MAX - 1 if b is None else min(int(got[b]), _I64_MAX - 1) for _, b in where]
    rows = red.gather(vec)                                                                   # collective 1
    n_all = sum(r[0] for r in rows)
    if n_global is not None and n_all != n_global:
        raise ValueError("sampled_select: n_global is not the sum of the ranks' sample counts")
    n_global = n_all
    if not all(r[1] for r in rows) or n_global == 0 or n_global < min_total:
        return None  # some rank holds too few samples to bracket anything
    ks = [n_global * num // den for num, den in fractions] if fractions is not None else [int(k) for k in ks]
    br = merge_brackets([(min(s64(r[2 + i]) for r in rows), max(s64(r[2 + nk + i]) for r in rows)) for i in range(nk)])
    thr = []
    for lo, hi in br:
        thr += [lo, hi + 1]
    # room for several times the expected number of samples inside this rank's share of the brackets (the union
    # over ranks is wider than one rank's own), at least n/64; more than n/8 is not worth keeping
    expect = nk * (2 * margin + 2) * stride
    cap = max(min(max(8 * expect, n_local // 64), n_local // 8), 1024)
    counts, nkept, stats = bracket_fn(thr, cap)
    extra = [int(x) for x in extra_fn(stats)] if extra_fn is not None else []
    nc = len(thr) + 1
    rows = red.gather([int(c) for c in counts] + [int(nkept), 1 if nkept > cap else 0] + extra)  # collective 2
    gc = [sum(r[i] for r in rows) for i in range(nc)]
    kept_all, overflow = sum(r[nc] for r in rows), sum(r[nc + 1] for r in rows)
    extra_rows = [r[nc + 2:] for r in rows]
    inside = sum(gc[1::2])
    if overflow or sum(gc) != n_global or kept_all != inside or inside == 0:
        return None  # brackets too heavy for the buffers (heavy duplicates, ranks with very different data)
    located = [locate_bracket(gc, k) for k in ks]
    if any(x is None for x in located):
        return None
    if br[0][0] == _I64_MIN or br[-1][1] == _I64_MAX - 1:
        # a bracket open at an end is as wide as int64: the passes run over all kept samples between their extremes
        rank_kept = [locate_in_brackets(gc, k) for k in ks]
        vals = radix_select(red, int(nkept), kept_minmax_fn, kept_count_fn, divisor, rank_kept, n_global=inside)
    else:
        vals = bracket_radix_select(red, kept_count_fn, br, located, divisor)              # collectives 3, 4
    return vals, stats, extra_rows


def _sampled_select_on_gpu(red, L, h, d, d_samples, n, n_global, divisor, ks, want_stats, fractions=None,
                           extra_fn=None, min_total=0):
    """sampled_select over this rank's device samples through the C ABI building blocks."""
    state = {}

    def sample_fn(stride):  # (the sample is taken together with its order statistics, below: one call, one round trip)
        state["stride"] = stride
        return n // stride

    def sample_select_fn(ranks, upper=None, slack=0):
        karr = np.array(ranks, dtype=np.uint64)
        uarr = np.array([1 if u else 0 for u in (upper or [0] * karr.size)], dtype=np.uint8)
        out = np.zeros(karr.size, dtype=np.int64)
        cnt = C.c_size_t()
        check(L.b200_select_sample_bounds(h, d, d_samples, n, state["stride"], karr.ctypes.data, uarr.ctypes.data, karr.size,
                                          int(slack), out.ctypes.data, C.byref(cnt)))
        return [int(x) for x in out]

    def bracket_fn(thr, cap):
        tarr = np.array(thr, dtype=np.int64)
        counts = np.zeros(tarr.size + 1, dtype=np.uint64)
        p, nkept, st = C.c_void_p(), C.c_uint64(), LatStats()
        # want_stats == "avg": without the sum of squares (the seven lines need the sum only)
        fn = L.b200_bracket_pass_avg if want_stats == "avg" else L.b200_bracket_pass
        check(fn(h, d, d_samples, n, divisor, tarr.ctypes.data, tarr.size, cap, counts.ctypes.data,
                 C.byref(p), C.byref(nkept), C.byref(st) if want_stats else None))
        state["kept"] = (p.value, min(nkept.value, cap))
        return counts, nkept.value, (st if want_stats else None)

    def kept_count_fn(key_base, shift, top, prefixes):
        p, cnt = state["kept"]
        counts = np.zeros(len(prefixes) * RADIX_BINS, dtype=np.uint64)
        parr = np.array(prefixes, dtype=np.uint64)
        if isinstance(key_base, (list, tuple)):  # a key base per histogram (bracket_radix_select)
            barr = np.array(key_base, dtype=np.uint64)
            check(L.b200_radix_count_based(h, d, p, cnt, barr.ctypes.data, shift, top, len(prefixes), parr.ctypes.data,
                                           counts.ctypes.data))
        else:
            check(L.b200_radix_count(h, d, p, cnt, key_base, shift, top, len(prefixes), parr.ctypes.data,
                                     counts.ctypes.data))
        return counts

    def kept_minmax_fn():
        p, cnt = state["kept"]
        mn, mx = C.c_int64(), C.c_int64()
        check(L.b200_minmax(h, d, p, cnt, C.byref(mn), C.byref(mx)))
        return mn.value, mx.value

    return sampled_select(red, n, n_global, divisor, ks, sample_fn, sample_select_fn, bracket_fn, kept_count_fn,
                          kept_minmax_fn, fractions=fractions, extra_fn=extra_fn, min_total=min_total)


def select_kth(d_samples: int, n: int, divisor: int, ks: Sequence[int], ctx: Optional[Context] = None,
               dev: Optional[int] = None, group=None, reduce_device=None, n_global: Optional[int] = None,
               bounds_ns: Optional[tuple] = None, try_sampled: bool = True):
    """k-th smallest (0-based) of trunc(sample/divisor) over all ranks' samples (sorted[k] of
    ssd_test/main.go:158-163) without sorting.  bounds_ns = (lo, hi) with lo <= every sample <= hi on every
    rank skips the min/max pass."""
    ctx = ctx or default_context()
    L, h, d = ctx._L, ctx._h, ctx._dev(dev)
    red = _Reducer(group, reduce_device)
    if not red.active:
        karr = np.array([int(k) for k in ks], dtype=np.uint64)
        out = np.zeros(karr.size, dtype=np.int64)
        if bounds_ns is None:
            check(L.b200_select_kth(h, d, d_samples, n, divisor, karr.ctypes.data, karr.size, out.ctypes.data))
        else:
            check(L.b200_select_kth_range(h, d, d_samples, n, divisor, int(bounds_ns[0]), int(bounds_ns[1]),
                                          karr.ctypes.data, karr.size, out.ctypes.data))
        return [int(x) for x in out]

    ntot = n_global if n_global is not None else int(red.sum_u64(np.array([n], dtype=np.uint64))[0])
    if (try_sampled and ntot >= SAMPLED_SELECT_MIN and all(0 <= int(k) < ntot for k in ks)
            and 0 < len(ks) <= 2 * SELECT_MAX_K):
        r = _sampled_select_on_gpu(red, L, h, d, d_samples, n, ntot, divisor, ks, False)
        if r is not None:
            return r[0]
    n_global = ntot

    def minmax_fn():
        if bounds_ns is not None:
            return int(bounds_ns[0]), int(bounds_ns[1])
        mn, mx = C.c_int64(), C.c_int64()
        check(L.b200_minmax(h, d, d_samples, n, C.byref(mn), C.byref(mx)))
        return mn.value, mx.value

    def count_fn(key_base, shift, top, prefixes):
        counts = np.zeros(len(prefixes) * RADIX_BINS, dtype=np.uint64)
        parr = np.array(prefixes, dtype=np.uint64)
        check(L.b200_radix_count(h, d, d_samples, n, key_base, shift, top, len(prefixes), parr.ctypes.data,
                                 counts.ctypes.data))
        return counts

    return radix_select(red, n, minmax_fn, count_fn, divisor, ks, n_global)


@dataclass
class Percentiles:
    """The seven lines ssd_test prints (benchmark-script/ssd_test/main.go:157-163), in sample units."""
    avg: int
    p20: int
    p50: int
    p90: int
    p99: int
    min: int
    max: int


def percentiles(d_samples_ns: int, n: int, divisor: int = NS_PER_US, ctx: Optional[Context] = None,
                dev: Optional[int] = None, group=None, reduce_device=None,
                dist: Optional[Distribution] = None) -> Percentiles:
    """avg / P20 / P50 / P90 / p99 / min / max of trunc(sample/divisor) exactly as ssd_test/main.go:147-163 prints
    them.  `dist` may be a Distribution already computed over the same samples WITH THE SAME divisor (any
    edges): its count, sum and min/max are reused instead of a second pass over the data."""
    if dist is None and not _Reducer(group, reduce_device).active:
        # one process: ranks and moments from one call (one pass over long inputs), ssd_test/main.go:147-163
        if n == 0:
            raise ZeroDivisionError("percentiles of an empty sample set (Go panics at ssd_test/main.go:157)")
        ctx_ = ctx or default_context()
        karr = np.array([n // 5, n // 2, (9 * n) // 10, (99 * n) // 100], dtype=np.uint64)
        out = np.zeros(4, dtype=np.int64)
        st = LatStats()
        # (the seven lines need the sum, not the sum of squares: b200_select_kth_avg)
        check(ctx_._L.b200_select_kth_avg(ctx_._h, ctx_._dev(dev), d_samples_ns, n, divisor, karr.ctypes.data, 4,
                                          out.ctypes.data, C.byref(st)))
        whole = merge_distribution(_Reducer(), np.zeros(0, dtype=np.uint64), st.count, st.min, st.max,
                                   _signed(st.sum, st.sum_hi), 0, 0, 0)
        return Percentiles(whole.avg_go_int, int(out[0]), int(out[1]), int(out[2]), int(out[3]), whole.min, whole.max)
    if dist is None:
        # several ranks: one sampled pass gives the ranks and the moments; the radix passes if it misses
        red = _Reducer(group, reduce_device)
        ctx_ = ctx or default_context()
        # (whether the ranks together hold enough samples for this route comes out of its first collective: every
        # rank takes the same decision)
        r = _sampled_select_on_gpu(red, ctx_._L, ctx_._h, ctx_._dev(dev), d_samples_ns, n, None, divisor, None, "avg",
                                   fractions=[(1, 5), (1, 2), (9, 10), (99, 100)],
                                   extra_fn=lambda st: _moment_words([], st.count, st.min, st.max,
                                                                     _signed(st.sum, st.sum_hi),
                                                                     (st.sumsq_hi << 64) | st.sumsq_lo, 0, 0),
                                   min_total=SAMPLED_SELECT_MIN)
        if r is not None:  # ranks, and every rank's moments from the same two collectives
            (p20, p50, p90, p99), _, rows = r
            _, cnt, mn, mx, total, sumsq, _, _ = _merge_moment_words(rows, 0)
            if cnt == 0:
                raise ZeroDivisionError("percentiles of an empty sample set (Go panics at ssd_test/main.go:157)")
            wrapped = total & _M64
            whole = Distribution(cnt, mn, mx, total, wrapped - (1 << 64) if wrapped >= _SIGN else wrapped, sumsq,
                                 np.zeros(0, dtype=np.uint64), 0, 0)
            return Percentiles(whole.avg_go_int, p20, p50, p90, p99, whole.min, whole.max)
    dist_ = dist if dist is not None else hist_fixed(d_samples_ns, n, divisor, OC_DEFAULT_LATENCY_BOUNDS,
                                                     HIST_OPENCENSUS, ctx=ctx, dev=dev, group=group,
                                                     reduce_device=reduce_device)
    size = dist_.count
    if size == 0:
        raise ZeroDivisionError("percentiles of an empty sample set (Go panics at ssd_test/main.go:157)")
    ks = [size // 5, size // 2, (9 * size) // 10, (99 * size) // 100]
    # raw-nanosecond bounds implied by the scaled min/max of the pass above (truncating division)
    lo = dist_.min * divisor if dist_.min > 0 else (dist_.min - 1) * divisor + 1
    hi = (dist_.max + 1) * divisor - 1 if dist_.max >= 0 else dist_.max * divisor
    lo, hi = max(lo, -(1 << 63)), min(hi, (1 << 63) - 1)
    # (a sampled pass was tried above when the ranks hold enough samples: not again)
    p20, p50, p90, p99 = select_kth(d_samples_ns, n, divisor, ks, ctx=ctx, dev=dev, group=group,
                                    reduce_device=reduce_device, n_global=size, bounds_ns=(lo, hi),
                                    try_sampled=dist is not None)
    return Percentiles(dist_.avg_go_int, p20, p50, p90, p99, dist_.min, dist_.max)


# ---- latency text files (execute_pb.sh:4,8 -> README.md:15-35) ------------------------------------------------

@dataclass
class ParsedLatencyFile:
    """Result of README.md:25-28's loop over one latency file."""
    nlines: int
    first_bad_line: Optional[int]   # where `float(line)` raises ValueError in the reference script
    nbad: int
    ninexact: int                   # lines outside the exactly-rounded domain (never produced by the pipeline)
    d_values: int                   # device float64 array, nlines entries
    d_status: int                   # device uint8 array


def go_duration_line(ns: int) -> str:
    """What one line of grpc_N.txt looks like: Go's Duration.String() after `tr 'ms' ' '` (execute_pb.sh:4,8)."""
    def fmt(v: int, unit_digits: int) -> str:
        whole, frac = divmod(v, 10 ** unit_digits)
        s = str(whole)
        if frac:
            s += "." + str(frac).rjust(unit_digits, "0").rstrip("0")
        return s
    neg, u = ns < 0, abs(ns)
    if u == 0:
        out = "0s"
    elif u < 1_000:
        out = f"{u}ns"
    elif u < 1_000_000:
        out = fmt(u, 3) + "µs"
    elif u < 1_000_000_000:
        out = fmt(u, 6) + "ms"
    else:
        h, rem = divmod(u, 3_600_000_000_000)
        m, rem = divmod(rem, 60_000_000_000)
        sec = fmt(rem, 9) + "s"
        out = (f"{h}h{m}m" if h else (f"{m}m" if m else "")) + sec
    out = ("-" if neg else "") + out
    return out.replace("m", " ").replace("s", " ")


def parse_latency_text(d_text: int, nbytes: int, max_lines: int, ctx: Optional[Context] = None,
                       dev: Optional[int] = None) -> ParsedLatencyFile:
    """README.md:25-28 on the device.  The caller frees d_values / d_status with ctx.free_device."""
    ctx = ctx or default_context()
    d_values = ctx.alloc_device(max(max_lines, 1) * 8, dev)
    d_status = ctx.alloc_device(max(max_lines, 1), dev)
    nl, fb, nb, ni = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
    try:
        check(ctx._L.b200_parse_latency_text(ctx._h, ctx._dev(dev), d_text, nbytes, d_values, d_status, max_lines,
                                             C.byref(nl), C.byref(fb), C.byref(nb), C.byref(ni)))
    except Exception:
        ctx.free_device(d_values, dev)
        ctx.free_device(d_status, dev)
        raise
    first = None if fb.value == (1 << 64) - 1 else int(fb.value)
    return ParsedLatencyFile(int(nl.value), first, int(nb.value), int(ni.value), d_values, d_status)


@dataclass
class FloatHistogram:
    counts: np.ndarray
    dropped_below: int
    dropped_above: int
    nan_count: int
    sum: float           # correctly rounded sum of the samples (double-double accumulation)
    n: int

    @property
    def mean(self) -> float:
        """README.md:30 `sum(x) / len(x)`."""
        return self.sum / self.n


def hist_fixed_f64(d_values: int, n: int, edges, mode: int = HIST_NUMPY, ctx: Optional[Context] = None,
                   dev: Optional[int] = None) -> FloatHistogram:
    """README.md:32 `plt.hist(x, bins)` (numpy.histogram) or the OpenCensus bucket rule on float64 samples."""
    ctx = ctx or default_context()
    edges = np.ascontiguousarray(edges, dtype=np.float64)
    nb = edges.size - 1 if mode == HIST_NUMPY else edges.size + 1
    counts = np.zeros(nb, dtype=np.uint64)
    below, above, nans = C.c_uint64(), C.c_uint64(), C.c_uint64()
    hi, lo = C.c_double(), C.c_double()
    check(ctx._L.b200_hist_fixed_f64(ctx._h, ctx._dev(dev), d_values, n, edges.ctypes.data, edges.size, mode,
                                     counts.ctypes.data, C.byref(below), C.byref(above), C.byref(nans), C.byref(hi),
                                     C.byref(lo)))
    return FloatHistogram(counts, int(below.value), int(above.value), int(nans.value), hi.value + lo.value, n)
