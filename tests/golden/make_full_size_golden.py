#!/usr/bin/env python
"""Full-size goldens for BASELINE.json configs 3 and 5, computed on the CPU only (no GPU, no product code).

    python tests/golden/make_full_size_golden.py [--configs 3,5] [--threads 8]

config3.json  CRC-32C of the 64 GiB splitmix64 stream (seed 0xC0FFEE) that `b200_fill_payload_device`
              regenerates on the GPU: the oracle generates and hashes it 1 GiB at a time on host threads
              (SSE4.2 path; the first piece is also hashed with the table-driven slicing-by-8 path as a
              cross-check), the per-piece CRCs are kept and folded with oracle_crc32c_combine.  The single
              read this stands for is main.go:135-140 (one NewReader drained once).
config5.json  10^9 synthetic latency samples (oracle_fill_latency_ns, seed 0xC0FFEE):
                README.md:21-32   numpy.histogram itself, applied to the distinct millisecond values with their
                                  multiplicities as weights (the same counts as over the 10^9 floats), and the sum
                metrics_exporter.go:28  OpenCensus buckets: the oracle's rule on each distinct value, weighted
                ssd_test/main.go:147-163  numpy.partition of the 10^9 microsecond values at size/5, size/2,
                                  9*size/10, 99*size/100; integer sum / size with Go's truncating division; min; max
Both files are small and committed; bench.py and the GPU tests compare the GPU results with them.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

SEED = 0xC0FFEE


def config3(threads: int, gib: int = 64):
    piece = 1 << 30
    npieces = gib

    def one(i):
        buf = oracle.payload(piece, seed=SEED, first_word=i * (piece // 8))
        c = oracle.crc32c(buf, impl="hw" if oracle.lib().oracle_crc32c_hw_available() else "slice8")
        if i == 0:
            assert c == oracle.crc32c(buf, impl="slice8"), "SSE4.2 and slicing-by-8 paths disagree"
        return c

    t0 = time.time()
    with ThreadPoolExecutor(threads) as ex:
        crcs = list(ex.map(one, range(npieces)))
    total = 0
    for c in crcs:
        total = oracle.combine(total, c, piece)
    shards = {}
    for n in (1, 2, 4, 8):
        per = npieces // n
        out = []
        for r in range(n):
            acc = 0
            for c in crcs[r * per:(r + 1) * per]:
                acc = oracle.combine(acc, c, piece)
            out.append(acc)
        shards[str(n)] = out
    return {"what": "CRC-32C of splitmix64(seed, word index) little-endian words, 64 GiB, and of its 1 GiB pieces",
            "seed": SEED, "bytes": gib << 30, "crc32c": total, "crc32c_hex": f"{total:#010x}",
            "piece_bytes": piece, "piece_crc32c": crcs, "shard_crc32c": shards,
            "made_by": "tests/golden/make_full_size_golden.py (oracle/oracle.c, CPU)", "seconds": round(time.time() - t0, 1)}


def config5(threads: int, n: int = 1_000_000_000):
    t0 = time.time()
    ns = np.empty(n, dtype=np.int64)
    step = 1 << 24

    def fill(lo):
        hi = min(lo + step, n)
        oracle.lib().oracle_fill_latency_ns(ns[lo:hi].ctypes.data, lo, hi - lo, SEED)

    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(fill, range(0, n, step)))
    assert int(ns.min()) >= 0
    # milliseconds (main.go:146): distinct values and multiplicities
    def count_ms(lo):
        return np.bincount(ns[lo:lo + step] // 1_000_000, minlength=10_001)

    with ThreadPoolExecutor(threads) as ex:
        ms_counts = sum(ex.map(count_ms, range(0, n, step)))
    vals = np.nonzero(ms_counts)[0]
    w = ms_counts[vals]
    readme, _ = np.histogram(vals.astype(np.float64), oracle.README_EDGES, weights=w.astype(np.float64))
    readme = [int(x) for x in readme]
    below = int(w[vals < 20].sum())
    above = int(w[vals > 95].sum())
    oc = [0] * (oracle.OC_BOUNDS.size + 1)
    for v, c in zip(vals, w):
        b = oracle.hist_opencensus(np.array([float(v)]))
        oc[int(np.nonzero(b)[0][0])] += int(c)
    ms_sum = int((vals.astype(object) * w.astype(object)).sum())
    ms_sumsq = int((vals.astype(object) ** 2 * w.astype(object)).sum())
    assert sum(readme) + below + above == n and sum(oc) == n
    # microseconds (ssd_test/main.go:74-80,147-163)
    us = ns // 1000
    del ns
    us_sum = 0
    for lo in range(0, n, step):
        us_sum += int(us[lo:lo + step].sum())
    us_min, us_max = int(us.min()), int(us.max())
    ks = [n // 5, n // 2, (9 * n) // 10, (99 * n) // 100]
    us.partition(ks)
    p = [int(us[k]) for k in ks]
    return {"what": "10^9 synthetic latency samples (oracle_fill_latency_ns): README.md:21-32 bins, "
                    "metrics_exporter.go:28 buckets, ssd_test/main.go:147-163 lines",
            "seed": SEED, "samples": n,
            "ms": {"readme15_counts": readme, "dropped_below": below, "dropped_above": above, "oc35_counts": oc,
                   "min": int(vals.min()), "max": int(vals.max()), "sum": ms_sum, "sumsq": ms_sumsq},
            "us": {"avg": us_sum // n, "p20": p[0], "p50": p[1], "p90": p[2], "p99": p[3], "min": us_min,
                   "max": us_max, "sum": us_sum, "ranks": ks},
            "made_by": "tests/golden/make_full_size_golden.py (oracle generator, numpy.histogram, numpy.partition)",
            "seconds": round(time.time() - t0, 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="3,5")
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    args = ap.parse_args()
    for c in args.configs.split(","):
        obj = config3(args.threads) if c == "3" else config5(args.threads)
        path = os.path.join(HERE, f"config{c}.json")
        with open(path, "w") as f:
            json.dump(obj, f, indent=1)
            f.write("\n")
        print(path, {k: v for k, v in obj.items() if k in ("crc32c_hex", "seconds", "us")})


if __name__ == "__main__":
    main()
