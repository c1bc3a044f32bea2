"""Generates tests/golden/golden.json.

The reference is Go (no toolchain here) and has no fixtures, so these vectors are produced by
independent Python statements of the same rules: a pure-Python bitwise CRC-32C, numpy.histogram (what
README.md:32's plt.hist calls), numpy.searchsorted for the OpenCensus bucket rule, and numpy sort +
floor indexing for ssd_test/main.go:147-163.  The synthetic inputs come from the oracle's counter-based
generators (pure-Python restated below so the fixture does not depend on the C code it later checks).
Run:  python tests/golden/make_golden.py
"""
import json
import os

import numpy as np

M = (1 << 64) - 1


def splitmix64(seed, index):
    z = (seed + (index + 1) * 0x9E3779B97F4A7C15) & M
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
    return z ^ (z >> 31)


def payload(nbytes, seed, first_word):
    nw = (nbytes + 7) // 8
    return b"".join(splitmix64(seed, first_word + i).to_bytes(8, "little") for i in range(nw))[:nbytes]


def crc32c(data):
    tab = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ (0x82F63B78 if c & 1 else 0)
        tab.append(c)
    s = 0xFFFFFFFF
    for b in data:
        s = (s >> 8) ^ tab[(s ^ b) & 0xFF]
    return s ^ 0xFFFFFFFF


def latency_ns(seed, i):
    z, z2 = splitmix64(seed, 2 * i), splitmix64(seed, 2 * i + 1)
    s = (z & 0xFFFF) + ((z >> 16) & 0xFFFF) + ((z >> 32) & 0xFFFF) + (z >> 48)
    ns = 20000000 + (s * 75000000) // 262141
    q = z2 & 0xFF
    if q < 64:
        ns = ns // 1000000 * 1000000
    elif q < 128:
        ns = ns // 1000 * 1000
    sel = (z2 >> 8) % 100
    if sel == 0:
        ns = 95000000 + (z2 >> 20) % (10000000000 - 95000000)
    elif sel == 1:
        ns = (z2 >> 20) % 20000000
    return ns


def main():
    seed = 0xC0FFEE
    crc_cases = []
    for nbytes, first_word, skip in [(0, 0, 0), (1, 0, 0), (15, 3, 0), (16, 0, 0), (511, 7, 0), (512, 0, 0),
                                     (513, 0, 5), (4096, 11, 0), (65536 + 3, 0, 1), (200000, 1 << 20, 13)]:
        d = payload(nbytes, seed, first_word)[skip:]
        crc_cases.append(dict(nbytes=nbytes, seed=seed, first_word=first_word, skip=skip, crc32c=crc32c(d)))
    n = 20000
    ns = np.array([latency_ns(seed, i) for i in range(n)], dtype=np.int64)
    ms = (ns // 1000000).astype(np.float64)
    h = np.histogram(ms, list(range(20, 100, 5)))[0]
    oc_bounds = [1, 2, 3, 4, 5, 6, 8, 10, 13, 16, 20, 25, 30, 40, 50, 65, 80, 100, 130, 160, 200, 250, 300, 400,
                 500, 650, 800, 1000, 2000, 5000, 10000, 20000, 50000, 100000]
    oc = np.bincount(np.searchsorted(np.array(oc_bounds, dtype=np.float64), ms, side="right"), minlength=35)
    us = np.sort(ns // 1000)
    pct = [int(us.sum()) // n, int(us[n // 5]), int(us[n // 2]), int(us[9 * n // 10]), int(us[99 * n // 100]),
           int(us[0]), int(us[-1])]
    out = dict(crc=crc_cases,
               latency=dict(n=n, seed=seed, numpy15_counts=[int(x) for x in h], numpy15_dropped=int(n - h.sum()),
                            oc35_counts=[int(x) for x in oc], us_percentiles=pct,
                            ns_first8=[int(x) for x in ns[:8]]))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
