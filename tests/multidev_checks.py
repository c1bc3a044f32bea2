"""Parity checks of the one-process-several-GPUs entry points (b200_crc32c_sharded, b200_hist_fixed_multi,
b200_select_kth_multi) on DISTINCT devices, against the CPU oracle.

Shared by tests/test_crc_gpu.py (which can only run it where two GPUs are visible) and by bench.py, whose rank 0
runs it before the timed region whenever the box shows at least two GPUs, so that these entry points meet real
second devices in the multi-GPU bench runs as well.  Test infrastructure: uses oracle/ as the checker.
"""
import ctypes as C
import importlib

import numpy as np

import oracle

pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c = pkg.crc32c


def ngpu() -> int:
    n = C.c_int()
    return n.value if pkg._lib.load().b200_device_count(C.byref(n)) == 0 else 0


def multi_device_parity(nd: int) -> dict:
    """Raises AssertionError on any mismatch; returns what was checked."""
    c = crc32c.Context(list(range(nd)))
    try:
        # BASELINE config 3 in its one-process form: shards on different GPUs, host-side GF(2) fold
        total = (96 << 20) + 40
        host = oracle.payload(total)
        bounds = [total * i // nd // 8 * 8 for i in range(nd)] + [total]
        ptrs, lens = [], []
        for dev in range(nd):
            a, b = bounds[dev], bounds[dev + 1]
            p = c.alloc_device(b - a, dev)
            c.h2d(p, host[a:b], dev)
            ptrs.append(p)
            lens.append(b - a)
        assert crc32c.ShardedChecksum(ptrs, lens, c) == oracle.crc32c(host), "sharded CRC differs from the oracle"
        # the multi-device latency entry points
        lat = pkg.latency
        n = 400_000
        ns = oracle.latency_ns(n)
        cuts = [n * i // nd for i in range(nd + 1)]
        dptr = []
        for dev in range(nd):
            part = np.ascontiguousarray(ns[cuts[dev]:cuts[dev + 1]])
            p = c.alloc_device(part.nbytes, dev)
            c.h2d(p, part, dev)
            dptr.append(p)
        devs = (C.c_int * nd)(*range(nd))
        pp = (C.c_void_p * nd)(*dptr)
        sz = (C.c_size_t * nd)(*[cuts[i + 1] - cuts[i] for i in range(nd)])
        counts = np.zeros(15, dtype=np.uint64)
        st = pkg._lib.LatStats()
        e = lat.README_EDGES
        pkg._lib.check(c._L.b200_hist_fixed_multi(c._h, devs, pp, sz, nd, 1_000_000, e.ctypes.data, e.size, 0,
                                                  counts.ctypes.data, C.byref(st)))
        ms = oracle.ns_to_ms(ns)
        ref, dropped = oracle.hist_numpy(ms.astype(np.float64))
        assert list(counts) == list(ref) and st.count == n, "multi-device histogram differs from the oracle"
        assert st.dropped_below + st.dropped_above == dropped and st.sum == int(ms.sum())
        assert (st.min, st.max) == (int(ms.min()), int(ms.max()))
        ks = np.array([n // 5, n // 2, 9 * n // 10, 99 * n // 100], dtype=np.uint64)
        out = np.zeros(4, dtype=np.int64)
        pkg._lib.check(c._L.b200_select_kth_multi(c._h, devs, pp, sz, nd, 1000, ks.ctypes.data, 4, out.ctypes.data))
        srt = np.sort(oracle.ns_to_us(ns))
        assert list(out) == [int(srt[int(k)]) for k in ks], "multi-device order statistics differ from the oracle"
        for dev in range(nd):
            c.free_device(ptrs[dev], dev)
            c.free_device(dptr[dev], dev)
    finally:
        c.close()
    return {"devices": nd, "entry_points": ["b200_crc32c_sharded", "b200_hist_fixed_multi", "b200_select_kth_multi"]}
