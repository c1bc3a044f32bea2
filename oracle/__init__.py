"""ctypes face of the CPU oracle (oracle/oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference`` legs
may import this module, and only as the checker or the timed CPU baseline.  The product package never
imports it.  Parity status: unpinned by the reference (no tests, CRC lives in un-vendored Go modules);
pinned against RFC 3720 CRC-32C vectors and numpy.histogram (tests/test_oracle.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("oracle.c", "oracle.h")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B"])
    return _SO


class Stats(C.Structure):
    _fields_ = [("count", C.c_uint64), ("min", C.c_int64), ("max", C.c_int64), ("sum", C.c_uint64),
                ("mean_welford", C.c_double), ("sum_sq_dev", C.c_double), ("mean_seq", C.c_double)]


class Percentiles(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("avg", "p20", "p50", "p90", "p99", "min", "max")]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        vp, u32, u64, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t
        for name in ("bitwise", "slice8", "hw"):
            f = getattr(L, f"oracle_crc32c_{name}")
            f.restype, f.argtypes = u32, [u32, vp, sz]
        L.oracle_crc32c_hw_available.restype = C.c_int
        L.oracle_crc32c_combine.restype, L.oracle_crc32c_combine.argtypes = u32, [u32, u32, u64]
        L.oracle_xpow8.restype, L.oracle_xpow8.argtypes = u32, [u64]
        L.oracle_gf2_mul.restype, L.oracle_gf2_mul.argtypes = u32, [u32, u32]
        L.oracle_verify_objects.restype = u64
        L.oracle_verify_objects.argtypes = [vp, vp, vp, sz, vp, vp, vp]
        L.oracle_verify_objects_mt.restype = u64
        L.oracle_verify_objects_mt.argtypes = [vp, vp, vp, sz, vp, vp, vp, C.c_int]
        L.oracle_splitmix64.restype, L.oracle_splitmix64.argtypes = u64, [u64, u64]
        L.oracle_fill_payload.restype, L.oracle_fill_payload.argtypes = None, [vp, u64, sz, u64]
        L.oracle_latency_ns.restype, L.oracle_latency_ns.argtypes = C.c_int64, [u64, u64]
        L.oracle_fill_latency_ns.restype, L.oracle_fill_latency_ns.argtypes = None, [vp, u64, sz, u64]
        L.oracle_hist_numpy.restype, L.oracle_hist_numpy.argtypes = u64, [vp, sz, vp, C.c_int, vp]
        L.oracle_hist_opencensus.restype, L.oracle_hist_opencensus.argtypes = None, [vp, sz, vp, C.c_int, vp]
        L.oracle_stats_i64.restype, L.oracle_stats_i64.argtypes = None, [vp, sz, C.POINTER(Stats)]
        L.oracle_percentiles_i64.restype = C.c_int
        L.oracle_percentiles_i64.argtypes = [vp, sz, C.POINTER(Percentiles)]
        L.oracle_select_kth_i64.restype = C.c_int
        L.oracle_select_kth_i64.argtypes = [vp, sz, vp, C.c_int, vp]
        L.oracle_ns_to_ms.restype, L.oracle_ns_to_ms.argtypes = None, [vp, sz, vp]
        L.oracle_ns_to_us.restype, L.oracle_ns_to_us.argtypes = None, [vp, sz, vp]
        _lib = L
    return _lib


SEED = 0xC0FFEE
OC_BOUNDS = np.array([1, 2, 3, 4, 5, 6, 8, 10, 13, 16, 20, 25, 30, 40, 50, 65, 80, 100, 130, 160, 200, 250, 300,
                      400, 500, 650, 800, 1000, 2000, 5000, 10000, 20000, 50000, 100000], dtype=np.float64)
README_EDGES = np.array(list(range(20, 100, 5)), dtype=np.float64)  # README.md:21-23


def _buf(b):
    """(address, nbytes) of bytes / bytearray / numpy array without copying."""
    if isinstance(b, np.ndarray):
        a = np.ascontiguousarray(b)
        return a, a.ctypes.data, a.nbytes
    a = np.frombuffer(b, dtype=np.uint8)
    return a, a.ctypes.data if a.size else 0, a.size


def crc32c(data, crc: int = 0, impl: str = "hw") -> int:
    keep, addr, n = _buf(data)
    return int(getattr(lib(), f"oracle_crc32c_{impl}")(crc, addr, n))


def combine(c1: int, c2: int, len2: int) -> int:
    return int(lib().oracle_crc32c_combine(c1, c2, len2))


def xpow8(n: int) -> int:
    return int(lib().oracle_xpow8(n))


def gf2_mul(a: int, b: int) -> int:
    return int(lib().oracle_gf2_mul(a, b))


def verify_objects(arena: np.ndarray, off, length, want=None, threads: int = 1):
    off = np.ascontiguousarray(off, dtype=np.uint64)
    length = np.ascontiguousarray(length, dtype=np.uint64)
    n = off.size
    got = np.zeros(n, dtype=np.uint32)
    ok = np.zeros(n, dtype=np.uint8)
    w = None if want is None else np.ascontiguousarray(want, dtype=np.uint32)
    bad = lib().oracle_verify_objects_mt(arena.ctypes.data, off.ctypes.data, length.ctypes.data, n,
                                         None if w is None else w.ctypes.data, got.ctypes.data, ok.ctypes.data,
                                         threads)
    return got, ok.astype(bool), int(bad)


def payload(nbytes: int, seed: int = SEED, first_word: int = 0) -> np.ndarray:
    """splitmix64 payload, nbytes rounded up to 8 internally then truncated."""
    nw = (nbytes + 7) // 8
    a = np.empty(nw, dtype=np.uint64)
    lib().oracle_fill_payload(a.ctypes.data, first_word, nw, seed)
    return a.view(np.uint8)[:nbytes]


def latency_ns(n: int, seed: int = SEED, first: int = 0) -> np.ndarray:
    a = np.empty(n, dtype=np.int64)
    lib().oracle_fill_latency_ns(a.ctypes.data, first, n, seed)
    return a


def ns_to_ms(ns: np.ndarray) -> np.ndarray:
    out = np.empty_like(ns)
    lib().oracle_ns_to_ms(ns.ctypes.data, ns.size, out.ctypes.data)
    return out


def ns_to_us(ns: np.ndarray) -> np.ndarray:
    out = np.empty_like(ns)
    lib().oracle_ns_to_us(ns.ctypes.data, ns.size, out.ctypes.data)
    return out


def hist_numpy(x, edges=README_EDGES):
    x = np.ascontiguousarray(x, dtype=np.float64)
    edges = np.ascontiguousarray(edges, dtype=np.float64)
    counts = np.zeros(edges.size - 1, dtype=np.uint64)
    dropped = lib().oracle_hist_numpy(x.ctypes.data, x.size, edges.ctypes.data, edges.size, counts.ctypes.data)
    return counts, int(dropped)


def hist_opencensus(x, bounds=OC_BOUNDS):
    x = np.ascontiguousarray(x, dtype=np.float64)
    bounds = np.ascontiguousarray(bounds, dtype=np.float64)
    counts = np.zeros(bounds.size + 1, dtype=np.uint64)
    lib().oracle_hist_opencensus(x.ctypes.data, x.size, bounds.ctypes.data, bounds.size, counts.ctypes.data)
    return counts


def stats_i64(v) -> Stats:
    v = np.ascontiguousarray(v, dtype=np.int64)
    s = Stats()
    lib().oracle_stats_i64(v.ctypes.data, v.size, C.byref(s))
    return s


def percentiles_i64(v) -> Percentiles:
    v = np.ascontiguousarray(v, dtype=np.int64)
    p = Percentiles()
    rc = lib().oracle_percentiles_i64(v.ctypes.data, v.size, C.byref(p))
    if rc:
        raise ZeroDivisionError("percentiles of an empty sample set (Go panics at ssd_test/main.go:157)")
    return p


def select_kth_i64(v, ks):
    v = np.ascontiguousarray(v, dtype=np.int64)
    ks = np.ascontiguousarray(ks, dtype=np.uint64)
    out = np.zeros(ks.size, dtype=np.int64)
    rc = lib().oracle_select_kth_i64(v.ctypes.data, v.size, ks.ctypes.data, ks.size, out.ctypes.data)
    if rc:
        raise ValueError(f"select_kth failed rc={rc}")
    return out
