"""ctypes binding of libb200crc.so (the C ABI in include/b200crc.h).

There is no fallback: if the shared library is missing or a GPU call fails, the error is raised.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libb200crc.so")

OK, ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_EMPTY, ERR_TOO_LARGE = 0, -1, -2, -3, -4, -5
HIST_NUMPY, HIST_OPENCENSUS = 0, 1
RADIX_BINS, SELECT_MAX_K = 4096, 4
NOALLOC_MAX_BYTES = 1 << 35


class B200Error(RuntimeError):
    def __init__(self, status: int, name: str, detail: str):
        super().__init__(f"{name} (status {status}): {detail}")
        self.status = status


class LatStats(C.Structure):
    _fields_ = [("count", C.c_uint64), ("min", C.c_int64), ("max", C.c_int64), ("sum", C.c_uint64),
                ("sum_hi", C.c_int64), ("sumsq_lo", C.c_uint64), ("sumsq_hi", C.c_uint64),
                ("dropped_below", C.c_uint64), ("dropped_above", C.c_uint64)]


_vp, _u32, _u64, _i64, _sz, _int = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int64, C.c_size_t, C.c_int
_pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes).  Kept in one table so tests can check it against include/b200crc.h.
SIGNATURES = {
    "b200_strerror": (C.c_char_p, [_int]),
    "b200_last_error": (C.c_char_p, []),
    "b200_device_count": (_int, [C.POINTER(_int)]),
    "b200_device_pci_bus_id": (_int, [_int, C.c_char_p, _int]),
    "b200_init": (_int, [_int, C.POINTER(_int), _pp]),
    "b200_shutdown": (_int, [_vp]),
    "b200_ctx_ndev": (_int, [_vp]),
    "b200_timer_start": (_int, [_vp, _int]),
    "b200_timer_stop": (_int, [_vp, _int, C.POINTER(C.c_float)]),
    "b200_alloc_pinned": (_int, [_vp, _sz, _pp]),
    "b200_free_pinned": (_int, [_vp, _vp]),
    "b200_alloc_device": (_int, [_vp, _int, _sz, _pp]),
    "b200_free_device": (_int, [_vp, _int, _vp]),
    "b200_memcpy_h2d": (_int, [_vp, _int, _vp, _vp, _sz]),
    "b200_memcpy_d2h": (_int, [_vp, _int, _vp, _vp, _sz]),
    "b200_crc32c_update_host": (_int, [_vp, _int, _u32, _vp, _sz, C.POINTER(_u32)]),
    "b200_crc32c_update_device": (_int, [_vp, _int, _u32, _vp, _sz, C.POINTER(_u32)]),
    "b200_crc32c_update_device_noalloc": (_int, [_int, _u32, _vp, _sz, C.POINTER(_u32)]),
    "b200_hasher_new": (_int, [_vp, _int, _pp]),
    "b200_hasher_write": (_int, [_vp, _vp, _sz]),
    "b200_hasher_reserve": (_int, [_vp, _pp, C.POINTER(_sz)]),
    "b200_hasher_commit": (_int, [_vp, _sz]),
    "b200_hasher_sum32": (_int, [_vp, C.POINTER(_u32)]),
    "b200_hasher_reset": (_int, [_vp]),
    "b200_hasher_free": (_int, [_vp]),
    "b200_crc32c_batch_device": (_int, [_vp, _int, _vp, _vp, _vp, _sz, _vp, _vp, _vp, C.POINTER(_u64)]),
    "b200_batch_create": (_int, [_vp, _int, _vp, _vp, _sz, _vp, _pp]),
    "b200_batch_run": (_int, [_vp, _vp]),
    "b200_batch_run_async": (_int, [_vp, _vp]),
    "b200_batch_wait": (_int, [_vp]),
    "b200_batch_results": (_int, [_vp, _vp, _vp, C.POINTER(_u64)]),
    "b200_batch_kernel_ms": (_int, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b200_batch_launches": (_int, [_vp, C.POINTER(_int)]),
    "b200_batch_free": (_int, [_vp]),
    "b200_crc32c_batch_host": (_int, [_vp, _int, _vp, _vp, _sz, _vp, _vp, _vp, C.POINTER(_u64)]),
    "b200_crc32c_sharded": (_int, [_vp, _vp, _vp, _int, C.POINTER(_u32)]),
    "b200_crc32c_stream_host": (_int, [_vp, _int, _u32, _vp, _sz, _sz, C.POINTER(_u32)]),
    "b200_crc32c_combine": (_u32, [_u32, _u32, _u64]),
    "b200_fill_payload_device": (_int, [_vp, _int, _vp, _u64, _sz, _u64]),
    "b200_fill_latency_ns_device": (_int, [_vp, _int, _vp, _u64, _sz, _u64]),
    "b200_hist_fixed": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _int, _vp, C.POINTER(LatStats)]),
    "b200_hist_fixed_pair": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _int, _vp, _vp, _int, _int, _vp, _vp, _vp]),
    "b200_select_kth": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _vp]),
    "b200_select_kth_stats": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _vp, _vp]),
    "b200_select_kth_avg": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _vp, _vp]),
    "b200_select_sample": (_int, [_vp, _int, _vp, _sz, _sz, _pp, C.POINTER(_sz)]),
    "b200_select_sample_bounds": (_int, [_vp, _int, _vp, _sz, _sz, _vp, _vp, _int, _u64, _vp, C.POINTER(_sz)]),
    "b200_bracket_pass": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _u64, _vp, _pp, C.POINTER(_u64), _vp]),
    "b200_bracket_pass_avg": (_int, [_vp, _int, _vp, _sz, _i64, _vp, _int, _u64, _vp, _pp, C.POINTER(_u64), _vp]),
    "b200_select_kth_range": (_int, [_vp, _int, _vp, _sz, _i64, _i64, _i64, _vp, _int, _vp]),
    "b200_hist_fixed_multi": (_int, [_vp, _vp, _vp, _vp, _int, _i64, _vp, _int, _int, _vp, C.POINTER(LatStats)]),
    "b200_select_kth_multi": (_int, [_vp, _vp, _vp, _vp, _int, _i64, _vp, _int, _vp]),
    "b200_parse_latency_text": (_int, [_vp, _int, _vp, _sz, _vp, _vp, _sz, C.POINTER(_u64), C.POINTER(_u64),
                                       C.POINTER(_u64), C.POINTER(_u64)]),
    "b200_hist_fixed_f64": (_int, [_vp, _int, _vp, _sz, _vp, _int, _int, _vp, C.POINTER(_u64), C.POINTER(_u64),
                                   C.POINTER(_u64), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "b200_minmax": (_int, [_vp, _int, _vp, _sz, C.POINTER(_i64), C.POINTER(_i64)]),
    "b200_radix_count": (_int, [_vp, _int, _vp, _sz, _u64, _int, _int, _int, _vp, _vp]),
    "b200_radix_count_based": (_int, [_vp, _int, _vp, _sz, _vp, _int, _int, _int, _vp, _vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load the CUDA library.  Raises if it has not been built (python __graft_entry__.py builds it)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise FileNotFoundError(
                f"{SO_PATH} is missing: build it with `python custom-go-client-benchmark_b200/build.py` "
                "(there is no CPU fallback)")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(status: int) -> None:
    if status != OK:
        L = load()
        raise B200Error(status, L.b200_strerror(status).decode(), L.b200_last_error().decode())
