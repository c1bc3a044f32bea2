"""GPU test of b200_crc32c_update_device_noalloc, the entry point for callers that cannot allocate device
memory (all of it belongs to their own pool, or they cannot afford cudaMalloc's device-wide sync).

It is called through the C ABI and must agree bit for bit — CRCs are integers, the tolerance is zero — with
the ordinary b200_crc32c_update_device on the same device buffers, on two calls in a row: the second call
reuses the library's fixed scratch array with the first call's leftovers still in it.  This file sorts last
on purpose so that a run that stops at the first failure has already run every other test.
"""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c = pkg.crc32c


@pytest.fixture(scope="module")
def buf():
    ctx = crc32c.Context([0])  # the CALLER owns and allocates the input; the entry point itself allocates nothing
    n = (48 << 20) + 77
    d = ctx.alloc_device(n)
    ctx.fill_payload(d, (n + 7) // 8 - 1, 0xC0FFEE)
    tail = np.arange(16, dtype=np.uint8)
    ctx.h2d(d + n - 16, tail)
    yield ctx, d, n
    ctx.free_device(d)
    ctx.close()


def noalloc(crc, dptr, n, dev=0):
    out = C.c_uint32(0xDEADBEEF)
    rc = pkg._lib.load().b200_crc32c_update_device_noalloc(dev, crc, dptr, n, C.byref(out))
    return rc, out.value


def test_two_calls_in_a_row_match_the_allocating_entry_point(buf):
    ctx, d, n = buf
    # first call: large, many segments -> fills much of the scratch
    rc, got1 = noalloc(0, d, n)
    assert rc == 0
    assert got1 == crc32c.UpdateDevice(0, d, n, ctx)
    # second call straight after: different length, unaligned start, non-zero running CRC; stale scratch
    # from the first call must not leak into it
    rc, got2 = noalloc(0x1234ABCD, d + 5, 3_000_001)
    assert rc == 0
    assert got2 == crc32c.UpdateDevice(0x1234ABCD, d + 5, 3_000_001, ctx)
    # and the first input again gives the first answer again
    rc, got3 = noalloc(0, d, n)
    assert rc == 0 and got3 == got1


@pytest.mark.parametrize("start,length", [(0, 0), (1, 1), (3, 15), (0, 512), (7, 513), (16, 8192), (9, 1 << 20),
                                          (0, 48 << 20)])
def test_edge_sizes_match(buf, start, length):
    ctx, d, n = buf
    for crc in (0, 0xFFFFFFFF):
        rc, got = noalloc(crc, d + start, length)
        assert rc == 0
        assert got == crc32c.UpdateDevice(crc, d + start, length, ctx)


def test_python_wrapper(buf):
    ctx, d, n = buf
    assert crc32c.UpdateDeviceNoAlloc(0, d + 2, 100_000) == crc32c.UpdateDevice(0, d + 2, 100_000, ctx)


def test_sizes_beyond_the_fixed_scratch_are_rejected(buf):
    ctx, d, n = buf
    rc, got = noalloc(0, d, pkg._lib.NOALLOC_MAX_BYTES + 1)
    assert rc == pkg._lib.ERR_TOO_LARGE
    assert got == 0xDEADBEEF  # output untouched
    with pytest.raises(pkg.B200Error) as e:
        crc32c.UpdateDeviceNoAlloc(0, d, pkg._lib.NOALLOC_MAX_BYTES + 1)
    assert e.value.status == pkg._lib.ERR_TOO_LARGE
    rc, _ = noalloc(0, None, 10)
    assert rc == pkg._lib.ERR_INVALID
