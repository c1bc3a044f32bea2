"""GPU test of b200_crc32c_update_device_noalloc, the entry point for callers that cannot allocate device
memory (all of it belongs to their own pool, or they cannot afford cudaMalloc's device-wide sync).

It is called through the C ABI and must agree bit for bit — CRCs are integers, the tolerance is zero — with
the ordinary b200_crc32c_update_device on the same device buffers AND with the CPU oracle on the same bytes, on
two calls in a row: the second call reuses the library's fixed scratch array with the first call's leftovers
still in it.  One test runs it in a process that never creates a b200_ctx at all, which is how the callers it is
meant for use it.  This file sorts last on purpose so that a run that stops at the first failure has already
run every other test.
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c = pkg.crc32c


@pytest.fixture(scope="module")
def buf():
    ctx = crc32c.Context([0])  # the CALLER owns and allocates the input; the entry point itself allocates nothing
    n = (48 << 20) + 77
    d = ctx.alloc_device(n)
    host = oracle.payload(n, seed=0xC0FFEE).copy()
    host[n - 16:] = np.arange(16, dtype=np.uint8)
    ctx.h2d(d, host)
    global HOST
    HOST = host
    yield ctx, d, n
    ctx.free_device(d)
    ctx.close()


HOST = None


def cpu(crc, start, length):
    return oracle.crc32c(HOST[start:start + length], crc)


def noalloc(crc, dptr, n, dev=0):
    out = C.c_uint32(0xDEADBEEF)
    rc = pkg._lib.load().b200_crc32c_update_device_noalloc(dev, crc, dptr, n, C.byref(out))
    return rc, out.value


def test_two_calls_in_a_row_match_the_allocating_entry_point(buf):
    ctx, d, n = buf
    # first call: large, many segments -> fills much of the scratch
    rc, got1 = noalloc(0, d, n)
    assert rc == 0
    assert got1 == crc32c.UpdateDevice(0, d, n, ctx) == cpu(0, 0, n)
    # second call straight after: different length, unaligned start, non-zero running CRC; stale scratch
    # from the first call must not leak into it
    rc, got2 = noalloc(0x1234ABCD, d + 5, 3_000_001)
    assert rc == 0
    assert got2 == crc32c.UpdateDevice(0x1234ABCD, d + 5, 3_000_001, ctx) == cpu(0x1234ABCD, 5, 3_000_001)
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
        assert got == crc32c.UpdateDevice(crc, d + start, length, ctx) == cpu(crc, start, length)


NEVER_INIT = r"""
import ctypes as C, os, sys
import numpy as np
import torch                                   # only to own the device buffer: the library allocates nothing
sys.path.insert(0, sys.argv[1])
import oracle
lib = C.CDLL(os.path.join(sys.argv[1], "custom-go-client-benchmark_b200", "libb200crc.so"))   # no b200_init anywhere
f = lib.b200_crc32c_update_device_noalloc
f.restype, f.argtypes = C.c_int, [C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32)]
n = (20 << 20) + 13
host = oracle.payload(n, seed=99)
x = torch.from_numpy(host.copy()).cuda()
torch.cuda.synchronize()
for crc, start, length in ((0, 0, n), (0xCAFEF00D, 3, 1_000_003), (0, 0, n), (0, 7, 0), (5, 1, 1)):
    out = C.c_uint32(0xDEADBEEF)
    rc = f(0, crc, x.data_ptr() + start, length, C.byref(out))
    assert rc == 0, rc
    assert out.value == oracle.crc32c(host[start:start + length], crc), (crc, start, length, hex(out.value))
assert torch.cuda.current_device() == 0
print("never-init ok")
"""


def test_in_a_process_that_never_creates_a_context():
    """The callers this entry point is for have no b200_ctx: the library is loaded and this one function is
    called, nothing else.  The device buffer belongs to the caller (torch here), the checker is the CPU oracle."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", NEVER_INIT, root], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "never-init ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


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
