"""Host-side mirror of the Go surface the reference's CRC path uses, over the C ABI.

Names follow Go's hash/crc32 as used by cloud.google.com/go/storage's Reader (the arithmetic behind
main.go:135-148 of the reference):  New() -> hash.Hash32 {Write, Sum, Sum32, Reset, Size, BlockSize},
Update(crc, p), Checksum(p), Castagnoli, plus the batched VerifyObjects and Combine the north star adds.
The same surface is provided for Go itself in go/crc32c (cgo).  Every byte is hashed on the GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import B200Error, check

Castagnoli = 0x82F63B78  # crc32.Castagnoli
Size = 4                 # crc32.Size


class BadCRCError(Exception):
    """storage.Reader's "storage: bad CRC on read: got %d, want %d" (surfaces at main.go:141-143)."""

    def __init__(self, index: int, got: int, want: int):
        super().__init__(f"storage: bad CRC on read: got {got}, want {want} (object {index})")
        self.index, self.got, self.want = index, got, want


class Context:
    """Owns the library context (streams, tables, scratch) for a set of devices."""

    def __init__(self, devices: Optional[Sequence[int]] = None):
        L = _lib.load()
        if devices is None:
            devices = [0]
        arr = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        check(L.b200_init(len(devices), arr, C.byref(h)))
        self._h, self._L, self.devices = h, L, list(devices)

    def close(self):
        if getattr(self, "_h", None):
            self._L.b200_shutdown(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- memory -------------------------------------------------------------------------------------
    def alloc_device(self, nbytes: int, dev: Optional[int] = None) -> int:
        p = C.c_void_p()
        check(self._L.b200_alloc_device(self._h, self._dev(dev), nbytes, C.byref(p)))
        return p.value

    def free_device(self, ptr: int, dev: Optional[int] = None):
        check(self._L.b200_free_device(self._h, self._dev(dev), ptr))

    def alloc_pinned(self, nbytes: int) -> np.ndarray:
        p = C.c_void_p()
        check(self._L.b200_alloc_pinned(self._h, nbytes, C.byref(p)))
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        a = np.frombuffer(buf, dtype=np.uint8)
        self._pinned = getattr(self, "_pinned", {})
        self._pinned[a.ctypes.data] = p.value
        return a

    def free_pinned(self, a: np.ndarray):
        p = self._pinned.pop(a.ctypes.data)
        check(self._L.b200_free_pinned(self._h, p))

    def h2d(self, dptr: int, host: np.ndarray, dev: Optional[int] = None):
        host = np.ascontiguousarray(host)
        check(self._L.b200_memcpy_h2d(self._h, self._dev(dev), dptr, host.ctypes.data, host.nbytes))

    def d2h(self, host: np.ndarray, dptr: int, dev: Optional[int] = None):
        check(self._L.b200_memcpy_d2h(self._h, self._dev(dev), host.ctypes.data, dptr, host.nbytes))

    def fill_payload(self, dptr: int, nwords: int, seed: int, first_word: int = 0, dev: Optional[int] = None):
        check(self._L.b200_fill_payload_device(self._h, self._dev(dev), dptr, first_word, nwords, seed))

    def fill_latency_ns(self, dptr: int, n: int, seed: int, first: int = 0, dev: Optional[int] = None):
        check(self._L.b200_fill_latency_ns_device(self._h, self._dev(dev), dptr, first, n, seed))

    def timer_start(self, dev: Optional[int] = None):
        check(self._L.b200_timer_start(self._h, self._dev(dev)))

    def timer_stop(self, dev: Optional[int] = None) -> float:
        ms = C.c_float()
        check(self._L.b200_timer_stop(self._h, self._dev(dev), C.byref(ms)))
        return ms.value

    def _dev(self, dev):
        return self.devices[0] if dev is None else dev


_default: Optional[Context] = None


def default_context() -> Context:
    global _default
    if _default is None:
        _default = Context([0])
    return _default


def _as_bytes(p) -> np.ndarray:
    if isinstance(p, np.ndarray):
        if p.dtype == np.uint8 and p.ndim == 1 and p.flags.c_contiguous:
            return p
        return np.ascontiguousarray(p).view(np.uint8).reshape(-1)
    return np.frombuffer(p, dtype=np.uint8)


def _addr(a: np.ndarray):
    return a.__array_interface__["data"][0] if a.size else None


def Update(crc: int, p, ctx: Optional[Context] = None, dev: Optional[int] = None) -> int:
    """crc32.Update(crc, castagnoliTable, p): Update(Update(0, a), b) == Update(0, a + b)."""
    ctx = ctx or default_context()
    a = _as_bytes(p)
    out = C.c_uint32()
    check(ctx._L.b200_crc32c_update_host(ctx._h, ctx._dev(dev), crc & 0xFFFFFFFF, a.ctypes.data if a.size else None,
                                         a.size, C.byref(out)))
    return out.value


def Checksum(p, ctx: Optional[Context] = None) -> int:
    """crc32.Checksum(p, castagnoliTable)."""
    return Update(0, p, ctx)


def UpdateDevice(crc: int, dptr: int, n: int, ctx: Optional[Context] = None, dev: Optional[int] = None) -> int:
    ctx = ctx or default_context()
    out = C.c_uint32()
    check(ctx._L.b200_crc32c_update_device(ctx._h, ctx._dev(dev), crc & 0xFFFFFFFF, dptr, n, C.byref(out)))
    return out.value


def UpdateDeviceNoAlloc(crc: int, dptr: int, n: int, dev: int = 0) -> int:
    """crc32.Update on a device buffer for callers that must not allocate device memory (their pool owns it
    all, or they cannot take cudaMalloc's device-wide synchronisation).  Uses the library's fixed built-in
    scratch: at most 32 GiB per call, calls are serialised, and the first call on a GPU does its set-up."""
    out = C.c_uint32()
    check(_lib.load().b200_crc32c_update_device_noalloc(dev, crc & 0xFFFFFFFF, dptr, n, C.byref(out)))
    return out.value


def Combine(crc1: int, crc2: int, len2: int) -> int:
    """CRC of A||B from crc(A), crc(B), len(B).  Pure GF(2) arithmetic on the host."""
    return int(_lib.load().b200_crc32c_combine(crc1 & 0xFFFFFFFF, crc2 & 0xFFFFFFFF, len2))


class Hash32:
    """hash.Hash32 as returned by crc32.New(crc32.MakeTable(crc32.Castagnoli))."""

    def __init__(self, ctx: Optional[Context] = None, dev: Optional[int] = None):
        self._ctx = ctx or default_context()
        self._h = C.c_void_p()
        check(self._ctx._L.b200_hasher_new(self._ctx._h, self._ctx._dev(dev), C.byref(self._h)))

    def Write(self, p) -> int:
        a = _as_bytes(p)
        check(self._ctx._L.b200_hasher_write(self._h, a.ctypes.data if a.size else None, a.size))
        return a.size  # Go: n == len(p), err == nil always

    def ReadFrom(self, r) -> int:
        """io.ReaderFrom: drain `r` (anything with readinto, or read) straight into the hasher's pinned
        staging buffer — the replacement for io.CopyBuffer(io.Discard, rc, buf) at main.go:140.  Returns the
        number of bytes consumed (Go returns (n int64, err error); errors raise here)."""
        L, total = self._ctx._L, 0
        while True:
            ptr, avail = C.c_void_p(), C.c_size_t()
            check(L.b200_hasher_reserve(self._h, C.byref(ptr), C.byref(avail)))
            view = memoryview((C.c_uint8 * avail.value).from_address(ptr.value)).cast("B")
            if hasattr(r, "readinto"):
                n = r.readinto(view)
            else:
                chunk = r.read(avail.value)
                n = len(chunk)
                view[:n] = chunk
            if not n:
                return total
            check(L.b200_hasher_commit(self._h, n))
            total += n

    def Sum32(self) -> int:
        out = C.c_uint32()
        check(self._ctx._L.b200_hasher_sum32(self._h, C.byref(out)))
        return out.value

    def Sum(self, b: bytes = b"") -> bytes:
        """Appends the big-endian checksum to b; does not change the state."""
        return bytes(b) + self.Sum32().to_bytes(4, "big")

    def Reset(self):
        check(self._ctx._L.b200_hasher_reset(self._h))

    def Size(self) -> int:
        return 4

    def BlockSize(self) -> int:
        return 1

    def close(self):
        if self._h:
            self._ctx._L.b200_hasher_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def New(ctx: Optional[Context] = None, dev: Optional[int] = None) -> Hash32:
    return Hash32(ctx, dev)


def VerifyObjects(bufs, want: Optional[Sequence[int]], ctx: Optional[Context] = None,
                  dev: Optional[int] = None):
    """VerifyObjects([]buf, []want) -> (bad []int, got []uint32).  Host buffers; flags, never aborts.

    bufs is a sequence of bytes-like objects, or a 2-D uint8 numpy array whose rows are the objects (the shape
    of the read loop's per-worker buffers, main.go:125,200) — that form needs no per-object Python work."""
    ctx = ctx or default_context()
    if isinstance(bufs, np.ndarray) and bufs.ndim == 2 and bufs.dtype == np.uint8 and bufs.strides[1] == 1:
        n = bufs.shape[0]
        base = bufs.__array_interface__["data"][0]
        ptrs = (base + np.arange(n, dtype=np.uint64) * np.uint64(bufs.strides[0])).astype(np.uint64)
        lens = np.full(n, bufs.shape[1], dtype=np.uint64)
        ptr_arg, keep = ptrs.ctypes.data, (bufs, ptrs)
    else:
        arrs = [_as_bytes(b) for b in bufs]
        n = len(arrs)
        ptr_arr = (C.c_void_p * max(n, 1))(*[_addr(a) for a in arrs])
        lens = np.array([a.size for a in arrs], dtype=np.uint64)
        ptr_arg, keep = ptr_arr, arrs
    w = None if want is None else np.ascontiguousarray(want, dtype=np.uint32)
    got = np.zeros(n, dtype=np.uint32)
    ok = np.zeros(n, dtype=np.uint8)
    nbad = C.c_uint64()
    check(ctx._L.b200_crc32c_batch_host(ctx._h, ctx._dev(dev), ptr_arg, lens.ctypes.data, n,
                                        None if w is None else w.ctypes.data, got.ctypes.data, ok.ctypes.data,
                                        C.byref(nbad)))
    del keep
    bad = [int(i) for i in np.nonzero(ok == 0)[0]]
    assert len(bad) == nbad.value
    return bad, got


def VerifyObjectsAt(ptrs, lens, want, ctx: Optional[Context] = None, dev: Optional[int] = None):
    """VerifyObjects on raw host addresses: ptrs[i] is the address of lens[i] bytes (what the cgo binding hands
    over for a [][]byte).  Returns (bad []int, got []uint32).  The caller keeps the buffers alive."""
    ctx = ctx or default_context()
    ptrs = np.ascontiguousarray(ptrs, dtype=np.uint64)
    lens = np.ascontiguousarray(lens, dtype=np.uint64)
    n = ptrs.size
    w = None if want is None else np.ascontiguousarray(want, dtype=np.uint32)
    got = np.zeros(n, dtype=np.uint32)
    ok = np.zeros(n, dtype=np.uint8)
    nbad = C.c_uint64()
    check(ctx._L.b200_crc32c_batch_host(ctx._h, ctx._dev(dev), ptrs.ctypes.data, lens.ctypes.data, n,
                                        None if w is None else w.ctypes.data, got.ctypes.data, ok.ctypes.data,
                                        C.byref(nbad)))
    return [int(i) for i in np.nonzero(ok == 0)[0]], got


class DeviceBatch:
    """A prepared VerifyObjects plan over a device-resident arena (BASELINE config 2)."""

    def __init__(self, off, length, want=None, ctx: Optional[Context] = None, dev: Optional[int] = None):
        self._ctx = ctx or default_context()
        self.off = np.ascontiguousarray(off, dtype=np.uint64)
        self.len = np.ascontiguousarray(length, dtype=np.uint64)
        self.n = self.off.size
        w = None if want is None else np.ascontiguousarray(want, dtype=np.uint32)
        self._h = C.c_void_p()
        check(self._ctx._L.b200_batch_create(self._ctx._h, self._ctx._dev(dev), self.off.ctypes.data,
                                             self.len.ctypes.data, self.n, None if w is None else w.ctypes.data,
                                             C.byref(self._h)))

    def run(self, base: int):
        check(self._ctx._L.b200_batch_run(self._h, base))

    def run_async(self, base: int):
        check(self._ctx._L.b200_batch_run_async(self._h, base))

    def wait(self):
        check(self._ctx._L.b200_batch_wait(self._h))

    def results(self):
        got = np.zeros(self.n, dtype=np.uint32)
        ok = np.zeros(self.n, dtype=np.uint8)
        nbad = C.c_uint64()
        check(self._ctx._L.b200_batch_results(self._h, got.ctypes.data, ok.ctypes.data, C.byref(nbad)))
        return got, ok.astype(bool), int(nbad.value)

    def kernel_ms(self):
        t, r = C.c_float(), C.c_float()
        check(self._ctx._L.b200_batch_kernel_ms(self._h, C.byref(t), C.byref(r)))
        return t.value, r.value

    def launches(self) -> int:
        n = C.c_int()
        check(self._ctx._L.b200_batch_launches(self._h, C.byref(n)))
        return n.value

    def close(self):
        if self._h:
            self._ctx._L.b200_batch_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def VerifyObjectsDevice(base: int, off, length, want=None, ctx: Optional[Context] = None, dev: Optional[int] = None):
    b = DeviceBatch(off, length, want, ctx, dev)
    try:
        b.run(base)
        return b.results()
    finally:
        b.close()


def ShardedChecksum(dptrs: Sequence[int], lens: Sequence[int], ctx: Context) -> int:
    """One object sharded over ctx.devices in message order (BASELINE config 3)."""
    n = len(dptrs)
    p = (C.c_void_p * n)(*dptrs)
    l = (C.c_uint64 * n)(*lens)
    out = C.c_uint32()
    check(ctx._L.b200_crc32c_sharded(ctx._h, p, l, n, C.byref(out)))
    return out.value


def StreamChecksum(host: np.ndarray, chunk: int = 256 << 20, crc: int = 0, ctx: Optional[Context] = None,
                   dev: Optional[int] = None) -> int:
    """Host-resident streaming mode (BASELINE config 4): chunked H2D overlapped with the CRC kernels."""
    ctx = ctx or default_context()
    a = _as_bytes(host)
    out = C.c_uint32()
    check(ctx._L.b200_crc32c_stream_host(ctx._h, ctx._dev(dev), crc, a.ctypes.data if a.size else None, a.size, chunk,
                                         C.byref(out)))
    return out.value


# ---- wire checksums: how GCS hands the expected CRC-32C to storage.Reader ----------------------

def EncodeXGoogHash(crc: int) -> str:
    """'crc32c=<base64 of the big-endian uint32>' as in the X-Goog-Hash response header (HTTP/JSON API)."""
    import base64
    return "crc32c=" + base64.b64encode((crc & 0xFFFFFFFF).to_bytes(4, "big")).decode()


def DecodeXGoogHash(header: str) -> Optional[int]:
    """Extract the CRC-32C from an X-Goog-Hash header value ('crc32c=AAAAAA==,md5=...'); None if absent."""
    import base64
    for part in header.split(","):
        part = part.strip()
        if part.startswith("crc32c="):
            raw = base64.b64decode(part[len("crc32c="):])
            if len(raw) != 4:
                raise ValueError("crc32c value must decode to 4 bytes")
            return int.from_bytes(raw, "big")
    return None


def DecodeFixed32(raw: bytes) -> int:
    """gRPC ObjectChecksums.crc32c is a protobuf fixed32: four little-endian bytes on the wire."""
    if len(raw) != 4:
        raise ValueError("fixed32 is 4 bytes")
    return int.from_bytes(raw, "little")


def EncodeFixed32(crc: int) -> bytes:
    return (crc & 0xFFFFFFFF).to_bytes(4, "little")
