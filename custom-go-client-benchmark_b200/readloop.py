"""The reference's read loop (main.go:119-156, `ReadObject`) restated over the GPU-backed hasher.

BASELINE config 1: N workers each re-read one object M times; every read drains an `io.ReadCloser` through a
2 MiB buffer while a running CRC-32C is kept, compares it with the object's expected checksum at EOF
("storage: bad CRC on read", surfacing at main.go:141-143) and records the read's latency in whole
milliseconds (main.go:145-146).  Here the network object is replaced by an in-memory reader, the buffer by
the hasher's pinned staging (io.ReaderFrom instead of io.CopyBuffer + io.Discard) and `stats.Record` by an
int64-nanosecond sample array that the latency kernels reduce on the device.
"""
from __future__ import annotations

import io
import time
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import crc32c, latency

MB = 1024 * 1024  # main.go:34


class InMemoryObject:
    """Stands in for `bucketHandle.Object(name)`: NewReader() returns a fresh reader over the same bytes that
    serves at most `max_read` bytes per Read call (a transport returns whatever it has, <= the 2 MiB buffer)."""

    def __init__(self, data, crc32c_want: int, max_read: int = 2 * MB):
        self.data = memoryview(data).cast("B")
        self.crc32c = crc32c_want  # what GCS reports in X-Goog-Hash / ObjectChecksums
        self.max_read = max_read

    def NewReader(self):
        obj = self

        class _Reader(io.RawIOBase):
            def __init__(self):
                self.pos = 0

            def readable(self):
                return True

            def readinto(self, b):
                n = min(len(b), obj.max_read, len(obj.data) - self.pos)
                b[:n] = obj.data[self.pos:self.pos + n]
                self.pos += n
                return n

        return _Reader()


@dataclass
class ReadLoopResult:
    latencies_ns: np.ndarray          # one time.Duration per read, all workers
    bytes_read: int
    reads: int
    errors: List[str] = field(default_factory=list)


def ReadObject(worker_id: int, obj: InMemoryObject, reads: int, ctx: Optional[crc32c.Context] = None,
               clock: Callable[[], int] = time.perf_counter_ns) -> ReadLoopResult:
    """One worker: main.go:119-156.  Returns at the first bad read, like the reference (main.go:141-143)."""
    h = crc32c.New(ctx)
    lat = np.zeros(reads, dtype=np.int64)
    total = 0
    try:
        for i in range(reads):
            start = clock()                                   # main.go:133
            rc = obj.NewReader()                              # main.go:135
            h.Reset()
            total += h.ReadFrom(rc)                           # main.go:140 (io.CopyBuffer(io.Discard, rc, buf))
            got = h.Sum32()                                   # storage.Reader's EOF check
            if got != obj.crc32c:
                err = f"while reading and discarding content: storage: bad CRC on read: got {got}, want {obj.crc32c}"
                return ReadLoopResult(lat[:i], total, i, [f"worker {worker_id}: {err}"])
            lat[i] = clock() - start                          # main.go:145
            rc.close()                                        # main.go:148
    finally:
        h.close()
    return ReadLoopResult(lat, total, reads)


def RunReadLoop(objects: Sequence[InMemoryObject], reads_per_worker: int, ctx: Optional[crc32c.Context] = None,
                clock: Callable[[], int] = time.perf_counter_ns) -> ReadLoopResult:
    """main.go:200-212: one goroutine per worker (errgroup); the first error is reported, all samples kept."""
    with ThreadPoolExecutor(max_workers=max(len(objects), 1)) as ex:
        parts = list(ex.map(lambda a: ReadObject(a[0], a[1], reads_per_worker, ctx, clock), enumerate(objects)))
    lat = np.concatenate([p.latencies_ns for p in parts]) if parts else np.zeros(0, dtype=np.int64)
    return ReadLoopResult(lat, sum(p.bytes_read for p in parts), sum(p.reads for p in parts),
                          [e for p in parts for e in p.errors])


def reduce_latencies(latencies_ns: np.ndarray, ctx: Optional[crc32c.Context] = None):
    """Upload the recorded samples and reduce them the three ways the reference does: README 15-bin histogram
    (README.md:21-32), OpenCensus distribution (metrics_exporter.go:22-34) — both on float64(ms) — and the
    ssd_test percentiles on microseconds (ssd_test/main.go:147-163)."""
    ctx = ctx or crc32c.default_context()
    a = np.ascontiguousarray(latencies_ns, dtype=np.int64)
    d = ctx.alloc_device(max(a.nbytes, 8))
    try:
        if a.size:
            ctx.h2d(d, a)
        readme, oc = latency.hist_readme_and_opencensus(d, a.size, ctx=ctx)  # one pass for both
        pct = latency.percentiles(d, a.size, latency.NS_PER_US, ctx=ctx) if a.size else None
    finally:
        ctx.free_device(d)
    return readme, oc, pct
