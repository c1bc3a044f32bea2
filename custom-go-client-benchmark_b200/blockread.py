"""The reference's local-SSD read benchmark (benchmark-script/ssd_test/main.go) restated over this library.

`ssd_test` opens `<dir>/Workload.<i>/0` with O_DIRECT for each of N threads (main.go:40-63), reads it block by
block in sequential or shuffled order (main.go:66-89, 118-129), appends every read's duration in microseconds to
one slice (main.go:79) and prints average / P20 / P50 / P90 / p99 / min / max of the sorted slice
(main.go:146-163).  Here:

  * each thread reads into its own pinned host buffer (page aligned, which O_DIRECT needs, and directly
    visible to the copy engine), so a block can go from the page-cache-bypassing read straight to the GPU;
  * durations are kept as int64 nanoseconds and reduced on the GPU by `latency.percentiles` (exact order
    statistics, no sort), with the reference's integer-microsecond semantics (`Duration.Microseconds()`);
  * optionally (`want_crc32c`) every block read is CRC-32C-verified on the GPU - the reference does not
    checksum local reads; this is the "feed" for the verify path from a block device.

GPUDirect Storage would skip the host buffer; libcufile does not work on the boxes this was built on (see
DESIGN.md), so the blocks travel host -> device through pinned memory.
"""
from __future__ import annotations

import os
import random
import time
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import crc32c, latency

ONE_KB = 1024  # main.go:33


class BlockReadError(RuntimeError):
    """An error the reference returns from runReadFileOperations (same message text)."""


@dataclass
class BlockReadResult:
    read_time_ns: np.ndarray                 # one sample per ReadAt, all threads (main.go:37 `readTime`, here ns)
    bytes_read: int
    percentiles: latency.Percentiles         # in microseconds, as ssd_test computes them
    bad_blocks: List[tuple] = field(default_factory=list)   # (thread, block index) whose CRC-32C differed

    def report(self) -> str:
        """The lines main.go:143,157-163 print."""
        p = self.percentiles
        ms = lambda us: 0.001 * float(us)  # MicroSecondsToMilliSecond, main.go:176-178
        return "\n".join(["read benchmark completed successfully!",
                          "Average: %.3f ms" % ms(p.avg), "P20: %.3f ms" % ms(p.p20), "P50: %.3f ms" % ms(p.p50),
                          "P90: %.3f ms" % ms(p.p90), "p99: %.3f ms" % ms(p.p99), "Min: %.3f ms" % ms(p.min),
                          "Max: %.3f ms" % ms(p.max)])


def access_pattern(total_io: int, read_type: str = "seq", rng: Optional[random.Random] = None) -> List[int]:
    """main.go:118-129: block indices 0..total-1, and for any read type but "seq" the reference's shuffle
    (for i: j = rand.Intn(i+1); swap) - one pattern shared by all threads."""
    offset = list(range(total_io))
    if read_type != "seq":
        rng = rng or random.Random()
        for i in range(total_io):
            j = rng.randrange(i + 1)
            offset[i], offset[j] = offset[j], offset[i]
    return offset


def _open_file(directory: str, index: int, file_size_kb: int, direct: bool) -> int:
    """main.go:40-63."""
    name = os.path.join(directory, "Workload." + str(index), "0")
    flags = os.O_RDONLY | (os.O_DIRECT if direct else 0)
    try:
        fd = os.open(name, flags, 0o600)
    except OSError as e:
        raise BlockReadError(f"while opening file: while opening file: {e}") from e
    if os.fstat(fd).st_size != file_size_kb * ONE_KB:
        os.close(fd)
        raise BlockReadError("while opening file: file present is not equal to given file-size")
    return fd


def RunReadFileOperations(directory: str, threads: int = 1, block_size_kb: int = 1024, file_size_kb: int = 5242880,
                          read_type: str = "seq", read_count: int = 1, ctx: Optional[crc32c.Context] = None,
                          direct: bool = True, want_crc32c: Optional[Sequence[Sequence[int]]] = None,
                          rng: Optional[random.Random] = None) -> BlockReadResult:
    """runReadFileOperations (main.go:91-173) with the flag defaults of main.go:19-35.

    want_crc32c[t][b], when given, is the expected CRC-32C of block b of thread t's file: every block read is then
    verified on the GPU from the pinned buffer it was read into (one `VerifyObjects` call per read, outside the
    timed region of that read).  `direct=False` drops O_DIRECT for file systems that refuse it (tmpfs)."""
    if directory == "":
        raise BlockReadError("you must set --dir flag")
    if threads <= 0:
        raise BlockReadError("threads count not valid")
    ctx = ctx or crc32c.default_context()
    fds = [-1] * threads
    bufs: List[np.ndarray] = []
    try:
        for i in range(threads):
            fds[i] = _open_file(directory, i, file_size_kb, direct)
        total_io = file_size_kb // block_size_kb
        if file_size_kb % block_size_kb != 0:
            raise BlockReadError("block-size should be multiple of file-size")  # the reference's wording, main.go:115
        offset = access_pattern(total_io, read_type, rng)
        block = block_size_kb * ONE_KB
        bufs = [ctx.alloc_pinned(block) for _ in range(threads)]

        def worker(t: int):
            """readAlreadyOpenedFile, main.go:66-89."""
            fd, buf = fds[t], bufs[t]
            view = memoryview(buf)
            samples = np.zeros(read_count * len(offset), dtype=np.int64)
            bad, k = [], 0
            for _ in range(read_count):
                for blk in offset:
                    start = time.perf_counter_ns()
                    n = os.preadv(fd, [view], blk * block)         # fileHandle.ReadAt(b, st)
                    samples[k] = time.perf_counter_ns() - start    # time.Since(readStart)
                    k += 1
                    if n != block:
                        raise BlockReadError("while reading file: error while reading")
                    if want_crc32c is not None:
                        b, _ = crc32c.VerifyObjects([buf], [int(want_crc32c[t][blk])], ctx)
                        if b:
                            bad.append((t, blk))
            return samples, bad

        with ThreadPoolExecutor(max_workers=threads) as ex:  # errgroup, main.go:131-141
            parts = list(ex.map(worker, range(threads)))
    finally:
        for fd in fds:
            if fd >= 0:
                os.close(fd)
        for b in bufs:
            ctx.free_pinned(b)
    read_time = np.concatenate([p[0] for p in parts])
    bad_blocks = [x for p in parts for x in p[1]]
    if read_time.size == 0:  # main.go:157 divides by len(readTime): Go panics
        raise ZeroDivisionError("no reads were made (file smaller than one block): the reference panics at "
                                "ssd_test/main.go:157")
    d = ctx.alloc_device(read_time.nbytes)
    try:
        ctx.h2d(d, read_time)
        pct = latency.percentiles(d, read_time.size, latency.NS_PER_US, ctx=ctx)  # main.go:146-163
    finally:
        ctx.free_device(d)
    return BlockReadResult(read_time, int(read_time.size) * block, pct, bad_blocks)
