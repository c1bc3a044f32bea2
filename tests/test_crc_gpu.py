"""GPU parity tests for the CRC-32C path, through the C ABI, against the CPU oracle (bit-exact)."""
import importlib
import json
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import oracle

pytestmark = pytest.mark.gpu
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c = pkg.crc32c

KATS = [(b"", 0), (b"a", 0xC1D04330), (b"123456789", 0xE3069283), (bytes(32), 0x8A9136AA),
        (b"\xff" * 32, 0x62A8AB43), (bytes(range(32)), 0x46DD794E), (bytes(range(31, -1, -1)), 0x113FDB5C)]


@pytest.fixture(scope="module")
def ctx():
    c = crc32c.Context([0])
    yield c
    c.close()


@pytest.fixture(scope="module")
def arena(ctx):
    """64 MiB + slack of splitmix payload on the device and on the host."""
    n = (64 << 20) + 4096
    host = oracle.payload(n)
    d = ctx.alloc_device(n)
    ctx.h2d(d, host)
    yield host, d
    ctx.free_device(d)


@pytest.mark.parametrize("data,want", KATS)
def test_known_answers_update_host(ctx, data, want):
    assert crc32c.Checksum(data, ctx) == want


def test_update_continuation_and_chunking(ctx):
    d = bytes(oracle.payload(300000))
    crc = 0
    for i in range(0, len(d), 37 * 1024 + 5):
        crc = crc32c.Update(crc, d[i:i + 37 * 1024 + 5], ctx)
    assert crc == oracle.crc32c(d)
    assert crc32c.Update(0x12345678, d[:1000], ctx) == oracle.crc32c(d[:1000], 0x12345678)


@settings(max_examples=60, deadline=None)
@given(st.integers(0, 4096 + 64), st.integers(0, 3 << 20), st.integers(0, 2**32 - 1))
def test_update_device_random_offsets_and_lengths(ctx, arena, start, length, crc):
    host, d = arena
    assert crc32c.UpdateDevice(crc, d + start, length, ctx) == oracle.crc32c(host[start:start + length], crc)


@pytest.mark.parametrize("length", [0, 1, 3, 15, 16, 17, 511, 512, 513, 1023, 1024, 8191, 8192, 8193, 65536,
                                    (1 << 20) - 1, (2 << 20), (2 << 20) + 1, 8 << 20, (64 << 20) + 999])
@pytest.mark.parametrize("start", [0, 1, 7, 15, 16, 33])
def test_update_device_edges(ctx, arena, start, length):
    host, d = arena
    assert crc32c.UpdateDevice(0, d + start, length, ctx) == oracle.crc32c(host[start:start + length])


def test_hash32_interface(ctx):
    h = crc32c.New(ctx)
    assert h.Size() == 4 and h.BlockSize() == 1
    assert h.Sum32() == 0
    d = bytes(oracle.payload(3_000_000))
    n = h.Write(d[:10])
    assert n == 10
    h.Write(d[10:70000])
    mid = h.Sum32()
    assert mid == oracle.crc32c(d[:70000])
    assert h.Sum32() == mid  # Sum32 does not disturb the state
    assert h.Sum(b"ab") == b"ab" + mid.to_bytes(4, "big")
    h.Write(d[70000:])
    assert h.Sum32() == oracle.crc32c(d)
    h.Reset()
    assert h.Sum32() == 0
    h.Write(b"123456789")
    assert h.Sum32() == 0xE3069283
    h.close()


def test_hash32_many_small_writes_and_flush_boundary(ctx):
    h = crc32c.New(ctx)
    d = bytes(oracle.payload((8 << 20) + 12345))  # crosses the 8 MiB staging buffer
    step = 2 << 20                                   # the read loop's buffer size (main.go:125)
    for i in range(0, len(d), step):
        h.Write(d[i:i + step])
    assert h.Sum32() == oracle.crc32c(d)
    h.close()


def test_hash32_read_from(ctx):
    import io
    d = bytes(oracle.payload((20 << 20) + 999))
    h = crc32c.New(ctx)
    assert h.ReadFrom(io.BytesIO(d)) == len(d)          # readinto path, crosses the staging buffer twice
    assert h.Sum32() == oracle.crc32c(d)

    class Dribble:                                       # transport that returns short reads, read() only
        def __init__(self, data):
            self.data, self.pos = data, 0

        def read(self, n):
            n = min(n, 70001)
            out = self.data[self.pos:self.pos + n]
            self.pos += len(out)
            return out

    h.Reset()
    h.Write(d[:100])
    assert h.ReadFrom(Dribble(d[100:3_000_000])) == 3_000_000 - 100
    assert h.Sum32() == oracle.crc32c(d[:3_000_000])
    h.close()


def _batch_check(ctx, host, d, off, ln):
    want, _, _ = oracle.verify_objects(host, off, ln, threads=8)
    got, ok, nbad = crc32c.VerifyObjectsDevice(d, off, ln, want, ctx=ctx)
    assert (got == want).all()
    assert nbad == 0 and ok.all()
    return want


def test_batch_uniform_2mib_objects(ctx, arena):
    host, d = arena
    n = 32
    off = np.arange(n, dtype=np.uint64) * (2 << 20)
    ln = np.full(n, 2 << 20, dtype=np.uint64)
    _batch_check(ctx, host, d, off, ln)


def test_batch_ragged_unaligned(ctx, arena):
    host, d = arena
    rng = np.random.default_rng(7)
    n = 300
    ln = rng.integers(1, 400_000, size=n).astype(np.uint64)
    off = rng.integers(0, (64 << 20) - 400_000, size=n).astype(np.uint64)  # overlapping, unaligned
    ln[:8] = [0, 1, 15, 16, 511, 512, 513, 4 << 20]
    _batch_check(ctx, host, d, off, ln)


def test_batch_many_tiny_objects(ctx, arena):
    host, d = arena
    rng = np.random.default_rng(11)
    n = 5000
    ln = rng.integers(0, 700, size=n).astype(np.uint64)
    off = rng.integers(0, 1 << 20, size=n).astype(np.uint64)
    _batch_check(ctx, host, d, off, ln)


@pytest.mark.parametrize("block,nobj,off0", [(4096, 4096, 0), (1024, 2000, 16), (512 * 3, 1333, 4096 + 48),
                                             (32768, 1100, 32), (4096, 63, 0), (4096, 1030, 8), (2048, 4097, 1 << 20),
                                             (4096, 1000, 0), (512, 3000, 0)])
def test_batch_of_equal_blocks_back_to_back(ctx, arena, block, nobj, off0):
    """Equal blocks back to back (pages, ssd_test's blocks): whole groups of 32 go to the fold kernel a lane per
    object and are finished there, the rest and every case it does not take (fewer than 1024 objects, blocks under
    1 KiB, an arena offset that is not 16-byte aligned) go to the row kernel; a corrupt block in either part is the one
    that is flagged."""
    host, d = arena
    off = np.uint64(off0) + np.arange(nobj, dtype=np.uint64) * np.uint64(block)
    ln = np.full(nobj, block, dtype=np.uint64)
    want = _batch_check(ctx, host, d, off, ln)
    for bad_at in (5, nobj - 1):
        w = want.copy()
        w[bad_at] ^= 0x80
        got, ok, nbad = crc32c.VerifyObjectsDevice(d, off, ln, w, ctx=ctx)
        assert nbad == 1 and not ok[bad_at] and ok.sum() == nobj - 1 and (got == want).all()


@pytest.mark.parametrize("longest", [700, 8192, 32768, 40000])
def test_batch_of_many_ragged_small_objects(ctx, arena, longest):
    """1024 or more objects of any offset and length, none above 32 KiB, go to the fold kernel a lane per object
    (sorted by stream length, right-aligned by zeros), heads and tails to a thread per object; with one object above
    32 KiB the row kernel takes the batch as before.  Lengths around every boundary of the cut (0, < 16, multiples of
    16 and of 256, one more, one less) are in; a corrupt object is the one that is flagged."""
    host, d = arena
    rng = np.random.default_rng(longest)
    n = 3000
    ln = rng.integers(0, longest + 1, size=n).astype(np.uint64)
    ln[:24] = [0, 1, 15, 16, 17, 31, 32, 33, 255, 256, 257, 271, 272, 511, 512, 513, 4095, 4096, 4097, 2, 3, 0, 16, longest]
    off = rng.integers(0, (64 << 20) - longest - 64, size=n).astype(np.uint64)  # overlapping, any alignment
    off[:4] = [0, 1, 15, 16]
    want = _batch_check(ctx, host, d, off, ln)
    w = want.copy()
    w[1234] ^= 1
    got, ok, nbad = crc32c.VerifyObjectsDevice(d, off, ln, w, ctx=ctx)
    assert nbad == 1 and not ok[1234] and ok.sum() == n - 1 and (got == want).all()


def test_batch_empty(ctx, arena):
    host, d = arena
    got, ok, nbad = crc32c.VerifyObjectsDevice(d, np.zeros(0, np.uint64), np.zeros(0, np.uint64), None, ctx=ctx)
    assert got.size == 0 and nbad == 0


def test_batch_misaligned_arena_is_rejected(ctx, arena):
    host, d = arena
    with pytest.raises(pkg.B200Error):
        crc32c.VerifyObjectsDevice(d + 4, np.zeros(1, np.uint64), np.full(1, 100, np.uint64), None, ctx=ctx)


def test_batch_flags_exactly_the_corrupt_object(ctx):
    n, size = 48, 1 << 20
    host = oracle.payload(n * size).copy()
    off = np.arange(n, dtype=np.uint64) * size
    ln = np.full(n, size, dtype=np.uint64)
    want, _, _ = oracle.verify_objects(host, off, ln, threads=8)
    host[17 * size + 777] ^= 0x01  # one flipped bit
    d = ctx.alloc_device(host.size)
    ctx.h2d(d, host)
    b = crc32c.DeviceBatch(off, ln, want, ctx)
    b.run(d)
    got, ok, nbad = b.results()
    assert nbad == 1 and list(np.nonzero(~ok)[0]) == [17]
    assert got[17] != want[17] and (np.delete(got, 17) == np.delete(want, 17)).all()
    # plan is reusable: heal the object and re-run
    host[17 * size + 777] ^= 0x01
    ctx.h2d(d, host)
    b.run(d)
    got, ok, nbad = b.results()
    assert nbad == 0 and (got == want).all()
    assert b.launches() == 2
    b.close()
    ctx.free_device(d)


@pytest.mark.parametrize("rps", ["16", "64", "1000"])
def test_rows_per_segment_is_not_a_semantic_knob(rps, arena):
    # fresh process-level knob: run in a subprocess so the env var is read
    import subprocess, sys
    code = (
        "import importlib, numpy as np, oracle\n"
        "pkg = importlib.import_module('custom-go-client-benchmark_b200')\n"
        "c = pkg.crc32c.Context([0])\n"
        "h = oracle.payload(5_000_000)\n"
        "d = c.alloc_device(h.size); c.h2d(d, h)\n"
        "assert pkg.crc32c.UpdateDevice(0, d + 5, 4_999_000, c) == oracle.crc32c(h[5:5 + 4_999_000])\n"
        "print('ok')\n")
    env = dict(os.environ, B200_ROWS_PER_SEG=rps)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.check_output([sys.executable, "-c", code], env=env, cwd=root, text=True)
    assert "ok" in out


@pytest.mark.parametrize("rps", ["128", "256", "512", "2048"])
def test_fold_and_row_kernels_mix_on_ragged_objects(rps):
    """Forces a segment length so that every object is part full segments (fold kernel) and part partial
    segment (row kernel), with unaligned starts; all must agree with the oracle."""
    import subprocess, sys
    code = (
        "import importlib, numpy as np, oracle\n"
        "pkg = importlib.import_module('custom-go-client-benchmark_b200')\n"
        "c = pkg.crc32c.Context([0])\n"
        "h = oracle.payload((96 << 20) + 64)\n"
        "d = c.alloc_device(h.size); c.h2d(d, h)\n"
        "rng = np.random.default_rng(5)\n"
        "n = 120\n"
        "ln = rng.integers(50_000, 6_000_000, size=n).astype(np.uint64)\n"
        "ln[:6] = [65536, 65536 * 2 + 16, 1 << 20, (1 << 20) + 511, (2 << 20) + 513, 4 << 20]\n"
        "off = rng.integers(0, (96 << 20) - 6_000_000, size=n).astype(np.uint64)\n"
        "want, _, _ = oracle.verify_objects(h, off, ln, threads=8)\n"
        "got, ok, nbad = pkg.crc32c.VerifyObjectsDevice(d, off, ln, want, ctx=c)\n"
        "assert nbad == 0 and (got == want).all(), np.nonzero(got != want)\n"
        "for s, l in [(0, 96 << 20), (3, (96 << 20) - 7), (17, 5_000_000)]:\n"
        "    assert pkg.crc32c.UpdateDevice(0x55AA55AA, d + s, l, c) == oracle.crc32c(h[s:s + l], 0x55AA55AA)\n"
        "print('ok')\n")
    env = dict(os.environ, B200_ROWS_PER_SEG=rps)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.check_output([sys.executable, "-c", code], env=env, cwd=root, text=True)
    assert "ok" in out


def test_row_kernel_alone_matches(arena):
    """B200_NO_FOLD=1 keeps everything on the row kernel (the A/B switch used in DESIGN.md)."""
    import subprocess, sys
    code = (
        "import importlib, numpy as np, oracle\n"
        "pkg = importlib.import_module('custom-go-client-benchmark_b200')\n"
        "c = pkg.crc32c.Context([0])\n"
        "h = oracle.payload(40_000_001)\n"
        "d = c.alloc_device(h.size); c.h2d(d, h)\n"
        "assert pkg.crc32c.UpdateDevice(0, d + 1, 40_000_000, c) == oracle.crc32c(h[1:])\n"
        "print('ok')\n")
    env = dict(os.environ, B200_NO_FOLD="1", B200_ROWS_PER_SEG="1024")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert "ok" in subprocess.check_output([sys.executable, "-c", code], env=env, cwd=root, text=True)


def test_verify_objects_host_buffers(ctx):
    rng = np.random.default_rng(3)
    sizes = [0, 1, 100, 4096, 2 << 20, 12345, (2 << 20) + 1, 40 << 20, 7]
    bufs = [bytes(oracle.payload(s, first_word=int(rng.integers(0, 1 << 30)))) for s in sizes]
    want = [oracle.crc32c(b) for b in bufs]
    bad, got = crc32c.VerifyObjects(bufs, want, ctx)
    assert bad == [] and list(got) == want
    want[3] ^= 1
    want[7] ^= 0x80000000
    bad, got = crc32c.VerifyObjects(bufs, want, ctx)
    assert bad == [3, 7]


def test_verify_objects_host_many(ctx):
    # enough 2 MiB objects to cycle the staging slabs several times (config 1's object size)
    n = 80
    base = oracle.payload(n * (2 << 20))
    bufs = [base[i * (2 << 20):(i + 1) * (2 << 20)] for i in range(n)]
    off = np.arange(n, dtype=np.uint64) * (2 << 20)
    want, _, _ = oracle.verify_objects(base, off, np.full(n, 2 << 20, np.uint64), threads=8)
    bad, got = crc32c.VerifyObjects(bufs, want, ctx)
    assert bad == [] and (got == want).all()
    # the 2-D array form (rows = objects), pinned and pageable
    bad, got = crc32c.VerifyObjects(base.reshape(n, 2 << 20), want, ctx)
    assert bad == [] and (got == want).all()
    pinned = ctx.alloc_pinned(16 * (1 << 20))
    pinned[:] = base[:16 << 20]
    w2, _, _ = oracle.verify_objects(base, np.arange(16, dtype=np.uint64) << np.uint64(20), np.full(16, 1 << 20, np.uint64))
    w2[5] ^= 1
    bad, got = crc32c.VerifyObjects(pinned.reshape(16, 1 << 20), w2, ctx)
    assert bad == [5]
    ctx.free_pinned(pinned)


def test_sharded_single_device(ctx, arena):
    host, d = arena
    n = 48 << 20
    assert crc32c.ShardedChecksum([d], [n], ctx) == oracle.crc32c(host[:n])


def test_sharded_emulated_split_matches_combine(ctx, arena):
    # the multi-GPU rule on one GPU: shard CRCs folded left to right with Combine
    host, d = arena
    n = 60 << 20
    for parts in (2, 4, 8):
        bounds = [n * i // parts for i in range(parts + 1)]
        acc = 0
        for a, b in zip(bounds[:-1], bounds[1:]):
            acc = crc32c.Combine(acc, crc32c.UpdateDevice(0, d + a, b - a, ctx), b - a)
        assert acc == oracle.crc32c(host[:n])


def test_stream_host_pinned_and_pageable(ctx):
    n = (96 << 20) + 4321
    src = oracle.payload(n)
    want = oracle.crc32c(src)
    assert crc32c.StreamChecksum(src, chunk=16 << 20, ctx=ctx) == want          # pageable: staged
    pinned = ctx.alloc_pinned(n)
    pinned[:] = src
    assert crc32c.StreamChecksum(pinned, chunk=16 << 20, ctx=ctx) == want       # pinned: direct H2D
    assert crc32c.StreamChecksum(pinned, chunk=256 << 20, ctx=ctx) == want      # config 4 chunk size
    assert crc32c.StreamChecksum(pinned[1:], chunk=5 << 20, ctx=ctx) == oracle.crc32c(src[1:])
    ctx.free_pinned(pinned)


def test_device_generator_matches_oracle(ctx):
    nw = 1 << 18
    d = ctx.alloc_device(nw * 8)
    ctx.fill_payload(d, nw, oracle.SEED, first_word=12345)
    back = np.empty(nw * 8, dtype=np.uint8)
    ctx.d2h(back, d)
    assert (back == oracle.payload(nw * 8, first_word=12345)).all()
    ctx.free_device(d)


def test_golden_fixture(ctx):
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
    for case in g["crc"]:
        data = oracle.payload(case["nbytes"], seed=case["seed"], first_word=case["first_word"])[case["skip"]:]
        assert crc32c.Checksum(data, ctx) == case["crc32c"], case


def test_full_size_config2_properties(ctx):
    """BASELINE config 2 at full size: 4096 x 2 MiB objects = 8 GiB, generated on the device.
    Size-independent checks: (a) a sample of objects against the oracle on regenerated payload,
    (b) the fold of all 4096 object CRCs with Combine equals the CRC of the arena taken as ONE object,
    (c) one flipped bit flags exactly one object."""
    nobj, size = 4096, 2 << 20
    total = nobj * size
    d = ctx.alloc_device(total)
    ctx.fill_payload(d, total // 8, oracle.SEED)
    off = np.arange(nobj, dtype=np.uint64) * size
    ln = np.full(nobj, size, dtype=np.uint64)
    b = crc32c.DeviceBatch(off, ln, None, ctx)
    b.run(d)
    got, ok, nbad = b.results()
    assert nbad == 0
    for i in (0, 1, 1234, 2047, 4095):
        assert got[i] == oracle.crc32c(oracle.payload(size, first_word=i * size // 8)), i
    acc = 0
    for c in got:
        acc = crc32c.Combine(acc, int(c), size)
    whole = crc32c.UpdateDevice(0, d, total, ctx)
    assert acc == whole
    # negative control: flip one bit in object 1234 on the device
    byte = np.empty(1, dtype=np.uint8)
    ctx.d2h(byte, d + 1234 * size + 99)
    byte ^= 0x20
    ctx.h2d(d + 1234 * size + 99, byte)
    b2 = crc32c.DeviceBatch(off, ln, got, ctx)
    b2.run(d)
    got2, ok2, nbad2 = b2.results()
    assert nbad2 == 1 and list(np.nonzero(~ok2)[0]) == [1234]
    b.close(); b2.close()
    ctx.free_device(d)


def test_c_abi_argument_errors(ctx):
    """Bad arguments come back as B200_ERR_INVALID with a message, never a crash."""
    import ctypes as C
    L, h = ctx._L, ctx._h
    out = C.c_uint32()
    inv = pkg._lib.ERR_INVALID
    assert L.b200_crc32c_update_host(h, 0, 0, None, 10, C.byref(out)) == inv          # null data, n > 0
    assert L.b200_crc32c_update_host(h, 0, 0, None, 0, C.byref(out)) == 0 and out.value == 0
    assert L.b200_crc32c_update_host(h, 7, 0, None, 0, C.byref(out)) == inv           # device not in the context
    assert L.b200_crc32c_update_device(h, 0, 0, None, 10, C.byref(out)) == inv
    assert L.b200_crc32c_update_host(None, 0, 0, None, 0, C.byref(out)) == inv
    hb = C.c_void_p()
    assert L.b200_batch_create(h, 0, None, None, 5, None, C.byref(hb)) == inv
    assert L.b200_batch_results(None, None, None, None) == inv
    assert L.b200_hasher_write(None, None, 0) == inv
    assert L.b200_crc32c_sharded(h, None, None, 1, C.byref(out)) == inv
    assert b"b200_" in L.b200_last_error()
    n = C.c_int()
    assert L.b200_device_count(C.byref(n)) == 0 and n.value >= 1
    assert L.b200_ctx_ndev(h) == 1
    # a batch that has not run yet has no results
    off = np.zeros(1, np.uint64); ln = np.full(1, 8, np.uint64)
    assert L.b200_batch_create(h, 0, off.ctypes.data, ln.ctypes.data, 1, None, C.byref(hb)) == 0
    assert L.b200_batch_results(hb, None, None, None) == inv
    assert L.b200_batch_free(hb) == 0


def _ngpu():
    import multidev_checks
    return multidev_checks.ngpu()


@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs in one process (bench.py runs the same checks on multi-GPU boxes)")
def test_sharded_object_over_two_devices_in_one_process():
    """BASELINE config 3 in its one-process form (shards on different GPUs, host-side GF(2) fold) and the
    multi-device latency entry points, on distinct devices: tests/multidev_checks.py."""
    import multidev_checks
    multidev_checks.multi_device_parity(min(_ngpu(), 4))


def test_verify_objects_host_with_an_object_larger_than_a_slab(ctx):
    """An object bigger than the 256 MiB device slab is streamed on its own, between ordinary slabs."""
    big = oracle.payload((300 << 20) + 5)
    bufs = [b"abc", big[:1000], big, big[7:4096 + 7], b""]
    want = [oracle.crc32c(b) for b in bufs]
    bad, got = crc32c.VerifyObjects(bufs, want, ctx)
    assert bad == [] and list(got) == want
    want[2] ^= 0x100
    bad, got = crc32c.VerifyObjects(bufs, want, ctx)
    assert bad == [2]


def test_concurrent_callers_share_one_context(ctx, arena):
    """48 goroutines call into the reference's storage client at once (main.go:36,200-212); the C ABI promises
    the same: any entry point from any thread, calls on one device serialise.  ctypes drops the GIL, so these
    threads really are inside the library together."""
    from concurrent.futures import ThreadPoolExecutor
    host, d = arena
    lat = pkg.latency
    ns = oracle.latency_ns(300_000)
    dns = ctx.alloc_device(ns.nbytes)
    ctx.h2d(dns, ns)
    ref_hist, _ = oracle.hist_numpy(oracle.ns_to_ms(ns).astype(np.float64))
    srt = np.sort(oracle.ns_to_us(ns))

    def worker(w):
        rng = np.random.default_rng(w)
        for it in range(12):
            kind = (w + it) % 5
            start, n = int(rng.integers(0, 4096)), int(rng.integers(1, 6 << 20))
            want = oracle.crc32c(host[start:start + n])
            if kind == 0:
                assert crc32c.UpdateDevice(0, d + start, n, ctx) == want
            elif kind == 1:
                assert crc32c.Checksum(host[start:start + n], ctx) == want
            elif kind == 2:
                h = crc32c.New(ctx)
                h.Write(host[start:start + n // 2]); h.Write(host[start + n // 2:start + n])
                assert h.Sum32() == want
                h.close()
            elif kind == 3:
                bufs = [host[start:start + n], host[start + 5:start + 5 + n // 3]]
                bad, got = crc32c.VerifyObjects(bufs, [want, oracle.crc32c(bufs[1]) ^ 1], ctx)
                assert bad == [1] and got[0] == want
            else:
                assert list(lat.hist_readme(dns, ns.size, ctx=ctx).counts) == list(ref_hist)
                k = int(rng.integers(0, ns.size))
                assert lat.select_kth(dns, ns.size, 1000, [k], ctx=ctx) == [int(srt[k])]
                assert crc32c.UpdateDeviceNoAlloc(0, d + start, n) == want
        return True

    with ThreadPoolExecutor(max_workers=8) as ex:
        assert all(ex.map(worker, range(8)))
    ctx.free_device(dns)


def test_single_bit_walk_across_every_structural_boundary(ctx):
    """CRC is linear, so a buffer with ONE bit set exercises exactly one weight of the whole scheme: heads,
    rows, lanes, segments (row kernel and fold kernel), tails.  Random payloads can mask a wrong weight that
    only a lone bit exposes."""
    n = (12 << 20) + 531
    d = ctx.alloc_device(n + 64)
    zeros = np.zeros(n + 64, dtype=np.uint8)
    ctx.h2d(d, zeros)
    base_crc = {s: oracle.crc32c(zeros[s:n]) for s in (0, 5)}
    seg = 2048 * 512
    positions = sorted({0, 1, 4, 5, 15, 16, 17, 511, 512, 513, 1023, 8191, 8192, 16383, 16384, 32767, 32768,
                        65535, 65536, 65537, (1 << 20) - 1, 1 << 20, (1 << 20) + 1, seg - 1, seg, seg + 1,
                        2 * seg + 300, 5 * seg - 1, 5 * seg, 11 * seg + 7, n - 532, n - 531, n - 530, n - 17, n - 16,
                        n - 2, n - 1} | {k * 32768 + 5 for k in range(1, 40, 7)} | {k * 16384 for k in range(1, 64, 9)})
    one = np.zeros(1, dtype=np.uint8)
    for pos in positions:
        for bit in (0, 7):
            one[0] = 1 << bit
            ctx.h2d(d + pos, one)
            for start in (0, 5):
                if pos < start:
                    continue
                host = zeros[start:n].copy()
                host[pos - start] = one[0]
                got = crc32c.UpdateDevice(0, d + start, n - start, ctx)
                assert got == oracle.crc32c(host), (pos, bit, start)
                assert got != base_crc[start]
            one[0] = 0
            ctx.h2d(d + pos, one)
    ctx.free_device(d)


@pytest.mark.parametrize("rps", ["256", "1024"])
def test_single_bit_walk_through_the_fold_kernel(rps):
    """Same walk with the segment length forced so the fold kernel takes the full segments: lone bits at stage
    (256 B), lane (rps*16 B) and segment (rps*512 B) boundaries, in the first/last word of the 64-word window."""
    import subprocess, sys
    code = (
        "import importlib, os, numpy as np, oracle\n"
        "pkg = importlib.import_module('custom-go-client-benchmark_b200')\n"
        "c = pkg.crc32c.Context([0])\n"
        "rps = int(os.environ['B200_ROWS_PER_SEG']); S = rps * 16; seg = rps * 512\n"
        "n = 5 * seg + 777\n"
        "z = np.zeros(n + 32, dtype=np.uint8)\n"
        "d = c.alloc_device(z.size); c.h2d(d, z)\n"
        "pos = sorted({0, 3, 4, 252, 255, 256, 259, 260, 511, 512, S - 1, S, S + 4, 2 * S - 1, 31 * S, 31 * S + 255, seg - 4,\n"
        "              seg - 1, seg, seg + 1, seg + S, 3 * seg + 17 * S + 300, 4 * seg - 1, 5 * seg - 1, 5 * seg, n - 1})\n"
        "one = np.zeros(1, dtype=np.uint8)\n"
        "for p in pos:\n"
        "    for start in (0, 3):\n"
        "        if p < start: continue\n"
        "        one[0] = 0x80; c.h2d(d + p, one)\n"
        "        h = z[start:n].copy(); h[p - start] = 0x80\n"
        "        assert pkg.crc32c.UpdateDevice(0xABCD1234, d + start, n - start, c) == oracle.crc32c(h, 0xABCD1234), (p, start)\n"
        "        one[0] = 0; c.h2d(d + p, one)\n"
        "print('ok')\n")
    env = dict(os.environ, B200_ROWS_PER_SEG=rps)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert "ok" in subprocess.check_output([sys.executable, "-c", code], env=env, cwd=root, text=True)
