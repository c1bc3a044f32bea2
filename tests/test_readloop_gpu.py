"""BASELINE config 1: the reference's read loop (main.go:119-156) over an in-memory reader, CRC-32C checked on
the GPU at every EOF, latencies reduced on the GPU — against the same loop done with the CPU oracle."""
import importlib

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, readloop = pkg.crc32c, pkg.readloop
MB = 1024 * 1024


@pytest.fixture(scope="module")
def ctx():
    c = crc32c.Context([0])
    yield c
    c.close()


def make_objects(n, size=2 * MB, max_read=2 * MB):
    objs = []
    for w in range(n):
        data = oracle.payload(size, first_word=w * size // 8)
        objs.append(readloop.InMemoryObject(data, oracle.crc32c(data), max_read))
    return objs


class FakeClock:
    """Deterministic time.Now(): successive (start, end) pairs differ by the oracle's synthetic latencies."""

    def __init__(self, first):
        self.i, self.t, self.first, self.started = 0, 0, first, False

    def __call__(self):
        if self.started:
            self.t += int(oracle.latency_ns(1, first=self.first + self.i)[0])
            self.i += 1
        self.started = not self.started
        return self.t


def test_read_loop_matches_cpu_loop(ctx):
    workers, reads = 6, 40
    objs = make_objects(workers, max_read=300_000)  # short reads: chunk boundaries must not matter
    parts = [readloop.ReadObject(w, objs[w], reads, ctx, FakeClock(w * reads)) for w in range(workers)]
    assert all(not p.errors and p.reads == reads and p.bytes_read == reads * 2 * MB for p in parts)
    lat = np.concatenate([p.latencies_ns for p in parts])
    assert (lat == oracle.latency_ns(workers * reads)).all()
    readme, oc, pct = readloop.reduce_latencies(lat, ctx)
    ms = oracle.ns_to_ms(lat)
    ref, dropped = oracle.hist_numpy(ms.astype(np.float64))
    assert list(readme.counts) == list(ref) and readme.dropped_below + readme.dropped_above == dropped
    assert list(oc.counts) == list(oracle.hist_opencensus(ms.astype(np.float64)))
    o = oracle.percentiles_i64(oracle.ns_to_us(lat))
    assert (pct.avg, pct.p20, pct.p50, pct.p90, pct.p99, pct.min, pct.max) == (o.avg, o.p20, o.p50, o.p90, o.p99,
                                                                                o.min, o.max)


def test_read_loop_concurrent_workers(ctx):
    objs = make_objects(8)
    res = readloop.RunReadLoop(objs, 5, ctx)
    assert res.errors == [] and res.reads == 40 and res.bytes_read == 40 * 2 * MB
    assert res.latencies_ns.size == 40 and (res.latencies_ns > 0).all()
    readme, oc, pct = readloop.reduce_latencies(res.latencies_ns, ctx)
    assert int(oc.counts.sum()) == 40 and pct.min <= pct.p50 <= pct.max


def test_bad_crc_stops_that_worker_with_the_reference_error(ctx):
    objs = make_objects(3, size=1 * MB + 123)
    bad = bytearray(objs[1].data)
    bad[500_000] ^= 0x40
    objs[1] = readloop.InMemoryObject(bytes(bad), objs[1].crc32c)
    res = readloop.RunReadLoop(objs, 4, ctx)
    assert len(res.errors) == 1 and "storage: bad CRC on read: got" in res.errors[0] and "worker 1" in res.errors[0]
    assert res.reads == 8  # the two good workers finished, the bad one stopped at its first read


def test_empty_object_and_empty_sample_set(ctx):
    obj = readloop.InMemoryObject(b"", 0)
    r = readloop.ReadObject(0, obj, 3, ctx)
    assert not r.errors and r.bytes_read == 0
    readme, oc, pct = readloop.reduce_latencies(np.zeros(0, dtype=np.int64), ctx)
    assert pct is None and int(oc.counts.sum()) == 0 and oc.count == 0
