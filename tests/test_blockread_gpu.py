"""The reference's local-SSD benchmark (benchmark-script/ssd_test/main.go) restated: block reads into pinned
buffers, per-read durations reduced on the GPU with ssd_test's percentile rule, blocks optionally CRC-verified."""
import importlib
import os
import random

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, blockread = pkg.crc32c, pkg.blockread


@pytest.fixture(scope="module")
def ctx():
    c = crc32c.Context([0])
    yield c
    c.close()


def make_workload(root, threads, file_kb):
    """<dir>/Workload.<i>/0 as the reference expects (main.go:41); returns the bytes of each file."""
    files = []
    for i in range(threads):
        os.makedirs(os.path.join(root, f"Workload.{i}"), exist_ok=True)
        data = oracle.payload(file_kb * 1024, first_word=i * file_kb * 128)
        with open(os.path.join(root, f"Workload.{i}", "0"), "wb") as f:
            f.write(bytes(data))
        files.append(data)
    return files


def supports_o_direct(root):
    try:
        os.close(os.open(os.path.join(root, "Workload.0", "0"), os.O_RDONLY | os.O_DIRECT))
        return True
    except OSError:
        return False


@pytest.mark.parametrize("read_type", ["seq", "rand"])
def test_block_reads_verify_and_reduce(ctx, tmp_path, read_type):
    threads, block_kb, file_kb, count = 3, 256, 2048, 2
    files = make_workload(str(tmp_path), threads, file_kb)
    nblk = file_kb // block_kb
    want = [[oracle.crc32c(f[b * block_kb * 1024:(b + 1) * block_kb * 1024]) for b in range(nblk)] for f in files]
    r = blockread.RunReadFileOperations(str(tmp_path), threads, block_kb, file_kb, read_type, count, ctx,
                                        direct=supports_o_direct(str(tmp_path)), want_crc32c=want,
                                        rng=random.Random(7))
    assert r.bad_blocks == [] and r.read_time_ns.size == threads * count * nblk
    assert r.bytes_read == threads * count * file_kb * 1024
    o = oracle.percentiles_i64(oracle.ns_to_us(r.read_time_ns))  # sort + floor-index picks, ssd_test/main.go:146-163
    p = r.percentiles
    assert (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) == (o.avg, o.p20, o.p50, o.p90, o.p99, o.min, o.max)
    lines = r.report().split("\n")
    assert lines[0] == "read benchmark completed successfully!" and lines[1] == "Average: %.3f ms" % (0.001 * o.avg)
    assert lines[5] == "p99: %.3f ms" % (0.001 * o.p99) and len(lines) == 8


def test_block_read_flags_a_corrupt_block(ctx, tmp_path):
    files = make_workload(str(tmp_path), 2, 1024)
    want = [[oracle.crc32c(f[b * 262144:(b + 1) * 262144]) for b in range(4)] for f in files]
    want[1][2] ^= 0x10
    r = blockread.RunReadFileOperations(str(tmp_path), 2, 256, 1024, "seq", 1, ctx,
                                        direct=supports_o_direct(str(tmp_path)), want_crc32c=want)
    assert r.bad_blocks == [(1, 2)]


def test_access_pattern_is_the_reference_shuffle():
    assert blockread.access_pattern(5, "seq") == [0, 1, 2, 3, 4]
    got = blockread.access_pattern(64, "random", random.Random(1))
    rng, ref = random.Random(1), list(range(64))
    for i in range(64):  # main.go:124-127
        j = rng.randrange(i + 1)
        ref[i], ref[j] = ref[j], ref[i]
    assert got == ref and sorted(got) == list(range(64))


def test_reference_error_cases(ctx, tmp_path):
    make_workload(str(tmp_path), 1, 1024)
    direct = supports_o_direct(str(tmp_path))
    with pytest.raises(blockread.BlockReadError, match="you must set --dir flag"):
        blockread.RunReadFileOperations("", ctx=ctx)
    with pytest.raises(blockread.BlockReadError, match="threads count not valid"):
        blockread.RunReadFileOperations(str(tmp_path), 0, ctx=ctx)
    with pytest.raises(blockread.BlockReadError, match="file present is not equal to given file-size"):
        blockread.RunReadFileOperations(str(tmp_path), 1, 256, 2048, ctx=ctx, direct=direct)
    with pytest.raises(blockread.BlockReadError, match="block-size should be multiple of file-size"):
        blockread.RunReadFileOperations(str(tmp_path), 1, 300, 1024, ctx=ctx, direct=direct)
    with pytest.raises(blockread.BlockReadError, match="while opening file"):
        blockread.RunReadFileOperations(str(tmp_path), 2, 256, 1024, ctx=ctx, direct=direct)  # no Workload.1
    empty = tmp_path / "empty"
    os.makedirs(empty / "Workload.0")
    open(empty / "Workload.0" / "0", "wb").close()
    with pytest.raises(ZeroDivisionError):  # no reads: main.go:157 divides by len(readTime)
        blockread.RunReadFileOperations(str(empty), 1, 256, 0, ctx=ctx, direct=direct)
