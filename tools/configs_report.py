#!/usr/bin/env python
"""Runs BASELINE.json configs 3, 4 and 5 at full size and prints one JSON object per config.

    python tools/configs_report.py [--configs 3,4,5] [--gib 64] [--samples 1000000000]
    torchrun --nproc-per-node N tools/configs_report.py ...      (one process per GPU)

config 3: one --gib GiB synthetic object sharded over the ranks (rank r owns bytes [r*L/N,(r+1)*L/N)), per-GPU
          CRCs folded with the GF(2) combine on rank 0.  Checks: fold of 1 GiB pieces == whole shard (GPU),
          first and last piece of rank 0 against the CPU oracle on regenerated payload.
config 4: 16 GiB pinned host buffer streamed in 256 MiB chunks (H2D overlapped with the CRC); PCIe-bound.
config 5: --samples latency samples split over the ranks; README 15-bin histogram, OpenCensus 35 buckets,
          exact p20/p50/p90/p99 by radix select; bin and digit counts all-reduced over NCCL.
The oracle is used only as the checker.  Timings are CUDA-event or wall-clock around synchronous C-ABI calls.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SEED = 0xC0FFEE


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="3,4,5")
    ap.add_argument("--gib", type=int, default=64)
    ap.add_argument("--samples", type=int, default=1_000_000_000)
    ap.add_argument("--stream-gib", type=int, default=16)
    args = ap.parse_args()
    want = {int(c) for c in args.configs.split(",")}
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as d
        torch.cuda.set_device(local)
        d.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = d
    pkg = importlib.import_module("custom-go-client-benchmark_b200")
    crc32c, lat = pkg.crc32c, pkg.latency
    import oracle
    ctx = crc32c.Context([local])

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    if 3 in want:
        L = args.gib << 30
        lo, hi = L * rank // world, L * (rank + 1) // world
        lo, hi = lo // 8 * 8, hi // 8 * 8
        n = hi - lo
        d = ctx.alloc_device(n)
        ctx.fill_payload(d, n // 8, SEED, first_word=lo // 8)
        crc = crc32c.UpdateDevice(0, d, n, ctx)  # warm-up + value
        barrier()
        reps = 5
        t0 = time.perf_counter()
        for _ in range(reps):
            crc = crc32c.UpdateDevice(0, d, n, ctx)
        wall_ms = (time.perf_counter() - t0) / reps * 1e3
        ms = max_over_ranks(wall_ms)
        # consistency: fold of 1 GiB pieces == whole shard
        piece = 1 << 30
        acc = 0
        for o in range(0, n, piece):
            ln = min(piece, n - o)
            acc = crc32c.Combine(acc, crc32c.UpdateDevice(0, d + o, ln, ctx), ln)
        assert acc == crc, "piecewise fold differs from single-object CRC"
        # oracle spot checks on regenerated payload (first and last 64 MiB of this rank's shard)
        chk = 64 << 20
        for o in (0, n - chk):
            ref = oracle.crc32c(oracle.payload(chk, first_word=(lo + o) // 8))
            assert crc32c.UpdateDevice(0, d + o, chk, ctx) == ref, "GPU CRC differs from the oracle"
        parts = [(crc, n)]
        if dist is not None:
            parts = [None] * world
            dist.all_gather_object(parts, (crc, n))
        if rank == 0:
            total = 0
            for c, ln in parts:
                total = crc32c.Combine(total, c, ln)
            print(json.dumps({"config": 3, "object_gib": args.gib, "n_gpus": world, "crc32c": f"{total:#010x}",
                              "ms": round(ms, 3), "GB/s": round(L / ms / 1e6, 1),
                              "checks": "piecewise fold == whole; oracle spot checks ok"}))
        ctx.free_device(d)

    if 4 in want and rank == 0:
        n = args.stream_gib << 30
        host = ctx.alloc_pinned(n)
        d = ctx.alloc_device(1 << 30)
        for o in range(0, n, 1 << 30):  # fill the pinned buffer from the device generator
            ctx.fill_payload(d, (1 << 30) // 8, SEED, first_word=o // 8)
            ctx.d2h(host[o:o + (1 << 30)], d)
        ctx.free_device(d)
        crc = crc32c.StreamChecksum(host, chunk=256 << 20, ctx=ctx)
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            crc = crc32c.StreamChecksum(host, chunk=256 << 20, ctx=ctx)
        ms = (time.perf_counter() - t0) / reps * 1e3
        ref = oracle.verify_objects(host, np.array([0], np.uint64), np.array([1 << 30], np.uint64))[0][0]
        assert crc32c.StreamChecksum(host[:1 << 30], chunk=256 << 20, ctx=ctx) == ref
        acc = 0
        for o in range(0, n, 4 << 30):
            acc = crc32c.Combine(acc, crc32c.StreamChecksum(host[o:o + (4 << 30)], chunk=128 << 20, ctx=ctx), 4 << 30)
        assert acc == crc
        print(json.dumps({"config": 4, "host_gib": args.stream_gib, "chunk_mib": 256, "crc32c": f"{crc:#010x}",
                          "ms": round(ms, 2), "GB/s": round(n / ms / 1e6, 2), "pcie_gen5_x16_raw_GBps": 64,
                          "checks": "first GiB vs oracle; fold of 4 GiB pieces (128 MiB chunks) == whole"}))
        ctx.free_pinned(host)
    barrier()

    if 5 in want:
        N = args.samples
        lo, hi = N * rank // world, N * (rank + 1) // world
        n = hi - lo
        d = ctx.alloc_device(n * 8)
        ctx.fill_latency_ns(d, n, SEED, first=lo)
        rd = f"cuda:{local}" if dist is not None else None
        lat.hist_readme(d, min(n, 1000), ctx=ctx, reduce_device=rd)  # warm-up
        barrier()
        t0 = time.perf_counter()
        h = lat.hist_readme(d, n, ctx=ctx, reduce_device=rd)
        t1 = time.perf_counter()
        oc = lat.hist_opencensus(d, n, ctx=ctx, reduce_device=rd)
        t2 = time.perf_counter()
        both = lat.hist_readme_and_opencensus(d, n, ctx=ctx, reduce_device=rd)
        t2c = time.perf_counter()
        assert list(both[0].counts) == list(h.counts) and list(both[1].counts) == list(oc.counts)
        p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx, reduce_device=rd)  # first call sizes the scratch buffers
        barrier()
        t2b = time.perf_counter()
        p2 = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx, reduce_device=rd)
        t3 = time.perf_counter()
        assert p2 == p
        hist_ms, oc_ms, pct_ms = (max_over_ranks(x * 1e3) for x in (t1 - t0, t2 - t1, t3 - t2b))
        assert int(h.counts.sum()) + h.dropped_below + h.dropped_above == N and int(oc.counts.sum()) == N
        # oracle check on a prefix of rank 0's slice
        m = min(n, 2_000_000)
        if rank == 0:
            hp = lat.hist_fixed(d, m, lat.NS_PER_MS, lat.README_EDGES, pkg._lib.HIST_NUMPY, ctx=ctx, group=None) \
                if dist is None else None
            if hp is not None:
                ref, _ = oracle.hist_numpy(oracle.ns_to_ms(oracle.latency_ns(m)).astype(np.float64))
                assert list(hp.counts) == list(ref)
            print(json.dumps({"config": 5, "samples": N, "n_gpus": world,
                              "readme15_counts": [int(x) for x in h.counts],
                              "dropped": h.dropped_below + h.dropped_above,
                              "oc35_counts": [int(x) for x in oc.counts], "mean_ms": oc.mean,
                              "us": {"avg": p.avg, "p20": p.p20, "p50": p.p50, "p90": p.p90, "p99": p.p99,
                                     "min": p.min, "max": p.max},
                              "hist_ms": round(hist_ms, 3), "oc_ms": round(oc_ms, 3),
                              "hist_and_oc_one_pass_ms": round(max_over_ranks((t2c - t2) * 1e3), 3),
                              "percentiles_ms": round(pct_ms, 3),
                              "hist_GBps_at_8B_per_sample": round(N * 8 / hist_ms / 1e6, 1),
                              "percentiles_GBps_at_8B_per_sample": round(N * 8 / pct_ms / 1e6, 1)}))
        ctx.free_device(d)
    barrier()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
