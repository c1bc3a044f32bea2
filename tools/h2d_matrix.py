#!/usr/bin/env python
"""Host->device copy bandwidth of every GPU alone, of every PAIR of GPUs copying at once, and of the first
1/2/4/8 GPUs together: which GPUs share a PCIe uplink (or a host-memory limit) shows up as a pair whose
concurrent rate per GPU is well below its solo rate.  One process, one thread per GPU, bare cudaMemcpy of
pinned buffers through the C ABI (b200_memcpy_h2d); no kernels.  Prints one JSON object.

    python tools/h2d_matrix.py [--mib 1024] [--reps 3] > gpurun_out/h2d_matrix.json
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    pkg = importlib.import_module("custom-go-client-benchmark_b200")
    import ctypes as C
    n = C.c_int()
    pkg._lib.check(pkg._lib.load().b200_device_count(C.byref(n)))
    nd = n.value
    ctx = pkg.crc32c.Context(list(range(nd)))
    nbytes = args.mib << 20
    host = [ctx.alloc_pinned(nbytes) for _ in range(nd)]
    for h in host:
        h[::4096] = 1  # touch every page
    dev = [ctx.alloc_device(nbytes, d) for d in range(nd)]

    def run(group):
        """GB/s per GPU of `group` copying concurrently (each its own pinned buffer), reps copies each."""
        out = {}
        start = threading.Barrier(len(group))

        def work(d):
            ctx.h2d(dev[d], host[d], d)  # warm
            start.wait()
            t0 = time.perf_counter()
            for _ in range(args.reps):
                ctx.h2d(dev[d], host[d], d)
            out[d] = nbytes * args.reps / (time.perf_counter() - t0) / 1e9
        th = [threading.Thread(target=work, args=(d,)) for d in group]
        for t in th:
            t.start()
        for t in th:
            t.join()
        return [round(out[d], 2) for d in group]

    res = {"gpus": nd, "mib": args.mib, "reps": args.reps}
    res["solo"] = [run([d])[0] for d in range(nd)]
    pairs = {}
    for a in range(nd):
        for b in range(a + 1, nd):
            pairs[f"{a},{b}"] = run([a, b])
    res["pairs_GBps_each"] = pairs
    res["first_k_together"] = {str(k): run(list(range(k))) for k in (1, 2, 4, 8) if k <= nd}
    res["first_k_aggregate"] = {k: round(sum(v), 1) for k, v in res["first_k_together"].items()}
    # pairs whose members lose more than a quarter of their solo rate: a shared uplink or memory path
    res["sharing_pairs"] = [p for p, (x, y) in pairs.items()
                            if x < 0.75 * res["solo"][int(p.split(",")[0])] or y < 0.75 * res["solo"][int(p.split(",")[1])]]
    print(json.dumps(res))
    ctx.close()


if __name__ == "__main__":
    main()
