"""Latency of the small all-reduces the multi-rank latency reduction uses (torchrun, one process per GPU)."""
import importlib, os, sys, time
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
lat = pkg.latency
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
g = dist.new_group(backend="gloo")
for name, red in (("nccl via device tensor", lat._Reducer(None, f"cuda:{local}")), ("gloo host tensor", lat._Reducer(g, None))):
    for n in (4, 16, 4096 * 4):
        a = np.arange(n, dtype=np.uint64)
        red.sum_u64(a)
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(50):
            red.sum_u64(a)
        dt = (time.perf_counter() - t0) / 50 * 1e3
        if rank == 0:
            print(f"{name:24s} sum_u64 of {n:6d} values: {dt:.3f} ms", flush=True)
dist.destroy_process_group()
