#!/usr/bin/env python
"""Benchmark of the CRC-32C verify path (BASELINE.json metric) on 1..8 B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One process per GPU; for N > 1 launch with torch.distributed.run (RANK / LOCAL_RANK / WORLD_SIZE env).
A step is one VerifyObjects pass over this rank's batch: 4096 objects x 2 MiB = 8 GiB, device resident
(BASELINE config 2; weak scaling: every GPU owns its own 8 GiB batch, objects are independent, so there is
no data-path collective — only the timing barrier and the max-over-ranks reduction use NCCL).
`value`   : whole-job GB/s with inputs resident in HBM, timed with CUDA events on the library's stream.
`e2e`     : the same batch through the host-buffer entry point b200_crc32c_batch_host (what a Go caller's
            VerifyObjects([]buf, []want) binds): pinned host buffers, H2D copies inside the timed region.
`roofline`: the dominant kernel (crc_fold_kernel; every segment of this config is a full one, so the row
            kernel does not run): algorithmic bytes (1 byte read per byte hashed) / its mean CUDA-event
            duration over the timed steps, against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
`cpu_baseline`: the oracle's SSE4.2 three-lane CRC-32C (C restatement of Go's amd64 path; Go is not installed
            on the box) on the box's host cores, over the same objects.
--impl reference times that CPU implementation alone, all host threads, same config/metric.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NOBJ = 4096
OBJ_BYTES = 2 << 20
SEED = 0xC0FFEE
METRIC = "crc32c_verify_throughput"
UNIT = "GB/s"
WORKLOAD = ("config 2: batched CRC32C verify of 4096 x 2 MiB device-resident objects per GPU "
            "(splitmix64 payload, seed 0xC0FFEE), bit-exact vs Go hash/crc32 Castagnoli")


def config(n_gpus):
    return {"workload": WORKLOAD, "objects_per_gpu": NOBJ, "object_bytes": OBJ_BYTES,
            "bytes_per_gpu_per_step": NOBJ * OBJ_BYTES, "sharding": f"objects x {n_gpus} gpus, no collective",
            "l2_policy": "inputs larger than L2 (8 GiB arena per GPU vs 126 MB L2), no flush needed"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (torch copy, read+write)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md); MEASURED_PEAKS.json absent"


def cuda_pci_bus_id(index: int):
    """PCI bus id of CUDA device `index` (NVML numbers GPUs differently under CUDA_VISIBLE_DEVICES)."""
    import ctypes
    lib = importlib.import_module("custom-go-client-benchmark_b200")._lib.load()
    buf = ctypes.create_string_buffer(32)
    return buf.value.decode() if lib.b200_device_pci_bus_id(index, buf, 32) == 0 else None


def nvml_handle(pynvml, index: int):
    bus = cuda_pci_bus_id(index)
    if bus:
        try:
            return pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            pass
    return pynvml.nvmlDeviceGetHandleByIndex(index)


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = nvml_handle(pynvml, index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # NVML missing: report that rather than invent numbers
            self.nv, self.err = None, repr(e)

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
                 "hw_thermal_slowdown": 0x40, "hw_power_brake": 0x80, "sync_boost": 0x10, "app_clocks": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.005)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": self.err}
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def bind_to_gpu_numa_node(index: int):
    """Pin this process to the CPUs local to its GPU before any pinned memory is allocated, so that the
    pinned host buffers of the e2e path land on the GPU's own NUMA node (8 ranks streaming 8 GiB each would
    otherwise cross sockets).  Best effort: returns a description for the JSON line."""
    try:
        bus = cuda_pci_bus_id(index)
        if not bus:
            return "not bound (no PCI bus id)"
        path = f"/sys/bus/pci/devices/{bus[-12:].lower()}/local_cpulist"
        cpus = set()
        for part in open(path).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} cpus local to {bus}"
    except Exception as e:
        return f"not bound ({type(e).__name__})"
    return "not bound"


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as d
        torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        d.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = d
    return rank, world, local, dist


def cpu_baseline_run(host_arena, off, ln, threads, budget_s=12.0):
    """Times the oracle (C restatement of Go's SSE4.2 Castagnoli path) over the batch; returns GB/s."""
    import oracle
    nbytes = int(ln.sum())
    reps, t_total, got = 0, 0.0, None
    while True:
        t0 = time.perf_counter()
        got, _, _ = oracle.verify_objects(host_arena, off, ln, None, threads=threads)
        t_total += time.perf_counter() - t0
        reps += 1
        if t_total > budget_s or reps >= 5:
            break
    return nbytes * reps / t_total / 1e9, got, reps


def run_reference(args):
    """--impl reference: the reference path's own CPU implementation (oracle port; the reference is Go and
    neither Go nor its modules exist on the box), all host threads, same metric/config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    cores = os.cpu_count() or 1
    nobj = NOBJ
    arena = oracle.payload(nobj * OBJ_BYTES, seed=SEED)
    off = np.arange(nobj, dtype=np.uint64) * OBJ_BYTES
    ln = np.full(nobj, OBJ_BYTES, dtype=np.uint64)
    # bound the run: a step is the whole batch unless that would take minutes in total
    t0 = time.perf_counter()
    oracle.verify_objects(arena, off, ln, None, threads=cores)
    one = time.perf_counter() - t0
    frac = 1.0
    if one * (args.steps + args.warmup) > 150:
        frac = max(150 / (one * (args.steps + args.warmup)), 1 / 64)
    m = max(int(nobj * frac), 1)
    for _ in range(args.warmup):
        oracle.verify_objects(arena, off[:m], ln[:m], None, threads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.verify_objects(arena, off[:m], ln[:m], None, threads=cores)
    dt = time.perf_counter() - t0
    ms = dt / args.steps * 1e3
    val = m * OBJ_BYTES / (ms * 1e-3) / 1e9
    kind = "port"
    sample = f"{m} of {nobj} objects x 2 MiB per step, {args.steps} steps, {cores} threads (pthreads over objects)"
    line = {"impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic", "config": config(args.gpus),
            "cpu_baseline": {"value": round(val, 3), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                             "sse42": bool(oracle.lib().oracle_crc32c_hw_available()),
                             "note": "C restatement of Go hash/crc32's amd64 Castagnoli path; Go unavailable on this box"},
            "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def latency_side_measurement(pkg, ctx, n=500_000_000, reps=5):
    """SURVEY 8's second row (per-read latency samples -> README histogram, ssd_test percentiles), measured
    beside the headline metric so the round's bench record holds it: n synthetic int64 ns samples resident in
    HBM, wall clock around the C-ABI calls (each ends with a device->host copy of the result).  Not part of the
    timed region above; never fails the bench."""
    try:
        lat = pkg.latency
        d = ctx.alloc_device(n * 8)
        ctx.fill_latency_ns(d, n, SEED)

        def ms(f):
            f()
            t0 = time.perf_counter()
            for _ in range(reps):
                f()
            return (time.perf_counter() - t0) / reps * 1e3
        h_ms = ms(lambda: lat.hist_readme(d, n, ctx=ctx))
        p_ms = ms(lambda: lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx))
        h = lat.hist_readme(d, n, ctx=ctx)
        p = lat.percentiles(d, n, lat.NS_PER_US, ctx=ctx)
        assert int(h.counts.sum()) + h.dropped_below + h.dropped_above == n == h.count
        assert p.min <= p.p20 <= p.p50 <= p.p90 <= p.p99 <= p.max
        ctx.free_device(d)
        gbs = lambda t: round(n * 8 / (t * 1e-3) / 1e9, 1)
        return {"samples": n, "bytes_per_sample": 8,
                "readme_histogram_with_moments": {"ms": round(h_ms, 3), "GB/s": gbs(h_ms)},
                "ssd_test_percentiles_avg_p20_p50_p90_p99_min_max": {"ms": round(p_ms, 3), "GB/s": gbs(p_ms)},
                "us": {"avg": p.avg, "p20": p.p20, "p50": p.p50, "p90": p.p90, "p99": p.p99, "min": p.min, "max": p.max},
                "note": "consistency only here (counts + dropped == n, ordered percentiles); parity with the oracle is "
                        "tests/test_latency_gpu.py"}
    except Exception as e:  # a side measurement must not cost the headline line
        return {"error": f"{type(e).__name__}: {e}"[:300]}


def run_b200(args):
    rank, world, local, dist = dist_setup(args.gpus)
    affinity = bind_to_gpu_numa_node(local) if world > 1 else "single rank: not bound"
    pkg = importlib.import_module("custom-go-client-benchmark_b200")
    crc32c = pkg.crc32c
    ctx = crc32c.Context([local])
    total = NOBJ * OBJ_BYTES
    d_arena = ctx.alloc_device(total)
    # rank r owns words [r*total/8, (r+1)*total/8) of the splitmix stream
    ctx.fill_payload(d_arena, total // 8, SEED, first_word=rank * (total // 8))
    off = np.arange(NOBJ, dtype=np.uint64) * OBJ_BYTES
    ln = np.full(NOBJ, OBJ_BYTES, dtype=np.uint64)

    # expected checksums: first pass on the GPU (parity against the oracle is checked below on rank 0
    # and exhaustively in tests/); the timed steps then VERIFY against them.
    b0 = crc32c.DeviceBatch(off, ln, None, ctx)
    b0.run(d_arena)
    want, _, _ = b0.results()
    b0.close()
    batch = crc32c.DeviceBatch(off, ln, want, ctx)

    def barrier():
        if dist is not None:
            dist.barrier()

    for _ in range(max(args.warmup, 0)):
        batch.run(d_arena)
    batch.kernel_ms()  # reset the per-run event window
    barrier()
    with ClockSampler(local) as clk:
        ctx.timer_start()
        for _ in range(args.steps):
            batch.run_async(d_arena)
        elapsed_ms = ctx.timer_stop()
    barrier()
    got, ok, nbad = batch.results()
    assert nbad == 0 and (got == want).all(), "verification failed inside the timed region"
    _, rows_ms = batch.kernel_ms()
    launches = batch.launches() * args.steps
    ms_per_step = elapsed_ms / args.steps
    if dist is not None:
        import torch
        t = torch.tensor([ms_per_step, rows_ms], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_per_step, rows_ms = float(t[0]), float(t[1])
    value = world * total / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: host buffers through the C-ABI VerifyObjects entry point ---------------------------------
    h_arena = ctx.alloc_pinned(total)
    ctx.d2h(h_arena, d_arena)
    bufs = h_arena.reshape(NOBJ, OBJ_BYTES)  # one pinned 2 MiB buffer per object, as the read loop holds them
    e2e_steps = max(1, min(args.steps, 5))
    bad, g = crc32c.VerifyObjects(bufs, want, ctx)  # warm-up (allocates the slabs)
    assert bad == [] and (g == want).all()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        bad, g = crc32c.VerifyObjects(bufs, want, ctx)
    e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
    barrier()
    assert bad == []
    if dist is not None:
        import torch
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t[0])
    e2e_val = world * total / (e2e_ms * 1e-3) / 1e9

    line = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = total / (rows_ms * 1e-3) / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                traffic = json.load(f).get("crc_fold_kernel_dram_bytes_per_launch")
        except Exception:
            pass
        line = {"metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_per_step, 5), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic", "config": config(world),
                "clocks": clk.summary(),
                "e2e": {"value": round(e2e_val, 2), "unit": UNIT, "h2d_bytes_per_step": total,
                        "d2h_bytes_per_step": NOBJ * 5 + 8, "steps": e2e_steps, "ms_per_step": round(e2e_ms, 3),
                        "api": "b200_crc32c_batch_host (VerifyObjects on pinned host buffers)",
                        "pcie_note": "PCIe Gen5 x16 bound (~64 GB/s raw per direction)", "cpu_affinity": affinity},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                             "frac": round(achieved / peak, 4), "traffic": traffic,
                             "frac_of_nominal_8TBps": round(achieved / 8000.0, 4), "kernel": "crc_fold_kernel",
                             "kernel_ms": round(rows_ms, 5), "algorithmic_bytes_per_launch": total,
                             "peak_source": peak_src + " — of measured"}}
        if world == 1:
            cores = os.cpu_count() or 1
            allc, got_cpu, reps = cpu_baseline_run(h_arena, off, ln, cores)
            assert (got_cpu == want).all(), "GPU CRCs differ from the CPU oracle"
            onec, _, _ = cpu_baseline_run(h_arena, off[:512], ln[:512], 1, budget_s=4.0)
            line["cpu_baseline"] = {"value": round(allc, 2), "unit": UNIT, "cores": cores, "kind": "port",
                                    "single_core": round(onec, 2),
                                    "sample": f"all {NOBJ} objects x 2 MiB, {reps} passes on {cores} threads; "
                                              "single_core: first 512 objects on 1 thread",
                                    "note": "oracle = C restatement of Go hash/crc32 amd64 Castagnoli "
                                            "(SSE4.2 crc32q, 3 lanes); Go unavailable on this box"}
    ctx.free_pinned(h_arena)
    ctx.free_device(d_arena)
    batch.close()
    if line is not None and world == 1:
        line["latency_reduction"] = latency_side_measurement(pkg, ctx)
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
