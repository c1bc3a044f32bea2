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
Beside the headline line's keys, at EVERY N:
`config3` : BASELINE config 3 — ONE 64 GiB object (the splitmix stream) sharded over the N ranks, rank r hashes bytes
            [r*L/N, (r+1)*L/N), the (crc, len) pairs are all-gathered and folded with the GF(2) combine on the host;
            strong scaling; the CRC is compared with tests/golden/config3.json (made on the CPU).
`config5` : BASELINE config 5 — 10^9 latency samples split over the N ranks: README histogram, OpenCensus buckets and
            the seven ssd_test lines with the counts gathered / all-reduced over NCCL (collectives counted); compared with tests/golden/config5.json.
`e2e.h2d_only` : a bare cudaMemcpy of the same pinned arena on every rank at once — the ceiling of the host-buffer path.
`multi_device_parity` : the one-process multi-GPU entry points checked on distinct devices (tests/multidev_checks.py).
At N = 1 also: `small_objects` (a million 4 KiB blocks, a million ragged objects under 8 KiB, 4096 ragged
objects up to 4 MiB; device-timed, a sample of each checked against the oracle), `text_codec` (20 M lines of latency
text parsed on the device, the first block against Python's float()), `latency_reduction` (5*10^8 samples: README
histogram with moments, the seven ssd_test lines).
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


def pick_device(local_rank, world):
    """Which visible GPU a rank takes.  The 8-GPU boxes of this pool feed GPUs 0-3 and GPUs 4-7 through two host
    paths of ~115-140 GB/s each (profiles/r02_h2d_matrix_8gpu.json: any two GPUs copy at 55.5 GB/s each, GPUs 0-3
    together at 28.8 each), so when fewer ranks than GPUs run they are dealt alternately to the two halves instead
    of filling the first half.  With one rank, or as many ranks as GPUs, this changes nothing that matters."""
    try:
        import ctypes
        n = ctypes.c_int()
        lib = importlib.import_module("custom-go-client-benchmark_b200")._lib.load()
        visible = n.value if lib.b200_device_count(ctypes.byref(n)) == 0 else 0
    except Exception:
        visible = 0
    if world <= 1 or visible <= world or visible % 2:
        return local_rank
    return (local_rank % 2) * (visible // 2) + local_rank // 2


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = pick_device(int(os.environ.get("LOCAL_RANK", "0")), world)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as d
        torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        # a rank that dies must fail the others within minutes, not hold the box until the job limit
        d.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(minutes=5))
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


def small_objects_side(crc32c, ctx, oracle, d_arena, h_arena, total, reps=20):
    """the row kernel's shapes — batches of many small or ragged objects resident in HBM — timed on
    the device beside the headline (same arena), each with a sample of its objects checked against the oracle.
    Never fails the bench."""
    try:
        rng = np.random.default_rng(SEED)
        peak, _ = measured_peaks()
        shapes = []
        n = 1 << 20
        shapes.append(("1Mi x 4 KiB", np.arange(n, dtype=np.uint64) * np.uint64(4096), np.full(n, 4096, dtype=np.uint64)))
        shapes.append(("1Mi x U[0,8 KiB), unaligned",
                       np.arange(n, dtype=np.uint64) * np.uint64(8192) + rng.integers(0, 16, size=n).astype(np.uint64),
                       rng.integers(0, 8192, size=n).astype(np.uint64)))
        ln = rng.integers(1, 4 << 20, size=4096).astype(np.uint64)
        off = np.concatenate([[0], np.cumsum(ln[:-1] + rng.integers(0, 64, size=4095).astype(np.uint64))]).astype(np.uint64)
        keep = off + ln <= np.uint64(total)
        shapes.append(("ragged U[1,4 MiB], unaligned", off[keep], ln[keep]))
        out = {}
        for name, off, ln in shapes:
            b = crc32c.DeviceBatch(off, ln, None, ctx)
            for _ in range(3):
                b.run(d_arena)
            got = b.results()[0].copy()
            ctx.timer_start()
            for _ in range(reps):
                b.run_async(d_arena)
            ms = ctx.timer_stop() / reps
            launches = b.launches()
            b.close()
            idx = rng.choice(len(ln), size=min(len(ln), 2048 if len(ln) > 5000 else 64), replace=False)
            cpu, _, _ = oracle.verify_objects(h_arena, off[idx], ln[idx], None, threads=os.cpu_count() or 1)
            nbytes = int(ln.sum())
            gbs = nbytes / (ms * 1e-3) / 1e9
            out[name] = {"objects": int(len(ln)), "bytes": nbytes, "ms": round(ms, 4), "GB/s": round(gbs, 1),
                         "frac_of_hbm_peak": round(gbs / peak, 4), "launches": launches,
                         "oracle_ok": bool((cpu == got[idx]).all()), "oracle_sample": int(len(idx))}
        return out
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"[:300]}


def text_codec_side(pkg, ctx, nlines=20_000_000, reps=5):
    """The text form of the samples (Go Duration.String() lines, what the README pipes to the histogram script)
    parsed on the device: lines, bytes, ms and GB/s, the first block of values against Python's float()."""
    try:
        import ctypes as C
        lat = pkg.latency
        ns = np.random.default_rng(SEED + 1).integers(20_000_000, 95_000_000, size=200_000)
        lines = [lat.go_duration_line(int(v)) for v in ns]
        chunk = ("\n".join(lines) + "\n").encode()
        data = np.frombuffer(chunk * (nlines // len(lines)), dtype=np.uint8)
        dt = ctx.alloc_device(data.size)
        ctx.h2d(dt, data)
        dv = ctx.alloc_device(nlines * 8)
        dst = ctx.alloc_device(nlines)
        o = [C.c_uint64() for _ in range(4)]
        call = lambda: pkg._lib.check(ctx._L.b200_parse_latency_text(ctx._h, 0, dt, data.size, dv, dst, nlines,
                                                                      *[C.byref(x) for x in o]))
        call()
        t0 = time.perf_counter()
        for _ in range(reps):
            call()
        ms = (time.perf_counter() - t0) / reps * 1e3
        head = np.empty(len(lines), dtype=np.float64)
        ctx.d2h(head, dv)
        want = np.array([float(x.strip()) for x in lines])
        for p in (dt, dv, dst):
            ctx.free_device(p)
        return {"lines": int(o[0].value), "bytes": int(data.size), "ms": round(ms, 3),
                "GB/s": round(data.size / (ms * 1e-3) / 1e9, 1), "Mlines/s": round(o[0].value / ms / 1e3, 1),
                "bad_lines": int(o[2].value), "first_block_equals_python_float": bool((head == want).all()),
                "timing": "wall clock around the C-ABI call (ends with a device->host copy of the 4-word summary)"}
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"[:300]}


GOLDEN = os.path.join(ROOT, "tests", "golden")


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


class Ranks:
    """The few collectives the bench itself needs (timing reductions, gathers), over NCCL."""

    def __init__(self, dist, rank, world, local):
        self.dist, self.rank, self.world, self.local = dist, rank, world, local

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def reduce(self, values, op="MAX"):
        if self.dist is None:
            return [float(v) for v in values]
        import torch
        t = torch.tensor([float(v) for v in values], dtype=torch.float64, device=f"cuda:{self.local}")
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return [float(x) for x in t]

    def gather_i64(self, values):
        """every rank's int64 vector, on every rank: [world][len(values)]"""
        if self.dist is None:
            return [[int(v) for v in values]]
        import torch
        t = torch.tensor([int(v) for v in values], dtype=torch.int64, device=f"cuda:{self.local}")
        out = [torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [[int(x) for x in o.cpu()] for o in out]


def multi_device_parity(rk):
    """Rank 0, before anything is timed: b200_crc32c_sharded / b200_hist_fixed_multi / b200_select_kth_multi on
    distinct GPUs against the oracle, whenever the box shows more than one (the test box has one GPU)."""
    out = None
    if rk.rank == 0:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import multidev_checks
            nd = min(multidev_checks.ngpu(), 4)
            out = "skipped: one GPU visible" if nd < 2 else dict(multidev_checks.multi_device_parity(nd), result="ok")
        except AssertionError as e:
            out = f"FAILED: {e}"
    rk.barrier()
    return out


def config3_block(rk, crc32c, ctx, oracle_mod):
    """ONE 64 GiB object sharded over the ranks (strong scaling), host-side GF(2) fold; checked against the golden."""
    g = golden("config3.json")
    L, W, r = int(g["bytes"]), rk.world, rk.rank
    lo, hi = L * r // W // 8 * 8, L * (r + 1) // W // 8 * 8
    n = hi - lo
    d = ctx.alloc_device(n)
    ctx.fill_payload(d, n // 8, SEED, first_word=lo // 8)
    crc = crc32c.UpdateDevice(0, d, n, ctx)  # warm-up, and this shard's value
    reps = 5
    rk.barrier()
    ctx.timer_start()
    for _ in range(reps):
        assert crc32c.UpdateDevice(0, d, n, ctx) == crc
    ms = ctx.timer_stop() / reps
    # the whole operation as a caller sees it: shard CRC, all-gather of (crc, len), fold
    rk.barrier()
    t0 = time.perf_counter()
    c = crc32c.UpdateDevice(0, d, n, ctx)
    parts = rk.gather_i64([c, n])
    total = 0
    for pc, pn in parts:
        total = crc32c.Combine(total, pc, pn)
    op_ms = (time.perf_counter() - t0) * 1e3
    # this rank's shard against the golden's 1 GiB pieces (CPU-made), and one piece regenerated by the oracle here
    piece = int(g["piece_bytes"])
    shard_ok = 1
    if lo % piece == 0 and hi % piece == 0:
        acc = 0
        for pc in g["piece_crc32c"][lo // piece:hi // piece]:
            acc = crc32c.Combine(acc, pc, piece)
        shard_ok = int(acc == crc)
    chk = 64 << 20
    spot_ok = int(crc32c.UpdateDevice(0, d + n - chk, chk, ctx) ==
                  oracle_mod.crc32c(oracle_mod.payload(chk, seed=SEED, first_word=(hi - chk) // 8)))
    ctx.free_device(d)
    ms, op_ms = rk.reduce([ms, op_ms], "MAX")
    shard_ok, spot_ok = rk.reduce([shard_ok, spot_ok], "MIN")
    out = {"workload": "config 3: one 64 GiB object (splitmix64 stream) sharded over the ranks, per-GPU CRCs folded "
                       "with the GF(2) combine on the host; no GPU<->GPU exchange", "scaling": "strong",
           "bytes": L, "n_gpus": W, "crc32c": f"{total:#010x}", "golden": g["crc32c_hex"],
           "golden_ok": bool(total == int(g["crc32c"]) and shard_ok and spot_ok),
           "shards_match_golden_pieces": bool(shard_ok), "oracle_spot_check_per_rank": bool(spot_ok),
           "ms": round(ms, 4), "GB/s": round(L / ms / 1e6, 1),
           "timing": f"CUDA events around {reps} b200_crc32c_update_device calls on the shard, max over ranks",
           "operation_ms": round(op_ms, 3),
           "operation": "wall clock: shard CRC + all_gather of (crc, len) + host fold, max over ranks"}
    # the same object on ONE GPU in the same run (rank 0; the others wait), for the strong-scaling efficiency
    if W > 1:
        n1 = None
        if r == 0:
            d = ctx.alloc_device(L)
            ctx.fill_payload(d, L // 8, SEED, first_word=0)
            whole = crc32c.UpdateDevice(0, d, L, ctx)
            ctx.timer_start()
            for _ in range(3):
                crc32c.UpdateDevice(0, d, L, ctx)
            n1 = ctx.timer_stop() / 3
            ctx.free_device(d)
            out["golden_ok"] = bool(out["golden_ok"] and whole == int(g["crc32c"]))
        rk.barrier()
        if r == 0:
            out["n1_ms_same_run"] = round(n1, 4)
            out["efficiency_vs_n1"] = round(n1 / (W * ms), 4)
    else:
        out["efficiency_vs_n1"] = 1.0
    return out


def config5_block(rk, pkg, ctx):
    """10^9 latency samples split over the ranks; bin counts / bracket counts all-reduced over NCCL."""
    lat = pkg.latency
    g = golden("config5.json")
    N, W, r = int(g["samples"]), rk.world, rk.rank
    lo, hi = N * r // W, N * (r + 1) // W
    n = hi - lo
    d = ctx.alloc_device(n * 8)
    ctx.fill_latency_ns(d, n, SEED, first=lo)
    rd = f"cuda:{rk.local}" if rk.dist is not None else None
    kw = dict(ctx=ctx, reduce_device=rd)
    calls = {"readme_histogram": lambda: lat.hist_readme(d, n, **kw),
             "opencensus_distribution": lambda: lat.hist_opencensus(d, n, **kw),
             "readme_and_opencensus_one_pass": lambda: lat.hist_readme_and_opencensus(d, n, **kw),
             "ssd_test_seven_lines": lambda: lat.percentiles(d, n, lat.NS_PER_US, **kw)}
    res, timing = {}, {}
    reps = 3
    for name, f in calls.items():
        res[name] = f()  # warm-up: sizes the scratch buffers
        rk.barrier()
        tally0 = dict(lat._Reducer.tally)
        t0 = time.perf_counter()
        for _ in range(reps):
            f()
        dt = (time.perf_counter() - t0) / reps * 1e3
        rk.barrier()
        ms, coll_ms = rk.reduce([dt, (lat._Reducer.tally["seconds"] - tally0["seconds"]) / reps * 1e3], "MAX")
        timing[name] = {"ms": round(ms, 3), "GB/s": round(N * 8 / ms / 1e6, 1),
                        "collectives": (lat._Reducer.tally["collectives"] - tally0["collectives"]) // reps,
                        "collectives_ms": round(coll_ms, 3)}
    ctx.free_device(d)
    h, oc, (h2, oc2), p = (res[k] for k in calls)
    gm, gu = g["ms"], g["us"]
    ok = ([int(x) for x in h.counts] == gm["readme15_counts"] and h.dropped_below == gm["dropped_below"] and
          h.dropped_above == gm["dropped_above"] and [int(x) for x in oc.counts] == gm["oc35_counts"] and
          [int(x) for x in h2.counts] == gm["readme15_counts"] and [int(x) for x in oc2.counts] == gm["oc35_counts"] and
          (h.count, h.min, h.max, h.sum, h.sumsq) == (N, gm["min"], gm["max"], gm["sum"], gm["sumsq"]) and
          (oc.count, oc.sum) == (N, gm["sum"]) and
          (p.avg, p.p20, p.p50, p.p90, p.p99, p.min, p.max) ==
          (gu["avg"], gu["p20"], gu["p50"], gu["p90"], gu["p99"], gu["min"], gu["max"]))
    return {"workload": "config 5: 10^9 synthetic latency samples split over the ranks: README.md:21-32 histogram, "
                        "metrics_exporter.go:28 distribution, ssd_test/main.go:147-163 lines", "scaling": "strong",
            "samples": N, "n_gpus": W, "bytes_per_sample": 8, "golden_ok": bool(ok),
            "reduction": "NCCL all-reduce of bin counts / moments / bracket ends and counts" if W > 1 else "none (one rank)",
            "timing": f"wall clock around {reps} synchronous calls (each ends with a device->host copy of its result "
                      "and, with several ranks, its all-reduces), max over ranks",
            "calls": timing,
            "us": {"avg": p.avg, "p20": p.p20, "p50": p.p50, "p90": p.p90, "p99": p.p99, "min": p.min, "max": p.max}}


def run_b200(args):
    rank, world, local, dist = dist_setup(args.gpus)
    rk = Ranks(dist, rank, world, local)
    affinity = bind_to_gpu_numa_node(local) if world > 1 else "single rank: not bound"
    pkg = importlib.import_module("custom-go-client-benchmark_b200")
    import oracle  # the checker (and, at N=1, the timed CPU baseline); never on the GPU path
    crc32c = pkg.crc32c
    ctx = crc32c.Context([local])
    multidev = multi_device_parity(rk)
    total = NOBJ * OBJ_BYTES
    d_arena = ctx.alloc_device(total)
    # rank r owns words [r*total/8, (r+1)*total/8) of the splitmix stream
    ctx.fill_payload(d_arena, total // 8, SEED, first_word=rank * (total // 8))
    off = np.arange(NOBJ, dtype=np.uint64) * OBJ_BYTES
    ln = np.full(NOBJ, OBJ_BYTES, dtype=np.uint64)

    # expected checksums: first pass on the GPU; the timed steps then VERIFY against them.  Every rank checks a
    # sample of ITS objects against the oracle on regenerated payload (rank 0 at N=1 checks all of them below).
    b0 = crc32c.DeviceBatch(off, ln, None, ctx)
    b0.run(d_arena)
    want, _, _ = b0.results()
    b0.close()
    probe = [0, 1, 1234, NOBJ // 2, NOBJ - 1] + [int(x) for x in np.random.default_rng(rank).integers(0, NOBJ, 11)]
    rank_ok = 1
    for i in probe:
        ref = oracle.crc32c(oracle.payload(OBJ_BYTES, seed=SEED, first_word=(rank * total + i * OBJ_BYTES) // 8))
        rank_ok &= int(ref == int(want[i]))
    rank_ok = int(rk.reduce([rank_ok], "MIN")[0])
    assert rank_ok, "GPU CRCs differ from the CPU oracle on some rank"
    batch = crc32c.DeviceBatch(off, ln, want, ctx)

    for _ in range(max(args.warmup, 0)):
        batch.run(d_arena)
    batch.kernel_ms()  # reset the per-run event window
    rk.barrier()
    with ClockSampler(local) as clk:
        ctx.timer_start()
        for _ in range(args.steps):
            batch.run_async(d_arena)
        elapsed_ms = ctx.timer_stop()
    rk.barrier()
    got, ok, nbad = batch.results()
    assert nbad == 0 and (got == want).all(), "verification failed inside the timed region"
    _, rows_ms = batch.kernel_ms()
    launches = batch.launches() * args.steps
    ms_per_step, rows_ms = rk.reduce([elapsed_ms / args.steps, rows_ms], "MAX")
    value = world * total / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: host buffers through the C-ABI VerifyObjects entry point ---------------------------------
    h_arena = ctx.alloc_pinned(total)
    ctx.d2h(h_arena, d_arena)
    # the ceiling first: a bare cudaMemcpy of the same pinned arena (no kernels), all ranks at once
    ctx.h2d(d_arena, h_arena)
    rk.barrier()
    t0 = time.perf_counter()
    for _ in range(2):
        ctx.h2d(d_arena, h_arena)
    h2d_ms = (time.perf_counter() - t0) / 2 * 1e3
    rk.barrier()
    bufs = h_arena.reshape(NOBJ, OBJ_BYTES)  # one pinned 2 MiB buffer per object, as the read loop holds them
    e2e_steps = max(1, min(args.steps, 5))
    bad, g = crc32c.VerifyObjects(bufs, want, ctx)  # warm-up (allocates the slabs)
    assert bad == [] and (g == want).all()
    rk.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        bad, g = crc32c.VerifyObjects(bufs, want, ctx)
    e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
    rk.barrier()
    assert bad == []
    # the same objects visited in shuffled order, so that no two neighbours are contiguous in host memory and
    # every 2 MiB buffer is its own copy (a caller holding 4096 separately allocated buffers)
    perm = np.random.default_rng(7).permutation(NOBJ)
    ptrs = (h_arena.ctypes.data + perm.astype(np.uint64) * np.uint64(OBJ_BYTES)).astype(np.uint64)
    bad, g = crc32c.VerifyObjectsAt(ptrs, ln, want[perm], ctx)
    assert bad == [] and (g == want[perm]).all()
    rk.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        bad, g = crc32c.VerifyObjectsAt(ptrs, ln, want[perm], ctx)
    sc_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
    rk.barrier()
    assert bad == []
    h2d_own_gbs = total / (h2d_ms * 1e-3) / 1e9
    e2e_ms, sc_ms, h2d_ms = rk.reduce([e2e_ms, sc_ms, h2d_ms], "MAX")
    h2d_per_rank = [x[0] / 1e3 for x in rk.gather_i64([int(h2d_own_gbs * 1e3)])]
    e2e_val = world * total / (e2e_ms * 1e-3) / 1e9
    h2d_val = world * total / (h2d_ms * 1e-3) / 1e9

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
                        "h2d_only": {"value": round(h2d_val, 2), "unit": UNIT, "per_rank": [round(x, 2) for x in h2d_per_rank],
                                     "what": "bare cudaMemcpy host->device of the same pinned 8 GiB arena, no kernels, "
                                             "every rank at once (the PCIe / host-memory ceiling of this path)"},
                        "frac_of_h2d": round(e2e_val / h2d_val, 4),
                        "scattered": {"value": round(world * total / (sc_ms * 1e-3) / 1e9, 2), "unit": UNIT,
                                      "ms_per_step": round(sc_ms, 3),
                                      "what": "the same 4096 buffers in shuffled order: one cudaMemcpyAsync per 2 MiB "
                                              "buffer, no merging of host-contiguous neighbours"},
                        "pcie_note": "PCIe Gen5 x16 bound (~64 GB/s raw per direction)", "cpu_affinity": affinity,
                        "cuda_device_of_rank0": local,
                        "placement": "ranks dealt alternately to the two halves of the visible GPUs (pick_device)"},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                             "frac": round(achieved / peak, 4), "traffic": traffic,
                             "frac_of_nominal_8TBps": round(achieved / 8000.0, 4), "kernel": "crc_fold_kernel",
                             "kernel_ms": round(rows_ms, 5), "algorithmic_bytes_per_launch": total,
                             "peak_source": peak_src + " — of measured"},
                "oracle_check": f"every rank: {len(probe)} of its objects against the CPU oracle on regenerated payload"
                                + ("; rank 0: all 4096" if world == 1 else ""),
                "multi_device_parity": multidev if isinstance(multidev, str) else multidev.get("result"),
                "multi_device_parity_detail": multidev}
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
    small = small_objects_side(crc32c, ctx, oracle, d_arena, h_arena, total) if world == 1 else None
    ctx.free_pinned(h_arena)
    ctx.free_device(d_arena)
    batch.close()
    side = {}
    for name, f in (("config3", lambda: config3_block(rk, crc32c, ctx, oracle)),
                    ("config5", lambda: config5_block(rk, pkg, ctx))):
        try:
            side[name] = f()
        except Exception as e:  # a side block must not cost the headline line (every rank fails alike or not at all)
            side[name] = {"error": f"{type(e).__name__}: {e}"[:300], "golden_ok": False}
    if line is not None:
        line.update(side)
        if world == 1:
            line["latency_reduction"] = latency_side_measurement(pkg, ctx)
            line["small_objects"] = small
            line["text_codec"] = text_codec_side(pkg, ctx)
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
