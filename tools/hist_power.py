"""SM clock and board power while the README histogram runs back to back for a few seconds (NVML sampled in a
thread): python tools/hist_power.py [nsamples] [seconds]"""
import ctypes as C, importlib, os, sys, threading, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
crc32c, lat = pkg.crc32c, pkg.latency
import pynvml
n = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000_000
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
ctx = crc32c.Context([0])
d = ctx.alloc_device(n * 8)
ctx.fill_latency_ns(d, n, 0xC0FFEE)
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
L, hc = ctx._L, ctx._h
edges = lat.README_EDGES
counts = np.zeros(40, dtype=np.uint64)
st = pkg._lib.LatStats()
mn, mx = C.c_int64(), C.c_int64()
cases = {"hist+moments": lambda: L.b200_hist_fixed(hc, 0, d, n, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, C.byref(st)),
         "hist only": lambda: L.b200_hist_fixed(hc, 0, d, n, 1000000, edges.ctypes.data, edges.size, 0, counts.ctypes.data, None),
         "minmax": lambda: L.b200_minmax(hc, 0, d, n, C.byref(mn), C.byref(mx))}
for name, f in cases.items():
    clk, pw, stop = [], [], threading.Event()

    def sample():
        while not stop.is_set():
            clk.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
            pw.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000)
            stop.wait(0.01)
    t = threading.Thread(target=sample); t.start()
    f(); reps = 0; t0 = time.perf_counter()
    while time.perf_counter() - t0 < secs:
        f(); reps += 1
    dt = time.perf_counter() - t0
    stop.set(); t.join()
    half = len(clk) // 2
    print(f"{name:14s} {dt/reps*1e3:7.3f} ms  {n*8*reps/dt/1e9:7.1f} GB/s   SM MHz median(2nd half) {np.median(clk[half:]):6.0f}  min {min(clk)}  power W median {np.median(pw[half:]):6.0f} max {max(pw):6.0f}")
