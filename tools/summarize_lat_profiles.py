"""Appends the latency-kernel section to profiles/<round>_summary.md from an ncu capture of tools/lat_timing.py.

usage: python tools/summarize_lat_profiles.py r01      (after tools/summarize_profiles.py r01)
Reads  gpurun_out/lat_prof.ncu-rep  (ncu --set full -k regex:"hist_kernel|sel_pass_kernel|radix_kernel|minmax_kernel"
       python tools/lat_timing.py 100000000)
Writes profiles/<round>_latency_kernels_metrics.csv and the "## Latency kernels" section of the summary.
"""
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rnd = sys.argv[1] if len(sys.argv) > 1 else "r01"
hdr = units = None
vals = []
for name in ("lat_prof.ncu-rep", "lat_prof2.ncu-rep"):  # the second, optional: hist_kernel launches only (all variants)
    rep = os.path.join(ROOT, "gpurun_out", name)
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    if hdr is None:
        hdr, units = rr[0], rr[1]
    assert rr[0] == hdr
    vals += rr[2:]
name_i = hdr.index("Kernel Name")


def short(n):
    n = n.split("(const")[0].split("(unsigned")[0]
    return n.replace("void ", "").replace("b200::<unnamed>::", "").replace("unnamed>::", "").replace("(bool)", "").replace("(int)", "").strip()


first = {}
for v in vals:  # first launch of each distinct kernel
    first.setdefault(short(v[name_i]), v)
keep = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic"]
with open(os.path.join(ROOT, "profiles", f"{rnd}_latency_kernels_metrics.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(["metric", "unit"] + list(first))
    for k in keep:
        if k in hdr:
            i = hdr.index(k)
            w.writerow([k, units[i]] + [v[i] for v in first.values()])


def g(v, k):
    return v[hdr.index(k)]


path = os.path.join(ROOT, "profiles", f"{rnd}_summary.md")
text = open(path).read().split("## Latency kernels")[0].rstrip() + "\n\n"
text += ("## Latency kernels (`python tools/lat_timing.py 100000000`, ncu --set full, first launch of each)\n\n"
         "10^8 int64 samples = 0.8 GB algorithmic bytes per launch (the select kernels of the sampled route read\n"
         "the 4·10^6-sample stratified sample or the ~1 % of kept samples instead).  Cold-cache and serialised:\n"
         "compare with the event-timed numbers in DESIGN.md §6 for the steady state.\n\n"
         "| kernel | duration | DRAM % | ALU % | issue % | warp-instr | regs | dyn smem |\n|---|---:|---:|---:|---:|---:|---:|---:|\n")
for n, v in first.items():
    text += (f"| {n} | {g(v, 'gpu__time_duration.sum')} {units[hdr.index('gpu__time_duration.sum')]} "
             f"| {float(g(v, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')):.1f} "
             f"| {float(g(v, 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active')):.1f} "
             f"| {float(g(v, 'smsp__issue_active.avg.pct_of_peak_sustained_active')):.1f} "
             f"| {float(g(v, 'smsp__inst_executed.sum').replace(',', '')):.0f} "
             f"| {g(v, 'launch__registers_per_thread')} "
             f"| {g(v, 'launch__shared_mem_per_block_dynamic')} {units[hdr.index('launch__shared_mem_per_block_dynamic')]} |\n")
open(path, "w").write(text)
print(text.split("## Latency kernels")[1])
