"""Turns the ncu outputs a gpurun call left in gpurun_out/ into the tracked summaries under profiles/.

usage: python tools/summarize_profiles.py r01
Reads  gpurun_out/launches.csv (ncu --metrics gpu__time_duration.sum launch list of `bench.py --steps 20 --warmup 3`)
       gpurun_out/prof_top.ncu-rep (ncu --set full of the dominant kernel, same command)
Writes profiles/<round>_launches.csv, profiles/<round>_<kernel>_metrics.csv, profiles/<round>_summary.md,
       profiles/traffic.json (dram bytes per launch of the row kernel, read by bench.py for roofline.traffic)
"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rnd = sys.argv[1] if len(sys.argv) > 1 else "r01"
src = os.path.join(ROOT, "gpurun_out")
dst = os.path.join(ROOT, "profiles")
os.makedirs(dst, exist_ok=True)

# ---- launch list -------------------------------------------------------------------------------------
lines = [l for l in open(os.path.join(src, "launches.csv")) if not l.startswith("==")]
rows = list(csv.DictReader(io.StringIO("".join(lines))))
with open(os.path.join(dst, f"{rnd}_launches.csv"), "w") as f:
    f.write("id,kernel,grid,block,duration_ns\n")
    for r in rows:
        name = r["Kernel Name"].split("(")[0].split("::")[-1].split("<")[0]
        f.write(f'{r["ID"]},{name},"{r["Grid Size"]}","{r["Block Size"]}",{r["Metric Value"]}\n')
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name = r["Kernel Name"].split("(")[0].split("::")[-1].split("<")[0]
    tot[name][0] += 1
    tot[name][1] += float(r["Metric Value"])
all_ns = sum(v[1] for v in tot.values())

# ---- full-set metrics of the row kernel -----------------------------------------------------------------
raw = subprocess.run(["ncu", "-i", os.path.join(src, "prof_top.ncu-rep"), "--page", "raw", "--csv"],
                     capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rr[0], rr[1], rr[2:]
keep = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second"]
top = max(tot.items(), key=lambda kv: kv[1][1])[0]
with open(os.path.join(dst, f"{rnd}_{top}_metrics.csv"), "w") as f:
    f.write("metric,unit," + ",".join(f"launch{i}" for i in range(len(vals))) + "\n")
    for k in keep:
        if k in hdr:
            i = hdr.index(k)
            f.write(f"{k},{units[i]}," + ",".join(v[i] for v in vals) + "\n")


def metric(k, row=0):
    i = hdr.index(k)
    v, u = float(vals[row][i].replace(",", "")), units[i]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1}.get(u)
    return v * scale if scale else v


traffic = sum(metric("dram__bytes_read.sum", r) + metric("dram__bytes_write.sum", r) for r in range(len(vals))) / len(vals)
json.dump({f"{top}_dram_bytes_per_launch": int(traffic), "round": rnd,
           "source": f"profiles/{rnd}_{top}_metrics.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum)"},
          open(os.path.join(dst, "traffic.json"), "w"), indent=1)

with open(os.path.join(dst, f"{rnd}_summary.md"), "w") as f:
    f.write(f"# {rnd}: ncu summary of `python bench.py --steps 20 --warmup 3` (1 x B200)\n\n")
    f.write("Launch list (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised —\n"
            "compare SHARES, not absolutes).  First 60 launches of the process.\n\n")
    f.write("| kernel | launches | total ms | share |\n|---|---:|---:|---:|\n")
    for k, (n, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| {k} | {n} | {ns / 1e6:.3f} | {100 * ns / all_ns:.1f} % |\n")
    step = sum(v[1] / v[0] for k, v in tot.items() if k.startswith("crc_"))
    f.write(f"\nPer step (one {top} + one crc_finalize_kernel): {top} share "
            f"{100 * tot[top][1] / tot[top][0] / step:.1f} % of the step's kernel time.\n\n")
    f.write(f"{top}, `ncu --set full --clock-control none` (first of {len(vals)} launches):\n\n")
    f.write("| metric | value |\n|---|---:|\n")
    for k in keep:
        if k in hdr:
            i = hdr.index(k)
            f.write(f"| {k} | {vals[0][i]} {units[i]} |\n")
    f.write(f"\nDRAM traffic per launch: {traffic / 1e9:.4f} GB for 8.5899 GB algorithmic bytes "
            f"(ratio {traffic / 8589934592:.4f}) — no re-reads.\n")
print(open(os.path.join(dst, f"{rnd}_summary.md")).read())
