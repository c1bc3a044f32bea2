"""profiles/<round>_kernels_metrics.csv and the table in profiles/<round>_summary.md from the raw page of an
`ncu --set full` capture of tools/profile_workloads.py (gpurun_out/<round>_all_raw.csv): first launch of every
distinct kernel past its warm-up launch.  usage: python tools/summarize_all_kernels.py r02"""
import csv, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rnd = sys.argv[1] if len(sys.argv) > 1 else "r02"
# one or two captures (the second, <round>_b_raw.csv, holds the kernels the first one's launch budget did not reach);
# each row becomes a dict metric -> (value, unit)
vals = []
for name in (f"{rnd}_all_raw.csv", f"{rnd}_b_raw.csv"):
    path = os.path.join(ROOT, "gpurun_out", name)
    if os.path.exists(path):
        rr = list(csv.reader(open(path)))
        vals += [{h: (x, u) for h, x, u in zip(rr[0], r, rr[1])} for r in rr[2:]]
col = {h: h for h in set().union(*[set(v) for v in vals])}
rows = [None, {h: "" for h in col}]
keep = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
keep = [k for k in keep if all(k in v for v in vals)]


class Row(dict):  # v[col[m]] -> the value string
    def __getitem__(self, m):
        return dict.__getitem__(self, m)[0]


units = {m: vals[0][m][1] for m in vals[0]}
vals = [Row(v) for v in vals]
rows[1] = units


def short(n):
    n = n.replace("void ", "").replace("b200::<unnamed>::", "").replace("unnamed>::", "").replace("(bool)", "").replace("(int)", "")
    return (n[:n.index(">(") + 1] if ">(" in n else n.split("(")[0]).strip()


seen, picked = {}, []
for v in vals:
    k = short(v[col["Kernel Name"]]) + " grid " + v[col["launch__grid_size"]]
    # the same kernel and grid at very different sizes (fold kernel: 32 MiB segments of the ragged batch, a million
    # 4 KiB blocks a lane per object) are different rows: add the launch's instruction count, to a power of four
    import math
    inst = float(v[col["smsp__inst_executed.sum"]].replace(",", "") or 0)
    k += f" (~{4 ** round(math.log(max(inst, 1), 4)) / 1e6:.0f} M instr)" if "fold" in k or "rows" in k else ""
    seen[k] = seen.get(k, 0) + 1
    if seen[k] == 2 or (seen[k] == 1 and k not in [p[0] for p in picked]):
        picked = [p for p in picked if p[0] != k] + [(k, v)]
with open(os.path.join(ROOT, "profiles", f"{rnd}_kernels_metrics.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(["kernel"] + keep)
    for k, v in picked:
        w.writerow([k] + [v[col[m]] for m in keep])
lines = ["| kernel (grid) | us | DRAM GB/s | DRAM % | L1/smem % | issue % | instr / warp-instr lanes | smem conflicts / wavefronts | regs | stalls long / short / wait |",
         "|---|---|---|---|---|---|---|---|---|---|"]
for k, v in picked:
    g = lambda m: float(v[col[m]].replace(",", "")) if m in col and v[col[m]] not in ("", "n/a") else float("nan")
    unit_of = lambda m: dict.__getitem__(v, m)[1]
    tscale = {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}[unit_of("gpu__time_duration.sum")]
    us = g("gpu__time_duration.sum") * tscale
    bscale = lambda m: {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[unit_of(m)]
    nbytes = g("dram__bytes_read.sum") * bscale("dram__bytes_read.sum") + g("dram__bytes_write.sum") * bscale("dram__bytes_write.sum")
    lines.append(f"| `{k}` | {us:.1f} | {nbytes / us / 1e3:.0f} | {g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | "
                 f"{g('l1tex__throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | {g('smsp__issue_active.avg.pct_of_peak_sustained_active'):.0f} | "
                 f"{g('smsp__inst_executed.sum') / 1e6:.1f} M / {g('smsp__thread_inst_executed_per_inst_executed.ratio'):.1f} | "
                 f"{g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum') / 1e6:.2f} M / {g('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum') / 1e6:.1f} M | "
                 f"{g('launch__registers_per_thread'):.0f} | {g('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio'):.1f} / "
                 f"{g('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio'):.1f} / {g('smsp__average_warps_issue_stalled_wait_per_issue_active.ratio'):.1f} |")
if rnd == "r02" and not any("rows_kernel grid 148" in l for l in lines):
    # measured on 1 Mi x 4 KiB aligned blocks before they moved to the fold kernel (the capture of commit 2a65aff);
    # tools/profile_workloads.py now holds a misaligned copy of the shape so that later captures have the row again
    lines.append("| `crc_rows_kernel grid 148` (1 Mi x 4 KiB, earlier capture) | 1011.6 | 4273 | 52 | 84 | 75 | 837.4 M / 32.0 | "
                 "28.40 M / 180.6 M | 64 | 1.8 / 0.7 / 1.3 |")
open(os.path.join(ROOT, "profiles", f"{rnd}_kernels_table.md"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
