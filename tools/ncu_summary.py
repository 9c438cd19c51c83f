"""Summarise an .ncu-rep (raw page) into one line per launch: usage python tools/ncu_summary.py rep [out.md]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H = rows[0]
want = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp_inst"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem%"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block")]
idx = {h: i for i, h in enumerate(H)}
lines = []
hdr = "| kernel | " + " | ".join(n for _, n in want) + " |"
lines.append(hdr)
lines.append("|" + "---|" * (len(want) + 1))
for r in rows[2:]:
    if len(r) < len(H):
        continue
    name = r[idx["Kernel Name"]].split("(")[0]
    vals = []
    for m, _ in want:
        if m in idx:
            u = rows[1][idx[m]]
            v = r[idx[m]]
            try:
                v = "%.1f" % float(v) if "." in v else v
            except ValueError:
                pass
            u = {"register/thread": "", "inst": ""}.get(u, u)
            vals.append(v + (" " + u if u and u not in ("%",) else ""))
        else:
            vals.append("-")
    lines.append("| " + name + " | " + " | ".join(vals) + " |")
out = "\n".join(lines)
print(out)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(out + "\n")
