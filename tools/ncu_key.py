#!/usr/bin/env python
"""Key counters of EVERY kernel launch in an ncu --set full report, as a text table kept under profiles/.

    python tools/ncu_key.py REPORT.ncu-rep OUT.txt [name-regex]
"""
import csv
import io
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"]


def main():
    rep, out = sys.argv[1:3]
    pat = re.compile(sys.argv[3]) if len(sys.argv) > 3 else None
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    lines = [f"# {rep.split('/')[-1]}: ncu --set full --clock-control none, one capture per launch (cold caches, serialised)"]
    for vals in rows[2:]:
        name = vals[col["Kernel Name"]]
        if pat and not pat.search(name):
            continue
        lines.append("")
        lines.append(f"## launch {vals[col['ID']]}: {name}")
        for h in hdr:
            if h in KEYS or ("issue_stalled" in h and h.endswith("per_issue_active.ratio") and float(vals[col[h]].replace(",", "") or 0) >= 0.2):
                lines.append(f"{h} [{units[col[h]]}] {vals[col[h]]}")
    open(out, "w").write("\n".join(lines) + "\n")
    print(len(lines), "lines ->", out)


if __name__ == "__main__":
    main()
