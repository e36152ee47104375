#!/usr/bin/env python
"""Record the issue-roofline inputs of one ncu --set full capture of a render kernel.

    python tools/ncu_issue.py REPORT.ncu-rep WORKLOAD_NAME OUT_KEY_FILE.txt

Reads the report here (no GPU needed), writes the key counters to OUT_KEY_FILE (kept under profiles/) and stores
{kernel, warp instructions per launch, lanes per instruction, DRAM bytes, SHA-1 of the kernel's SASS in the library that is
built right now} under WORKLOAD_NAME in profiles/r02_issue_counts.json.  bench.py prints the issue-slot roofline from that
entry only while the SASS hash of the kernel that ran still equals the stored one."""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def main():
    rep, workload, out_txt = sys.argv[1:4]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    col = {h: i for i, h in enumerate(hdr)}
    kernel = vals[col["Kernel Name"]]
    lines = [f"# {os.path.basename(rep)}: {kernel}", f"# workload {workload}; ncu --set full --clock-control none, one launch"]
    for h in hdr:
        if h in KEYS or ("issue_stalled" in h and h.endswith("per_issue_active.ratio")):
            lines.append(f"{h} [{units[col[h]]}] {vals[col[h]]}")
    open(out_txt, "w").write("\n".join(lines) + "\n")
    num = lambda k: float(vals[col[k]].replace(",", ""))
    dram = num("dram__bytes_read.sum") * UNIT[units[col["dram__bytes_read.sum"]]] + num("dram__bytes_write.sum") * UNIT[units[col["dram__bytes_write.sum"]]]
    # the ncu name "k_render_pq<8, 0, 0>(...)" -> the demangled prefix cu++filt prints
    m = re.search(r"(\w+)<(\d+), (\d+), (\d+)>", kernel)
    want = f"void {m.group(1)}<(int){m.group(2)}, (int){m.group(3)}, (unsigned int){m.group(4)}>"
    hashes = json.load(open(os.path.join(ROOT, "raytracer_project_b200", "_build", "sass_hashes.json")))
    built = [v for k, v in hashes.items() if k.startswith(want)]
    assert len(built) == 1, want
    path = os.path.join(ROOT, "profiles", "r02_issue_counts.json")
    db = json.load(open(path)) if os.path.exists(path) else {}
    db[workload] = {"kernel": want, "warp_inst_per_launch": num("smsp__inst_executed.sum"),
                    "lanes_per_instruction": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
                    "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"), "dram_bytes": dram,
                    "kernel_ms_under_ncu": num("gpu__time_duration.sum"), "sass_sha1": built[0]["sha1"],
                    "source": "profiles/" + os.path.basename(out_txt)}
    json.dump(db, open(path, "w"), indent=1, sort_keys=True)
    print(workload, db[workload])


if __name__ == "__main__":
    main()
