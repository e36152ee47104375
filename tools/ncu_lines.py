#!/usr/bin/env python
"""Per-source-line view of an ncu capture (CPU box).

ncu's own source page loses the file/line correlation for heavily inlined device code, so this joins
  * `ncu -i REP --page source --csv --print-source sass`  (per-SASS-instruction counters), with
  * `nvdisasm -g CUBIN`                                   (SASS with //## File "...", line N markers)
by instruction offset, and aggregates warp instructions, thread instructions, active-lane average and
stall samples per source line (innermost inlined location).

  python tools/ncu_lines.py REP.ncu-rep LIB.so 'k_renderILb0' [--top 40]
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import tempfile

ap = argparse.ArgumentParser()
ap.add_argument("rep")
ap.add_argument("lib")
ap.add_argument("func", help="substring of the mangled kernel name")
ap.add_argument("--top", type=int, default=40)
ap.add_argument("--by", default="inst", choices=["inst", "samples"])
ap.add_argument("--launch", type=int, default=None, help="index of the launch inside a report that holds several")
a = ap.parse_args()

tmp = tempfile.mkdtemp(prefix="ncul_")
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(a.lib)], cwd=tmp, check=True, capture_output=True)
lines_of = {}  # offset -> (file, line)
for cub in sorted(os.listdir(tmp)):
    txt = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    cur_fn, loc = None, ("?", 0)
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            cur_fn = m.group(1)
            continue
        if cur_fn is None or a.func not in cur_fn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            loc = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            lines_of[int(m.group(1), 16)] = (loc, m.group(2).strip())
    if lines_of:
        break

sel = [] if a.launch is None else ["--launch-skip", str(a.launch), "--launch-count", "1"]
out = subprocess.run(["ncu", "-i", a.rep, *sel, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {n: i for i, n in enumerate(hdr)}
base = None
agg = collections.defaultdict(lambda: [0, 0, 0])  # inst, thread inst, samples
tot = [0, 0, 0]
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    addr = int(r[0], 16)
    if base is None:
        base = addr
    loc, _ = lines_of.get(addr - base, (("?", 0), ""))
    v = [int(r[col["Instructions Executed"]] or 0), int(r[col["Thread Instructions Executed"]] or 0),
         int(r[col["# Samples"]] or 0)]
    for k in range(3):
        agg[loc][k] += v[k]
        tot[k] += v[k]
key = 0 if a.by == "inst" else 2
print(f"total warp-inst {tot[0]:.3e}  thread-inst {tot[1]:.3e}  avg lanes {tot[1] / max(tot[0], 1):.2f}  samples {tot[2]}")
print(f"{'file:line':28s} {'inst%':>6s} {'lanes':>6s} {'smp%':>6s}")
for loc, v in sorted(agg.items(), key=lambda kv: -kv[1][key])[:a.top]:
    print(f"{loc[0] + ':' + str(loc[1]):28s} {100 * v[0] / tot[0]:6.2f} {v[1] / max(v[0], 1):6.1f} {100 * v[2] / max(tot[2], 1):6.2f}")
