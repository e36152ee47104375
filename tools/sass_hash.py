#!/usr/bin/env python
"""SHA-1 of the SASS text of every kernel in libzenith_b200.so (addresses and encodings stripped), keyed by the demangled
kernel name.  build.sh writes raytracer_project_b200/_build/sass_hashes.json with it; bench.py uses the hashes to tie
instruction counts taken from an ncu capture (profiles/r02_issue_counts.json) to the code that is actually running: a
kernel change invalidates the recorded count instead of silently keeping it.

    python tools/sass_hash.py [path/to/lib.so] [out.json]
"""
import hashlib
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA_BIN = os.environ.get("CUDA_BIN", "/usr/local/cuda/bin")


def sass_hashes(lib: str) -> dict:
    text = subprocess.run([os.path.join(CUDA_BIN, "cuobjdump"), "-sass", lib], check=True, capture_output=True, text=True).stdout
    funcs, cur, name = {}, [], None
    for line in text.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                funcs[name] = cur
            name, cur = m.group(1), []
            continue
        if name is None:
            continue
        ins = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?)\s*/\*", line)  # "/*0010*/  INSTR ;  /* 0x... */"
        if ins:
            cur.append(ins.group(1).strip())
    if name:
        funcs[name] = cur
    mangled = list(funcs)
    dem = subprocess.run([os.path.join(CUDA_BIN, "cu++filt")] + mangled, check=True, capture_output=True, text=True).stdout.splitlines()
    out = {}
    for mname, dname in zip(mangled, dem):
        body = "\n".join(funcs[mname]).encode()
        out[dname.strip()] = {"sha1": hashlib.sha1(body).hexdigest(), "instructions": len(funcs[mname]), "mangled": mname}
    return out


if __name__ == "__main__":
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "raytracer_project_b200", "libzenith_b200.so")
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "raytracer_project_b200", "_build", "sass_hashes.json")
    h = sass_hashes(lib)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w") as f:
        json.dump(h, f, indent=0, sort_keys=True)
    print(f"{len(h)} kernels -> {out}")
