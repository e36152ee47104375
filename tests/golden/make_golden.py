#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/ref_harness, built by oracle/Makefile from /root/reference) on the seeded scenes of
raytracer_project_b200/scenes.py.  Needs /root/reference; run in the build container only:

    python tests/golden/make_golden.py [name ...]

The reference's RNG is unseedable (common.hpp:29-34), so radiance fixtures are statistical:
per-pixel mean and variance-of-the-mean over independent batches.
Primary-hit id fixtures are deterministic.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from raytracer_project_b200 import scenes  # noqa: E402

HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def run_harness(args, cwd):
    out = subprocess.run([HARNESS] + args, cwd=cwd, check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def read_ids(path):
    with open(path, "rb") as f:
        w, h = np.frombuffer(f.read(8), np.int32)
        n = int(w) * int(h)
        leaf = np.frombuffer(f.read(4 * n), np.uint32).reshape(h, w)
        tri = np.frombuffer(f.read(4 * n), np.uint32).reshape(h, w)
        t = np.frombuffer(f.read(8 * n), np.float64).reshape(h, w)
    return leaf, tri, t


def read_stats(path):
    with open(path, "rb") as f:
        w, h = np.frombuffer(f.read(8), np.int32)
        n = int(w) * int(h) * 3
        mean = np.frombuffer(f.read(4 * n), np.float32).reshape(h, w, 3)
        var = np.frombuffer(f.read(4 * n), np.float32).reshape(h, w, 3)
    return mean, var


def scratch_dir():
    """cwd for the harness: camera::initialize creates ./assets/hdr_maps (camera.hpp:123-140)."""
    d = tempfile.mkdtemp(prefix="zb_ref_")
    os.makedirs(os.path.join(d, "assets", "hdr_maps"), exist_ok=True)
    return d


def make_ids(name, scene_text, cfg, extra_files=None):
    d = scratch_dir()
    sp, cp, op = (os.path.join(d, x) for x in ("s.zscene", "s.cfg", "ids.bin"))
    open(sp, "w").write(scene_text)
    open(cp, "w").write(cfg.to_text())
    for fn, data in (extra_files or {}).items():
        open(os.path.join(d, fn), "wb" if isinstance(data, bytes) else "w").write(data)
    info = run_harness(["ids", sp, cp, op], d)
    leaf, tri, t = read_ids(op)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_ids.npz"), leaf=leaf, tri=tri, t=t.astype(np.float32))
    print(name, "ids", info, "hit fraction", float((leaf != 0xFFFFFFFF).mean()))


def make_stats(name, scene_text, cfg, batches, extra_files=None):
    d = scratch_dir()
    sp, cp, op = (os.path.join(d, x) for x in ("s.zscene", "s.cfg", "stats.bin"))
    open(sp, "w").write(scene_text)
    open(cp, "w").write(cfg.to_text())
    for fn, data in (extra_files or {}).items():
        open(os.path.join(d, fn), "wb" if isinstance(data, bytes) else "w").write(data)
    info = run_harness(["stats", sp, cp, op, str(batches)], d)
    mean, var = read_stats(op)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_stats.npz"), mean=mean, var_of_mean=var,
                        spp=cfg.samples_per_pixel, batches=batches)
    print(name, "stats", info)


def weekend_ids():
    txt, cfg = scenes.weekend(seed=1, width=1200, height=675)
    make_ids("weekend_1200x675", txt, cfg)


def weekend_stats():
    # 16 x 256 spp = 4096 spp at 300x169 (SURVEY 8(c) parity procedure)
    txt, cfg = scenes.weekend(seed=1, width=300, height=169, spp=256)
    make_stats("weekend_300x169", txt, cfg, 16)


TARGETS = {"weekend_ids": weekend_ids, "weekend_stats": weekend_stats}

if __name__ == "__main__":
    names = sys.argv[1:] or list(TARGETS)
    for n in names:
        TARGETS[n]()
