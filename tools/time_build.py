#!/usr/bin/env python
"""BLAS build times (GPU box): host binned SAH inside the flattener against the two device builders inside
zb_scene_upload, on bench.py's mesh configs.  One JSON line per config."""
import argparse, json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from raytracer_project_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--configs", default="c3")
a = ap.parse_args()
ctx = api.Context(0)
for name in a.configs.split(","):
    txt, cfg, files = bench.make_scene(name)
    with tempfile.TemporaryDirectory() as d:
        path = scenes.materialize(d, txt, cfg, files)
        t0 = time.perf_counter(); flat_dev = api.flatten_zscene(path=path, gpu_blas=True); t1 = time.perf_counter()
        flat_host = api.flatten_zscene(path=path); t2 = time.perf_counter()
    tris = int(flat_dev.contents.n_tris)
    out = {"config": name, "triangles": tris, "flatten_without_blas_s": round(t1 - t0, 3), "flatten_with_host_sah_s": round(t2 - t1, 3),
           "host_sah_build_s": round((t2 - t1) - (t1 - t0), 3)}
    def upload_ms(flat, reps=3):
        best = 1e30
        for _ in range(reps):
            torch.cuda.synchronize(); t = time.perf_counter()
            sc = ctx.upload(flat); torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t) * 1e3)
            sc.close()
        return round(best, 2)
    out["upload_host_tree_ms"] = upload_ms(flat_host)
    for b, key in ((1, "upload_device_sah_ms"), (0, "upload_device_lbvh_ms")):
        ctx.set_tuning("gpu_builder", b)
        out[key] = upload_ms(flat_dev)
    ctx.set_tuning("gpu_builder", 1)
    # upload of the same flat scene with the builds taken out = H2D of triangles + the rest
    out["note"] = "device build time = upload_device_* minus the H2D part both share (see upload_host_tree_ms, which also copies the host nodes)"
    api.free_flat(flat_dev); api.free_flat(flat_host)
    print(json.dumps(out), flush=True)
