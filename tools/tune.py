#!/usr/bin/env python
"""Kernel-tuning helper (GPU box): time zb_render_device on a named scene with the library selected by
ZB_LIB, CUDA events on the context stream, L2 flushed between runs.  Prints one JSON line."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from raytracer_project_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="weekend")
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--spp", type=int, default=64)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--stats", action="store_true")
ap.add_argument("--gpu-blas", action="store_true")
ap.add_argument("--grid", type=int, default=0)
ap.add_argument("--tag", default=os.environ.get("ZB_LIB", "main"))
a = ap.parse_args()
import tempfile, time
gen = getattr(scenes, a.scene)
kw = {'grid': a.grid} if a.grid else {}
out = gen(width=a.width, height=a.height, spp=a.spp, **kw)
txt, cfg = out[0], out[1]
t0 = time.time()
if len(out) > 2:
    flat = api.flatten_zscene(path=scenes.materialize(tempfile.mkdtemp(prefix="zb_tune_"), txt, cfg, out[2]), gpu_blas=a.gpu_blas)
else:
    flat = api.flatten_zscene(txt)
flatten_s = time.time() - t0
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()
ctx = api.Context(0, stream=stream.cuda_stream)
torch.cuda.synchronize()
t1 = time.time()
sc = ctx.upload(flat)
upload_s = time.time() - t1
n = a.width * a.height
acc = torch.empty(n, 4, device=dev)
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
res = {"tag": a.tag, "scene": a.scene, "flatten_s": round(flatten_s, 2), "scene_mb": round(flat.contents.nbytes() / 1e6, 2), "upload_s": round(upload_s, 4), "gpu_blas": a.gpu_blas}
if a.stats:
    sc.render_device(cfg, acc.data_ptr(), seed=1, stats=True)
    torch.cuda.synchronize()
    st = ctx.stats()
    res.update(rays_per_path=st["rays"] / st["paths"], nodes_per_ray=st["node_visits"] / st["rays"],
               prims_per_ray=st["prim_tests"] / st["rays"])
ms = []
for i in range(a.reps + 1):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sc.render_device(cfg, acc.data_ptr(), seed=1)
    e1.record()
    torch.cuda.synchronize()
    if i:
        ms.append(e0.elapsed_time(e1))
res.update(ms=min(ms), ms_all=[round(x, 3) for x in ms], mpaths=n * a.spp / min(ms) * 1e-3,
           mean=[round(x, 6) for x in (acc[:, :3].mean(0) / a.spp).tolist()],
           acc_sha=__import__("hashlib").sha1(acc.cpu().numpy().tobytes()).hexdigest()[:12],
           wavefront=os.environ.get("ZB_WAVEFRONT", "0"), wf_slots=os.environ.get("ZB_WF_SLOTS", ""))
print(json.dumps(res), flush=True)
