#!/usr/bin/env python
"""Time zb_render_device in its three modes (beauty, beauty + AOVs, light passes) on a named scene (GPU box);
ZB_PQ=0 selects the tile-bound k_render.  Prints one JSON line."""
import argparse, json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from raytracer_project_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="weekend")
ap.add_argument("--spp", type=int, default=64)
a = ap.parse_args()
out = getattr(scenes, a.scene)(width=1920, height=1080, spp=a.spp)
txt, cfg = out[0], out[1]
flat = api.flatten_zscene(path=scenes.materialize(tempfile.mkdtemp(), txt, cfg, out[2])) if len(out) > 2 else api.flatten_zscene(txt)
ctx = api.Context(0, stream=torch.cuda.current_stream().cuda_stream)
sc = ctx.upload(flat)
n = 1920 * 1080
acc = torch.empty(n, 4, device="cuda")
aux = torch.empty(3 * n, 4, device="cuda")
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = True
res = {"scene": a.scene, "spp": a.spp, "pq": os.environ.get("ZB_PQ", "1")}
for name, kw in (("beauty", {}), ("aov", {"aux_ptr": aux.data_ptr()}), ("light", {"aux_ptr": aux.data_ptr(), "light_pass": True})):
    ms = []
    for i in range(4):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sc.render_device(cfg, acc.data_ptr(), seed=1, **kw)
        e1.record()
        torch.cuda.synchronize()
        if i:
            ms.append(e0.elapsed_time(e1))
    res[name + "_ms"] = round(min(ms), 3)
print(json.dumps(res))
