import sys, time, os
sys.path.insert(0, '/root/repo')
import torch, numpy as np
import bench
from raytracer_project_b200 import api, scenes
import tempfile
txt, cfg, files = bench.make_scene('c4', spp=8)
with tempfile.TemporaryDirectory() as d:
    flat = api.flatten_zscene(path=scenes.materialize(d, txt, cfg, files))
ctx = api.Context(0)
out = np.empty((cfg.image_height*cfg.image_width, 3))
for k in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); sc = ctx.upload(flat); t1 = time.perf_counter()
    sc.render(cfg, seed=1, beauty_out=out); t2 = time.perf_counter()
    sc.close(); t3 = time.perf_counter()
    torch.cuda.synchronize(); t4 = time.perf_counter()
    ctx.group_barrier(); t5 = time.perf_counter()
    print(f"upload {1e3*(t1-t0):.2f} render {1e3*(t2-t1):.2f} close {1e3*(t3-t2):.2f} torch sync {1e3*(t4-t3):.2f} barrier {1e3*(t5-t4):.2f}", flush=True)
