#!/usr/bin/env python
"""GPU box: dump zb_primary_ids of the golden scenes to gpurun_out/ids_<name>.npz for offline diffing."""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from raytracer_project_b200 import api, scenes
from tests.golden.golden_scenes import GOLDEN_SCENES

names = sys.argv[1:] or list(GOLDEN_SCENES)
ctx = api.Context(0)
for name in names:
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    d = tempfile.mkdtemp()
    flat = api.flatten_zscene(path=scenes.materialize(d, txt, cfg, files))
    sc = ctx.upload(flat)
    leaf, prim, t = sc.primary_ids(cfg)
    np.savez_compressed(f"gpurun_out/ids_{name}.npz", leaf=leaf, prim=prim, t=t)
    sc.close()
    print(name, "done")
