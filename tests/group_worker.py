"""One rank of a multi-PROCESS render group (zb_group_export / zb_group_join), used by tests/test_multi_gpu.py and
runnable by hand:  python tests/group_worker.py RANK N DEVICE DIR [partition]
Handles travel through files in DIR (the product leaves the transport to the caller: torch.distributed in bench.py).
Every rank renders collectively into ONE frame in shared memory; rank 0 then renders the same frame alone on a
plain context and writes `result.json`."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracer_project_b200 import api, scenes  # noqa: E402

rank, n, device, d = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
part = sys.argv[5] if len(sys.argv) > 5 else "auto"
W, H, SPP = 203, 90, 24
txt, cfg = scenes.weekend(seed=1, width=W, height=H, spp=SPP)
cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = cfg.use_reflection = cfg.use_refraction = True
flat = api.flatten_zscene(txt)
ctx = api.Context(device)
h = ctx.group_export(W * H)
with open(os.path.join(d, f"h{rank}.tmp"), "wb") as f:
    f.write(h)
os.rename(os.path.join(d, f"h{rank}.tmp"), os.path.join(d, f"h{rank}.bin"))
t0 = time.time()
while not all(os.path.exists(os.path.join(d, f"h{r}.bin")) for r in range(n)):
    time.sleep(0.01)
    assert time.time() - t0 < 120, "peers did not show up"
handles = [open(os.path.join(d, f"h{r}.bin"), "rb").read() for r in range(n)]
ctx.group_join(rank, handles, f"/zb_test_{os.path.basename(d)}_{part}")
sc = ctx.upload(flat)
keys = ("beauty", "albedo", "normal", "z_depth", "reflection", "refraction")
frames = {}
for k in keys:
    p = os.path.join(d, f"frame_{k}.bin")
    if rank == 0:
        with open(p + ".tmp", "wb") as f:
            f.truncate(W * H * 24)
        os.rename(p + ".tmp", p)
    while not os.path.exists(p):
        time.sleep(0.01)
    frames[k] = np.memmap(p, dtype=np.float64, mode="r+", shape=(H, W, 3))
for rep in range(3):  # several frames: the double-buffered sums and re-recorded events must stay consistent
    sc.render_into(cfg, frames, seed=5, partition=part)
sc.close()
ctx.group_leave()
if rank == 0:
    single = api.Context(device)
    s1 = single.upload(flat)
    ref = s1.render(cfg, seed=5)
    s1.close()
    single.close()
    res = {k: bool(np.array_equal(np.asarray(frames[k]), ref[k])) for k in keys}
    res["mean"] = float(ref["beauty"].mean())
    with open(os.path.join(d, "result.json"), "w") as f:
        json.dump(res, f)
ctx.close()
