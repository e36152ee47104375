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


def scene_digest(scene_text, cfg, files):
    """sha256 over the scene text, config and asset bytes: the GPU test regenerates the scene from the same
    seeded generator and refuses to compare if it does not reproduce what the reference rendered."""
    import hashlib
    h = hashlib.sha256()
    h.update(scene_text.encode())
    h.update(cfg.to_text().encode())
    for k in sorted(files or {}):
        if k.endswith(".zbtex"):
            continue  # derived from the same arrays; the reference never reads them
        v = files[k]
        h.update(k.encode())
        h.update(v if isinstance(v, (bytes, bytearray)) else v.encode())
    return h.hexdigest()


def make_ids(name, scene_text, cfg, files=None):
    d = scratch_dir()
    scenes.materialize(d, scene_text, cfg, files)
    op = os.path.join(d, "ids.bin")
    info = run_harness(["ids", "s.zscene", "s.cfg", op], d)
    leaf, tri, t = read_ids(op)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_ids.npz"), leaf=leaf, tri=tri, t=t.astype(np.float32),
                        digest=scene_digest(scene_text, cfg, files))
    print(name, "ids", info, "hit fraction", float((leaf != 0xFFFFFFFF).mean()))


def make_stats(name, scene_text, cfg, batches, files=None):
    d = scratch_dir()
    scenes.materialize(d, scene_text, cfg, files)
    op = os.path.join(d, "stats.bin")
    info = run_harness(["stats", "s.zscene", "s.cfg", op, str(batches)], d)
    mean, var = read_stats(op)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_stats.npz"), mean=mean, var_of_mean=var,
                        spp=cfg.samples_per_pixel, batches=batches, digest=scene_digest(scene_text, cfg, files))
    print(name, "stats", info)


def read_render(path, want=4):
    with open(path, "rb") as f:
        w, h = np.frombuffer(f.read(8), np.int32)
        n = int(w) * int(h) * 3
        return [np.frombuffer(f.read(8 * n), np.float64).reshape(h, w, 3) for _ in range(6)][:want]


def make_aov(name, scene_text, cfg, files=None):
    """First-hit AOV buffers (albedo, view-space normal, z-depth; camera.hpp:464-488, :518-541)."""
    d = scratch_dir()
    cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = True
    scenes.materialize(d, scene_text, cfg, files)
    op = os.path.join(d, "render.bin")
    info = run_harness(["render", "s.zscene", "s.cfg", op], d)
    _, albedo, normal, z = read_render(op)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_aov.npz"), albedo=albedo.astype(np.float32),
                        normal=normal.astype(np.float32), z_depth=z.astype(np.float32),
                        spp=cfg.samples_per_pixel, digest=scene_digest(scene_text, cfg, files))
    print(name, "aov", info)


def make_light(name, scene_text, cfg, files=None):
    """Reflection / refraction light passes (camera.hpp:492-517) of the reference at high spp."""
    d = scratch_dir()
    cfg.use_reflection = cfg.use_refraction = True
    scenes.materialize(d, scene_text, cfg, files)
    op = os.path.join(d, "render.bin")
    info = run_harness(["render", "s.zscene", "s.cfg", op], d)
    bufs = read_render(op, 6)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_light.npz"), beauty=bufs[0].astype(np.float32),
                        reflection=bufs[4].astype(np.float32), refraction=bufs[5].astype(np.float32),
                        spp=cfg.samples_per_pixel, digest=scene_digest(scene_text, cfg, files))
    print(name, "light", info)


def quirks_light():
    txt, cfg, files = scenes.quirks(seed=5, width=200, height=112, spp=2048)
    make_light("quirks_200x112", txt, cfg, files)


def read_finish(path):
    with open(path, "rb") as f:
        w, h = np.frombuffer(f.read(8), np.int32)
        img = np.frombuffer(f.read(8 * int(w) * int(h) * 3), np.float64).reshape(h, w, 3)
        avg, mx, exposure = np.frombuffer(f.read(24), np.float64)
        hist = np.frombuffer(f.read(1024), np.int32)
    return img, float(avg), float(mx), float(exposure), hist


def write_image(path, img):
    with open(path, "wb") as f:
        f.write(np.array([img.shape[1], img.shape[0]], np.int32).tobytes() + np.ascontiguousarray(img, np.float64).tobytes())


def finish():
    """Finishing chain of the reference (camera.hpp:142-207, color_processing.hpp, bloom.hpp) on the seeded HDR frame."""
    from raytracer_project_b200 import api
    from tests.golden.finish_cases import CASES, hdr_image
    d = scratch_dir()
    write_image(os.path.join(d, "in.bin"), hdr_image())
    out = {}
    for name, (pass_kind, kw) in CASES.items():
        with open(os.path.join(d, "post.cfg"), "w") as f:
            f.write(api.post_text(pass_kind, **kw))
        info = run_harness(["finish", "in.bin", "post.cfg", "out.bin"], d)
        img, avg, mx, exposure, hist = read_finish(os.path.join(d, "out.bin"))
        out[name + "_image"] = img
        out[name + "_stats"] = np.array([avg, mx, exposure])
        out["histogram"] = hist
        print("finish", name, info)
    np.savez_compressed(os.path.join(GOLDEN, "finish_96x60.npz"), **out)


def read_rgb8(path):
    with open(path, "rb") as f:
        w, h = np.frombuffer(f.read(8), np.int32)
        return np.frombuffer(f.read(int(w) * int(h) * 3), np.uint8).reshape(h, w, 3)


def export():
    """Image export of the reference (camera.hpp:299-343, :701-780): the PNG bytes of each pass kind."""
    from raytracer_project_b200 import api
    from tests.golden.finish_cases import EXPORT_CASES, hdr_image
    d = scratch_dir()
    write_image(os.path.join(d, "in.bin"), hdr_image())
    out = {}
    for name, (rp, kw) in EXPORT_CASES.items():
        with open(os.path.join(d, "post.cfg"), "w") as f:
            f.write(api.post_text(0, **kw) + f"render_pass {rp}\n")
        info = run_harness(["export", "in.bin", "post.cfg", "out.bin"], d)
        out[name] = read_rgb8(os.path.join(d, "out.bin")).copy()
        out[name + "_file"] = np.array(os.path.basename(info["file"]))
        print("export", name, info)
    np.savez_compressed(os.path.join(GOLDEN, "export_96x60.npz"), **out)


def wire():
    """BVH debug view of the reference (bvh.hpp:56-110) at level 0 — the root box, which does not depend on the
    reference's random tree — and, for the record only, at the leaf level of ONE reference tree."""
    txt, cfg, files = scenes.wire_demo(spp=256)
    d = scratch_dir()
    scenes.materialize(d, txt, cfg, files)
    out = {"digest": np.array(scene_digest(txt, cfg, files)), "spp": cfg.samples_per_pixel}
    for name, level, thick in (("level0", 0, 0.05), ("level0_thin", 0, 0.01), ("leaves", -1, 0.05)):
        with open(os.path.join(d, "wire.cfg"), "w") as f:
            f.write(cfg.to_text() + f"bvh_debug 1\ndebug_bvh_level {level}\nbvh_thickness {thick}\n")
        info = run_harness(["render", "s.zscene", "wire.cfg", "out.bin"], d)
        out[name] = read_render(os.path.join(d, "out.bin"))[0].astype(np.float32)
        out[name + "_params"] = np.array([level, thick])
        print("wire", name, info)
    np.savez_compressed(os.path.join(GOLDEN, "wire_240x135.npz"), **out)


def probe_rays(cfg, n, seed):
    """Known-answer rays for a scene: a third camera rays through random image positions (un-normalised, as
    get_ray builds them), the rest random origins in a ball around the look-at point with random directions of
    random length (epsilons scale with |dir|).  float32, identical on both sides."""
    rng = np.random.default_rng(seed)
    lf, la = np.array(cfg.lookfrom, float), np.array(cfg.lookat, float)
    dist = np.linalg.norm(lf - la)
    n_cam = n // 3
    w = (lf - la) / dist
    u = np.cross(np.array(cfg.vup, float), w)
    u /= np.linalg.norm(u)
    v = np.cross(w, u)
    half_h = np.tan(np.radians(cfg.vfov) / 2) * cfg.focus_dist
    half_w = half_h * cfg.image_width / cfg.image_height
    sx, sy = rng.uniform(-1, 1, n_cam), rng.uniform(-1, 1, n_cam)
    d_cam = -cfg.focus_dist * w[None] + (sx * half_w)[:, None] * u[None] + (sy * half_h)[:, None] * v[None]
    o_cam = np.repeat(lf[None], n_cam, 0)
    m = n - n_cam
    g = rng.normal(size=(m, 3))
    o_rnd = la[None] + 0.8 * dist * g / np.linalg.norm(g, axis=1, keepdims=True) * rng.random((m, 1)) ** (1 / 3)
    g = rng.normal(size=(m, 3))
    d_rnd = g / np.linalg.norm(g, axis=1, keepdims=True) * rng.uniform(0.3, 3.0, (m, 1))
    return (np.concatenate([o_cam, o_rnd]).astype(np.float32), np.concatenate([d_cam, d_rnd]).astype(np.float32))


def make_probe(name, scene_text, cfg, files=None, n=6000, seed=99):
    d = scratch_dir()
    scenes.materialize(d, scene_text, cfg, files)
    o, dr = probe_rays(cfg, n, seed)
    with open(os.path.join(d, "rays.bin"), "wb") as f:
        f.write(np.int32(n).tobytes() + o.tobytes() + dr.tobytes())
    info = run_harness(["probe", "s.zscene", "s.cfg", "rays.bin", "probe.bin"], d)
    raw = open(os.path.join(d, "probe.bin"), "rb").read()
    rec = np.frombuffer(raw[4:], np.float64).reshape(n, 19)
    np.savez_compressed(os.path.join(GOLDEN, f"{name}_probe.npz"), o=o, d=dr, rec=rec,
                        digest=scene_digest(scene_text, cfg, files))
    print(name, "probe", info, "hit fraction", float(np.isfinite(rec[:, 0]).mean()))


SHIPPED_IMAGES = ["bump_maps/concrete_bump_map.jpg", "bump_maps/scratches_bump_map.jpg", "bump_maps/water_bump_map.jpg",
                  "bump_maps/wood_bump_map.jpg", "textures/fine-wood.jpg", "textures/wood_box.png"]


def texels():
    """stb_image decode of every shipped texture by the reference harness (texture.hpp:31): size, sha256 and a few probe
    texels of the RGB8 buffer, so the host's decode can be checked texel for texel without the harness."""
    import hashlib
    d = scratch_dir()
    out = {}
    for rel in SHIPPED_IMAGES:
        src = os.path.join(GOLDEN, "assets", rel)
        dump = os.path.join(d, "t.zbtex")
        info = run_harness(["decode", src, dump, "0"], d)
        raw = open(dump, "rb").read()
        w, h, is_float = np.frombuffer(raw[4:16], np.uint32)
        px = np.frombuffer(raw[16:], np.uint8).reshape(int(h), int(w), 3)
        out[rel] = {"width": int(w), "height": int(h), "sha256": hashlib.sha256(px.tobytes()).hexdigest(),
                    "probe": px[::max(int(h) // 7, 1), ::max(int(w) // 5, 1)].reshape(-1, 3)[:24].tolist()}
        print("texels", rel, info, out[rel]["sha256"][:12])
    with open(os.path.join(GOLDEN, "shipped_texels.json"), "w") as f:
        json.dump(out, f, indent=1)


def kat():
    """Per-function known answers (oracle/ref_kat.inc): the reference's refract / reflect / reflectance / texture::value /
    get_bumped_normal / get_sphere_uv on the KAT scene, and get_background_color for the three environment modes."""
    cases, digests = [], {}
    for env in ("solid", "hdr", "sun"):
        txt, cfg, files = scenes.kat_scene(env)
        d = scratch_dir()
        scenes.materialize(d, txt, cfg, files)
        print("kat", env, run_harness(["kat", "s.zscene", "s.cfg", "kat.json"], d))
        got = json.load(open(os.path.join(d, "kat.json")))["cases"]
        cases += [c for c in got if c["op"] == "background" or env == "solid"]
        digests[env] = scene_digest(txt, cfg, files)
    with open(os.path.join(GOLDEN, "kat.json"), "w") as f:
        json.dump({"digests": digests, "cases": cases}, f)
    print("kat cases:", len(cases))


def quirks_aov():
    txt, cfg, files = scenes.quirks(seed=5, width=320, height=180, spp=512)
    cfg.z_depth_max_dist = 15.0
    make_aov("quirks_320x180", txt, cfg, files)


def weekend_ids():
    txt, cfg = scenes.weekend(seed=1, width=1200, height=675)
    make_ids("weekend_1200x675", txt, cfg)


def weekend_probe():
    txt, cfg = scenes.weekend(seed=1, width=1200, height=675)
    make_probe("weekend", txt, cfg)


def weekend_stats():
    # 16 x 256 spp = 4096 spp at 300x169 (SURVEY 8(c) parity procedure)
    txt, cfg = scenes.weekend(seed=1, width=300, height=169, spp=256)
    make_stats("weekend_300x169", txt, cfg, 16)


# Reduced-size instances of configs C2-C4 and the quirk scene: same generators as the bench, sized so the
# reference finishes 4096 spp in minutes on 8 cores.  GOLDEN_SCENES is shared with tests/test_gpu_parity.py.
from tests.golden.golden_scenes import GOLDEN_SCENES, STATS_BATCHES  # noqa: E402


def scene_targets():
    t = {}
    for name, (gen, kw, ids_wh, stats_wh, ids_kw) in GOLDEN_SCENES.items():
        def ids(name=name, gen=gen, kw={**kw, **ids_kw}, wh=ids_wh):
            txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **kw)
            make_ids(f"{name}_{wh[0]}x{wh[1]}", txt, cfg, files)

        def stats(name=name, gen=gen, kw=kw, wh=stats_wh):
            txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=256, **kw)
            make_stats(f"{name}_{wh[0]}x{wh[1]}", txt, cfg, STATS_BATCHES, files)
        def probe(name=name, gen=gen, kw={**kw, **ids_kw}, wh=ids_wh):
            txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **kw)
            make_probe(name, txt, cfg, files)
        t[f"{name}_ids"], t[f"{name}_probe"] = ids, probe
        if stats_wh:
            t[f"{name}_stats"] = stats
    return t


TARGETS = {"texels": texels, "kat": kat, "finish": finish, "export": export, "wire": wire, "quirks_aov": quirks_aov, "quirks_light": quirks_light, "weekend_ids": weekend_ids, "weekend_probe": weekend_probe, "weekend_stats": weekend_stats, **scene_targets()}

if __name__ == "__main__":
    names = sys.argv[1:] or list(TARGETS)
    for n in names:
        TARGETS[n]()
