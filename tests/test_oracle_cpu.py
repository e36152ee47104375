"""CPU checks of the oracle (test infrastructure): oracle/_ref/ref_harness — the unmodified reference compiled
by oracle/Makefile — must reproduce the committed deterministic golden vectors (per-ray hit records and
primary-hit ids).  Skipped where the harness binary is absent (it is built only where /root/reference exists,
and travels to the GPU box as a prebuilt file)."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from raytracer_project_b200 import scenes
from tests.golden.golden_scenes import GOLDEN_SCENES
from tests.golden.make_golden import read_ids, scene_digest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
needs_harness = pytest.mark.skipif(not os.path.exists(HARNESS), reason="oracle/_ref/ref_harness not built")


def _scratch(txt, cfg, files):
    d = tempfile.mkdtemp(prefix="zb_oracle_")
    os.makedirs(os.path.join(d, "assets", "hdr_maps"))  # camera::initialize side effect (camera.hpp:123-140)
    scenes.materialize(d, txt, cfg, files)
    return d


@needs_harness
@pytest.mark.parametrize("name", ["quirks", "cornell"])
def test_oracle_reproduces_probe_vectors(name):
    gold = np.load(os.path.join(GOLDEN, f"{name}_probe.npz"))
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    d = _scratch(txt, cfg, files)
    n = len(gold["o"])
    with open(os.path.join(d, "rays.bin"), "wb") as f:
        f.write(np.int32(n).tobytes() + gold["o"].tobytes() + gold["d"].tobytes())
    subprocess.run([HARNESS, "probe", "s.zscene", "s.cfg", "rays.bin", "probe.bin"], cwd=d, check=True, capture_output=True)
    rec = np.frombuffer(open(os.path.join(d, "probe.bin"), "rb").read()[4:], np.float64).reshape(n, 19)
    # leaf ids, t and hit points are deterministic functions of the ray; the reference's BVH has random split
    # axes (bvh.hpp:17) but closest-hit answers do not depend on the tree (exact ties aside)
    same = (rec[:, 1] == gold["rec"][:, 1])
    assert same.mean() > 0.9995
    assert np.allclose(rec[same][:, [0, 4, 5, 6]], gold["rec"][same][:, [0, 4, 5, 6]], rtol=1e-12, atol=1e-12, equal_nan=True)


@needs_harness
def test_oracle_reproduces_quirks_ids():
    name = "quirks"
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
    gold = np.load(os.path.join(GOLDEN, f"{name}_{wh[0]}x{wh[1]}_ids.npz"))
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    d = _scratch(txt, cfg, files)
    subprocess.run([HARNESS, "ids", "s.zscene", "s.cfg", "ids.bin"], cwd=d, check=True, capture_output=True)
    leaf, tri, t = read_ids(os.path.join(d, "ids.bin"))
    assert (leaf == gold["leaf"]).mean() > 0.9999


def test_golden_fixtures_are_complete():
    need = ["weekend_1200x675_ids.npz", "weekend_300x169_stats.npz", "weekend_probe.npz", "quirks_320x180_aov.npz", "kat.json",
            "shipped_texels.json"]
    for name, (_, _, ids_wh, stats_wh, _) in GOLDEN_SCENES.items():
        need += [f"{name}_{ids_wh[0]}x{ids_wh[1]}_ids.npz", f"{name}_probe.npz"]
        if stats_wh:
            need.append(f"{name}_{stats_wh[0]}x{stats_wh[1]}_stats.npz")
    for f in need:
        assert os.path.exists(os.path.join(GOLDEN, f)), f


# ------------------------------------------------------------------ the CPU restatement (oracle/zoracle.cpp)
# Pinned on outputs of the reference itself: the committed fixtures below were produced by the unmodified
# reference (oracle/_ref/ref_harness).  The restatement is compiled here if it is not there yet (g++ only).
ZORACLE = os.path.join(ROOT, "oracle", "_build", "zoracle")


@pytest.fixture(scope="module")
def zoracle():
    if not os.path.exists(ZORACLE):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, capture_output=True)
    return ZORACLE


def _golden_scene(name, wh, spp=1, ids=False):
    if name == "weekend":
        txt, cfg = scenes.weekend(seed=1, width=wh[0], height=wh[1], spp=spp)
        return txt, cfg, None
    gen, kw, _, _, ids_kw = GOLDEN_SCENES[name]
    return getattr(scenes, gen)(width=wh[0], height=wh[1], spp=spp, **{**kw, **(ids_kw if ids else {})})


ALL_SCENES = ["weekend", "quirks", "cornell", "mesh", "exterior"]


def _ids_wh(name):
    return (1200, 675) if name == "weekend" else GOLDEN_SCENES[name][2]


@pytest.mark.parametrize("name", ALL_SCENES)
def test_restatement_matches_reference_hit_records(zoracle, name):
    """Per-ray hit_record of the restatement vs the reference's: leaf id, face id, t, front_face, p, normal,
    (u, v), albedo, emitted and the background colour of misses, to 1e-9 (both compute in double)."""
    gold = np.load(os.path.join(GOLDEN, f"{name}_probe.npz"))
    txt, cfg, files = _golden_scene(name, _ids_wh(name), ids=True)
    d = _scratch(txt, cfg, files)
    n = len(gold["o"])
    with open(os.path.join(d, "rays.bin"), "wb") as f:
        f.write(np.int32(n).tobytes() + gold["o"].tobytes() + gold["d"].tobytes())
    subprocess.run([zoracle, "probe", "s.zscene", "s.cfg", "rays.bin", "probe.bin"], cwd=d, check=True, capture_output=True)
    rec = np.frombuffer(open(os.path.join(d, "probe.bin"), "rb").read()[4:], np.float64).reshape(n, 19)
    g = gold["rec"]
    same = rec[:, 1] == g[:, 1]
    # the only stochastic leaf is the Cornell fog (constant_medium.hpp:64)
    assert same.mean() >= (0.995 if name == "cornell" else 0.9995), same.mean()
    hit = same & np.isfinite(g[:, 0])
    cols = [0, 2, 3, 4, 5, 6, 7, 8, 9, 12, 13, 14, 15, 16, 17]
    assert np.allclose(rec[hit][:, cols], g[hit][:, cols], rtol=1e-9, atol=1e-9)
    # u, v are defined for spheres and cubes only: triangle::hit never writes them (triangle.hpp:75-79)
    uv_ok = np.isclose(rec[hit][:, 10:12], g[hit][:, 10:12], rtol=1e-9, atol=1e-9).all(axis=1)
    assert uv_ok.mean() > 0.97
    miss = same & ~np.isfinite(g[:, 0])
    if miss.any():
        assert np.allclose(rec[miss][:, 12:15], g[miss][:, 12:15], rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("name", ALL_SCENES)
def test_restatement_matches_reference_primary_ids(zoracle, name):
    wh = _ids_wh(name)
    gold = np.load(os.path.join(GOLDEN, f"{name}_{wh[0]}x{wh[1]}_ids.npz"))
    txt, cfg, files = _golden_scene(name, wh, ids=True)
    d = _scratch(txt, cfg, files)
    subprocess.run([zoracle, "ids", "s.zscene", "s.cfg", "ids.bin"], cwd=d, check=True, capture_output=True)
    leaf, tri, t = read_ids(os.path.join(d, "ids.bin"))
    assert (leaf == gold["leaf"]).mean() >= 0.9999
    both = (leaf == gold["leaf"]) & (leaf != 0xFFFFFFFF)
    assert (tri[both] == gold["tri"][both]).mean() >= 0.9999
    tol = 1e-9 if gold["t"].dtype == np.float64 else 1e-6  # the Weekend fixture keeps t as float32
    assert np.allclose(t[both], gold["t"][both], rtol=tol, atol=tol)


@pytest.mark.parametrize("name,spp,tol", [("weekend", 128, 0.04), ("quirks", 128, 0.05), ("mesh", 64, 0.03)])
def test_restatement_radiance_is_consistent_with_reference(zoracle, name, spp, tol):
    """render_rows / ray_color restated: a low-spp render must sit at its own Monte-Carlo noise distance from
    the reference's 4096-spp mean, with the image mean within 0.5 % (the RNG streams differ by construction)."""
    wh = (300, 169) if name == "weekend" else GOLDEN_SCENES[name][3]
    gold = np.load(os.path.join(GOLDEN, f"{name}_{wh[0]}x{wh[1]}_stats.npz"))
    txt, cfg, files = _golden_scene(name, wh, spp=spp)
    d = _scratch(txt, cfg, files)
    subprocess.run([zoracle, "render", "s.zscene", "s.cfg", "img.bin", "7"], cwd=d, check=True, capture_output=True)
    raw = open(os.path.join(d, "img.bin"), "rb").read()
    img = np.frombuffer(raw[8:8 + wh[0] * wh[1] * 24], np.float64).reshape(wh[1], wh[0], 3)
    ref = gold["mean"].astype(np.float64)
    rrmse = float(np.sqrt(((img - ref) ** 2).sum() / (ref ** 2).sum()))
    assert rrmse < tol, rrmse
    assert abs(img.mean() / ref.mean() - 1.0) < 0.005


def test_restatement_light_passes_are_consistent_with_reference(zoracle):
    """Reflection / refraction light passes (camera.hpp:492-517): restatement at 512 spp vs the reference's
    2048-spp buffers — image means within 2 %, rRMSE at the Monte-Carlo noise level of the two estimates."""
    gold = np.load(os.path.join(GOLDEN, "quirks_200x112_light.npz"))
    txt, cfg, files = scenes.quirks(seed=5, width=200, height=112, spp=512)
    cfg.use_reflection = cfg.use_refraction = True
    d = _scratch(txt, cfg, files)
    subprocess.run([zoracle, "render", "s.zscene", "s.cfg", "img.bin", "3"], cwd=d, check=True, capture_output=True)
    raw = open(os.path.join(d, "img.bin"), "rb").read()
    bufs = [np.frombuffer(raw[8 + k * 200 * 112 * 24:8 + (k + 1) * 200 * 112 * 24], np.float64).reshape(112, 200, 3) for k in range(6)]
    for key, img in (("reflection", bufs[4]), ("refraction", bufs[5])):
        ref = gold[key].astype(np.float64)
        assert ref.mean() > 1e-3, key  # the scene exercises both buffers
        rrmse = float(np.sqrt(((img - ref) ** 2).sum() / (ref ** 2).sum()))
        print(key, "mean", img.mean(), ref.mean(), "rRMSE", rrmse)
        assert abs(img.mean() / ref.mean() - 1.0) < 0.02, (key, img.mean(), ref.mean())
        assert rrmse < 0.12, (key, rrmse)


# ------------------------------------------------------------------ finishing chain
def _run_finish(binary, img, pass_kind, kw, d):
    from raytracer_project_b200 import api
    from tests.golden.make_golden import read_finish, write_image
    write_image(os.path.join(d, "in.bin"), img)
    with open(os.path.join(d, "post.cfg"), "w") as f:
        f.write(api.post_text(pass_kind, **kw))
    subprocess.run([binary, "finish", "in.bin", "post.cfg", "out.bin"], cwd=d, check=True, capture_output=True)
    return read_finish(os.path.join(d, "out.bin"))


def test_restatement_finishing_chain_matches_reference(zoracle, tmp_path):
    """analyze_framebuffer, auto-exposure, exposure, bloom, sharpening and process() restated in zoracle against
    the reference's final_framebuffer for every parameter set (both compute in double, HSV in float)."""
    from tests.golden.finish_cases import CASES, hdr_image
    gold = np.load(os.path.join(GOLDEN, "finish_96x60.npz"))
    img = hdr_image()
    for name, (pass_kind, kw) in CASES.items():
        out, avg, mx, exposure, hist = _run_finish(zoracle, img, pass_kind, kw, str(tmp_path))
        assert np.allclose([avg, mx, exposure], gold[name + "_stats"], rtol=1e-6), name
        assert np.array_equal(hist, gold["histogram"]), name
        close = np.isclose(out, gold[name + "_image"], rtol=1e-6, atol=1e-7)
        assert close.mean() == 1.0, (name, close.mean(), np.abs(out - gold[name + "_image"]).max())


# ------------------------------------------------------------------ image export
def test_restatement_export_matches_reference_png_bytes(zoracle, tmp_path):
    """process_framebuffer_to_image restated in zoracle against the pixels of the PNGs the reference wrote
    (save_render_pass, camera.hpp:299-343).  Both in double: the 8-bit values agree except where 255.999 * v sits on
    an integer boundary."""
    from raytracer_project_b200 import api
    from tests.golden.finish_cases import EXPORT_CASES, hdr_image
    from tests.golden.make_golden import read_rgb8, write_image
    gold = np.load(os.path.join(GOLDEN, "export_96x60.npz"))
    d = str(tmp_path)
    write_image(os.path.join(d, "in.bin"), hdr_image())
    for name, (rp, kw) in EXPORT_CASES.items():
        with open(os.path.join(d, "post.cfg"), "w") as f:
            f.write(api.post_text(0, **kw) + f"render_pass {rp}\n")
        subprocess.run([zoracle, "export", "in.bin", "post.cfg", "out.bin"], cwd=d, check=True, capture_output=True)
        out = read_rgb8(os.path.join(d, "out.bin")).astype(np.int32)
        ref = gold[name].astype(np.int32)
        diff = np.abs(out - ref)
        assert diff.max() <= 1 and (diff != 0).mean() < 1e-3, (name, diff.max(), (diff != 0).mean())
