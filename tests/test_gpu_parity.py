"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against
golden fixtures produced by the UNMODIFIED reference (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from raytracer_project_b200 import api, scenes

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def weekend_scene(gpu_ctx):
    txt, _ = scenes.weekend(seed=1)
    flat = api.flatten_zscene(txt)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    yield sc
    sc.close()


def test_native_library_is_loaded(gpu_ctx):
    info = gpu_ctx.device_info()
    assert info["cc"][0] >= 10 and info["sm_count"] > 0
    assert any("libzenith_b200.so" in l for l in open("/proc/self/maps"))


def test_weekend_primary_ids_match_reference(weekend_scene):
    """north_star: jitter and DOF disabled, primary-hit id buffer identical in >= 99.99 % of pixels;
    mismatches only at epsilon-grazing ties (silhouette pixels)."""
    gold = np.load(os.path.join(GOLDEN, "weekend_1200x675_ids.npz"))
    _, cfg = scenes.weekend(seed=1, width=1200, height=675)
    leaf, prim, t = weekend_scene.primary_ids(cfg)
    same = leaf == gold["leaf"]
    frac = same.mean()
    print(f"primary-id agreement {frac * 100:.5f} % ({(~same).sum()} of {same.size} differ)")
    assert frac >= 0.9999
    hit = same & (gold["leaf"] != 0xFFFFFFFF)
    rel = np.abs(t[hit] - gold["t"][hit]) / np.maximum(gold["t"][hit], 1e-6)
    # fp32 t against the reference's double t: ~1e-7 typically; at grazing incidence the root is
    # sqrt-conditioned (disc -> 0), so the tail is bounded separately
    assert np.quantile(rel, 0.999) < 2e-4, np.quantile(rel, 0.999)
    assert rel.max() < 5e-3, rel.max()
    # every mismatch sits on a silhouette: one of its 8 neighbours in the reference image has the GPU's id
    ys, xs = np.nonzero(~same)
    g = np.pad(gold["leaf"], 1, mode="edge")
    for y, x in zip(ys, xs):
        assert leaf[y, x] in g[y:y + 3, x:x + 3], (y, x)


def test_weekend_radiance_converges_to_reference(weekend_scene):
    """north_star: converged radiance within 1 % relative RMSE of the reference's double-precision
    output and per-pixel means within 3 sigma.  Reference: 16 x 256 spp = 4096 spp (fixture)."""
    gold = np.load(os.path.join(GOLDEN, "weekend_300x169_stats.npz"))
    ref, var_ref = gold["mean"].astype(np.float64), gold["var_of_mean"].astype(np.float64)
    _, cfg = scenes.weekend(seed=1, width=300, height=169, spp=16384)
    img = weekend_scene.render(cfg, seed=7)["beauty"]
    assert np.isfinite(img).all()
    rrmse = np.sqrt(((img - ref) ** 2).sum() / (ref ** 2).sum())
    print(f"rRMSE vs 4096-spp reference: {rrmse * 100:.3f} %  mean gpu {img.mean(axis=(0, 1))} ref {ref.mean(axis=(0, 1))}")
    assert rrmse <= 0.01
    # GPU estimate has 4x the samples => var_gpu ~ var_ref / 4 for the same per-sample variance
    sigma = np.sqrt(1.25 * var_ref + 1e-12)
    out = np.abs(img - ref) > 3 * sigma
    print(f"outside 3 sigma: {out.mean() * 100:.3f} % of pixel channels")
    assert out.mean() <= 0.015  # Student-t with 15 dof puts ~0.9 % outside 3 sigma by chance
    # no bias: image mean agrees to a fraction of a percent
    assert np.allclose(img.mean(axis=(0, 1)), ref.mean(axis=(0, 1)), rtol=3e-3)


def test_render_is_deterministic_and_partitions_agree(weekend_scene):
    _, cfg = scenes.weekend(seed=1, width=160, height=90, spp=32)
    a = weekend_scene.render(cfg, seed=3)["beauty"]
    b = weekend_scene.render(cfg, seed=3)["beauty"]
    assert np.array_equal(a, b)
    c = weekend_scene.render(cfg, seed=4)["beauty"]
    assert not np.array_equal(a, c)
    # batched launches: same samples; per-launch partial sums are exact (64-bit fixed point) but are
    # rounded to fp32 when they are added to the float4 accumulation buffer between launches
    d = weekend_scene.render(cfg, seed=3, samples_per_launch=5)["beauty"]
    assert np.allclose(a, d, rtol=2e-6, atol=1e-7)
    # sample slices: partial means add up to the full frame up to fp32 summation order
    parts = [weekend_scene.render(cfg, seed=3, sample_range=r)["beauty"] for r in ((0, 8), (8, 32))]
    assert np.allclose(parts[0] + parts[1], a, rtol=1e-5, atol=1e-6)
    # tile stripes: disjoint pixels, union is bit-identical to the single-GPU frame
    tiles = [weekend_scene.render(cfg, seed=3, tile=(r, 3))["beauty"] for r in range(3)]
    assert np.array_equal(tiles[0] + tiles[1] + tiles[2], a)
    assert ((tiles[0] != 0).any(axis=2) & (tiles[1] != 0).any(axis=2)).sum() == 0


def test_stats_counters(weekend_scene, gpu_ctx):
    _, cfg = scenes.weekend(seed=1, width=200, height=112, spp=8)
    weekend_scene.render(cfg, seed=1, stats=True)
    st = gpu_ctx.stats()
    assert st["paths"] == 200 * 112 * 8
    assert 1.5 < st["rays"] / st["paths"] < 4.0          # reference: 2.60 rays/path
    assert st["node_visits"] > st["rays"] and st["prim_tests"] > 0
    print({k: v for k, v in st.items()})


def test_aovs(weekend_scene):
    _, cfg = scenes.weekend(seed=1, width=160, height=90, spp=16)
    cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = True
    cfg.z_depth_max_dist = 30.0
    r = weekend_scene.render(cfg, seed=1)
    assert r["albedo"].min() >= 0 and r["albedo"].max() <= 1.0 + 1e-6
    assert r["normal"].min() >= -1e-6 and r["normal"].max() <= 1 + 1e-6
    sky = r["normal"][0, 0]
    assert np.allclose(sky, [0.5, 0.5, 1.0], atol=1e-6)      # miss writes (.5,.5,1) (camera.hpp:522-525)
    assert np.allclose(r["z_depth"][0, 0], 0.0) and r["z_depth"].max() > 0.3


# ------------------------------------------------------------------ configs C2-C4 + quirk scene
from tests.golden.golden_scenes import GOLDEN_SCENES, Q13_LEAVES  # noqa: E402
from tests.golden.make_golden import scene_digest  # noqa: E402


STATS_SCENES = [n for n, v in GOLDEN_SCENES.items() if v[3] and os.path.exists(os.path.join(GOLDEN, f"{n}_{v[3][0]}x{v[3][1]}_stats.npz"))]
SMALL_SCENES = [n for n in GOLDEN_SCENES if n not in ("mesh_full", "big")]  # kernel-vs-kernel tests: seconds to flatten


def _load_scene(gpu_ctx, tmp_path, name, wh, spp, extra=None):
    gen, kw, _, _, _ = GOLDEN_SCENES[name]
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=spp, **{**kw, **(extra or {})})
    path = scenes.materialize(str(tmp_path), txt, cfg, files)
    flat = api.flatten_zscene(path=path)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    return sc, cfg, scene_digest(txt, cfg, files)


@pytest.mark.parametrize("name", list(GOLDEN_SCENES))
def test_scene_primary_ids_match_reference(name, gpu_ctx, tmp_path):
    """Primary-hit leaf ids (and mesh face ids) against the reference on cubes, triangles, OBJ meshes,
    transformed instances and textured objects; mismatches only at silhouettes / grazing ties."""
    _, _, wh, _, ids_kw = GOLDEN_SCENES[name]
    gold = np.load(os.path.join(GOLDEN, f"{name}_{wh[0]}x{wh[1]}_ids.npz"))
    sc, cfg, digest = _load_scene(gpu_ctx, tmp_path, name, wh, 1, ids_kw)
    assert digest == str(gold["digest"]), "the regenerated scene differs from the one the reference rendered"
    leaf, prim, t = sc.primary_ids(cfg)
    sc.close()
    same = leaf == gold["leaf"]
    q13 = np.zeros_like(same)
    if name in Q13_LEAVES:  # the reference's mirrored rotate_y box (tests/golden/golden_scenes.py)
        q13 = ~same & np.isin(leaf, list(Q13_LEAVES[name])) & (t < gold["t"])
        print(f"{name}: {q13.sum()} pixels explained by the reference's rotate_y bounding box (q13)")
        assert q13.sum() < 1e-3 * same.size
    same = same | q13
    print(f"{name}: leaf-id agreement {same.mean() * 100:.5f} % ({(~same).sum()} of {same.size} differ)")
    assert same.mean() >= 0.9999
    same = same & ~q13
    hit = same & (gold["leaf"] != 0xFFFFFFFF)
    rel = np.abs(t[hit] - gold["t"][hit]) / np.maximum(gold["t"][hit], 1e-6)
    # a handful of grazing / shared-edge hits land on the neighbouring facet (different plane, same pixel)
    assert np.quantile(rel, 0.999) < 2e-4 and np.quantile(rel, 0.9999) < 5e-3 and rel.max() < 0.1, (np.quantile(rel, 0.999), rel.max())
    # mesh faces: same triangle wherever the same leaf was hit, except on shared edges (inclusive edge rule,
    # triangle.hpp:62) where either neighbour is a legitimate answer
    face_same = (prim == gold["tri"])[hit]
    print(f"{name}: face-id agreement {face_same.mean() * 100:.4f} %")
    # the shipped bowl.obj carries coincident (duplicated) faces: another face at the SAME t is a tie, not a miss
    tie = face_same | (rel < 1e-5)
    print(f"{name}: face-id agreement counting equal-t ties {tie.mean() * 100:.4f} %")
    assert tie.mean() >= 0.999 and face_same.mean() >= (0.995 if name == "models" else 0.999)
    ys, xs = np.nonzero(~same & ~q13)
    g = np.pad(gold["leaf"], 1, mode="edge")
    for y, x in zip(ys, xs):
        assert leaf[y, x] in g[y:y + 3, x:x + 3], (y, x)


@pytest.mark.parametrize("name", STATS_SCENES)
def test_scene_radiance_converges_to_reference(name, gpu_ctx, tmp_path):
    """Converged radiance against 16 x 256 spp of the reference.  Bar (north_star): image-level relative
    RMSE <= 1 % beyond the Monte-Carlo noise of the two estimates, per-pixel means within 3 sigma, no bias."""
    _, _, _, wh, _ = GOLDEN_SCENES[name]
    gold = np.load(os.path.join(GOLDEN, f"{name}_{wh[0]}x{wh[1]}_stats.npz"))
    ref, var_ref = gold["mean"].astype(np.float64), gold["var_of_mean"].astype(np.float64)
    sc, cfg, digest = _load_scene(gpu_ctx, tmp_path, name, wh, 256)
    assert digest == str(gold["digest"]), "the regenerated scene differs from the one the reference rendered"
    batches = int(gold["batches"])
    cfg.samples_per_pixel = 65536 if batches >= 64 else 16384  # 4x the reference's samples
    img = sc.render(cfg, seed=11)["beauty"]
    sc.close()
    assert np.isfinite(img).all()
    noise2 = 1.25 * var_ref.sum() / (ref ** 2).sum()   # reference noise + GPU noise at 4x the samples
    rrmse = np.sqrt(((img - ref) ** 2).sum() / (ref ** 2).sum())
    print(f"{name}: rRMSE {rrmse * 100:.3f} % (noise floor {np.sqrt(noise2) * 100:.3f} %)  mean gpu "
          f"{img.mean(axis=(0, 1))} ref {ref.mean(axis=(0, 1))}")
    if batches >= 64:
        # the plain north_star bar on 64 x 256 spp reference fixtures (round 2).  Where even 16 384 reference samples per
        # pixel leave more than 0.4 % of noise (Cornell: a small light, fog, no light sampling in the reference) the images
        # are compared as k x k block means, which divides the noise of both by k and keeps any bias
        k = 1
        while np.sqrt(noise2) / k > 0.004 and k < 8:
            k += 1
        h2, w2 = (img.shape[0] // k) * k, (img.shape[1] // k) * k
        blk = lambda a: a[:h2, :w2].reshape(h2 // k, k, w2 // k, k, 3).mean(axis=(1, 3))
        rr = np.sqrt(((blk(img) - blk(ref)) ** 2).sum() / (blk(ref) ** 2).sum())
        print(f"{name}: rRMSE of {k}x{k} block means {rr * 100:.3f} %")
        assert rr <= 0.01, (rr, k)
    else:
        assert rrmse <= np.sqrt(0.01 ** 2 + noise2)
    sigma = np.sqrt(1.25 * var_ref + 1e-12)
    out = np.abs(img - ref) > 3 * sigma
    print(f"{name}: outside 3 sigma: {out.mean() * 100:.3f} % of pixel channels")
    assert out.mean() <= (0.01 if batches >= 64 else 0.02)  # 0.27 % expected by chance
    m_gpu, m_ref = img.mean(axis=(0, 1)), ref.mean(axis=(0, 1))
    sem = np.sqrt(1.25 * var_ref.sum(axis=(0, 1))) / (ref.shape[0] * ref.shape[1])
    assert np.all(np.abs(m_gpu - m_ref) <= 0.005 * m_ref + 4 * sem), (m_gpu, m_ref, sem)


def test_quirks_aovs_match_reference(gpu_ctx, tmp_path):
    """First-hit AOVs (albedo, view-space normal, z-depth; camera.hpp:464-488) against the reference's buffers.
    Both sides average min(aux, spp) = 64 jittered samples, so only silhouette pixels may differ by more than
    the jitter noise."""
    gold = np.load(os.path.join(GOLDEN, "quirks_320x180_aov.npz"))
    txt, cfg, files = scenes.quirks(seed=5, width=320, height=180, spp=512)
    cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = True
    cfg.z_depth_max_dist = 15.0
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files))
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    r = sc.render(cfg, seed=5)
    sc.close()
    for key in ("albedo", "normal", "z_depth"):
        diff = np.abs(r[key] - gold[key].astype(np.float64)).max(axis=2)
        print(f"{key}: mean |diff| {diff.mean():.5f}, 99th pct {np.quantile(diff, 0.99):.4f}, max {diff.max():.3f}")
        # checker / silhouette pixels carry binomial jitter noise (64 samples); no systematic offset allowed
        signed = (r[key] - gold[key].astype(np.float64)).mean()
        assert diff.mean() < 8e-3 and np.quantile(diff, 0.95) < 0.05 and abs(signed) < 1e-3, (key, signed)


# ------------------------------------------------------------------ per-ray known answers (zb_probe_rays)
import ctypes as C  # noqa: E402


def _leaf_has_uv(flat):
    """leaf id -> True when the reference fills rec.u / rec.v for that object (spheres and cubes; triangles leave
    them stale, triangle.hpp:75-79)."""
    f = flat.contents
    items = np.ctypeslib.as_array(C.cast(f.items, C.POINTER(C.c_uint32)), (f.n_items, 8))
    inst = (np.ctypeslib.as_array(C.cast(f.instances, C.POINTER(C.c_uint32)), (f.n_instances, 32))
            if f.n_instances else np.zeros((1, 32), np.uint32))
    has = np.zeros(f.n_leaves, bool)
    for row in items:
        kind, leaf, aux = row[4] & 0xFF, row[6], row[7]
        if kind == 3:
            kind = {0: 0, 1: 1, 2: 2, 3: 2}[int(inst[aux, 24])]
        has[leaf] = kind in (0, 1)
    return has


PROBE_SCENES = ["weekend"] + list(GOLDEN_SCENES)


@pytest.mark.parametrize("name", PROBE_SCENES)
def test_probe_hit_records_match_reference(name, gpu_ctx, tmp_path):
    """6000 rays per scene (camera rays + random rays, many starting inside objects): leaf id, t, hit point,
    normal, front_face (Appendix A q1/q2/q5), u/v, material::get_albedo, emitted and the background colour of
    misses against the reference's hit_record for the same float32 rays."""
    gold = np.load(os.path.join(GOLDEN, f"{name}_probe.npz"))
    if name == "weekend":
        txt, cfg = scenes.weekend(seed=1, width=1200, height=675)
        files = {}
    else:
        gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
        txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files))
    has_uv = _leaf_has_uv(flat)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    r = sc.probe(cfg, gold["o"], gold["d"])
    sc.close()
    ref = gold["rec"]
    ref_hit = np.isfinite(ref[:, 0])
    ref_leaf = np.where(ref_hit, ref[:, 1], -1).astype(np.int64)
    gpu_leaf = np.where(r["leaf_id"] == 0xFFFFFFFF, -1, r["leaf_id"].astype(np.int64))
    same = ref_leaf == gpu_leaf
    print(f"{name}: leaf agreement {same.mean() * 100:.3f} % of {len(same)} rays ({ref_hit.mean() * 100:.1f} % hits)")
    assert same.mean() >= 0.998
    h = same & ref_hit
    scale = np.maximum(np.abs(ref[h, 4:7]).max(axis=1), 1.0)
    rel_t = np.abs(r["t"][h] - ref[h, 0]) / np.maximum(np.abs(ref[h, 0]), 1e-6)
    dp = np.abs(r["p"][h] - ref[h, 4:7]).max(axis=1) / scale
    dn = np.abs(r["normal"][h] - ref[h, 7:10]).max(axis=1)
    ff = r["front_face"][h] == ref[h, 3].astype(np.uint32)
    print(f"  t rel q99.9 {np.quantile(rel_t, 0.999):.2e}  p q99.9 {np.quantile(dp, 0.999):.2e}  "
          f"normal q99 {np.quantile(dn, 0.99):.2e}  front_face {ff.mean() * 100:.3f} %")
    assert np.quantile(rel_t, 0.999) < 5e-4 and np.quantile(dp, 0.999) < 1e-3
    assert np.quantile(dn, 0.99) < 5e-3 and ff.mean() >= 0.999
    uvm = h.copy()
    uvm[h] = has_uv[ref_leaf[h]]
    if uvm.any():
        du = np.abs(r["u"][uvm] - ref[uvm, 10])
        du = np.minimum(du, 1.0 - du)
        dv = np.abs(r["v"][uvm] - ref[uvm, 11])
        print(f"  uv: {uvm.sum()} rays, u q99 {np.quantile(du, 0.99):.2e} v q99 {np.quantile(dv, 0.99):.2e}")
        assert np.quantile(du, 0.99) < 2e-3 and np.quantile(dv, 0.99) < 2e-3
    da = np.abs(r["albedo"][h] - ref[h, 12:15]).max(axis=1)
    de = np.abs(r["emitted"][h] - ref[h, 15:18]).max(axis=1)
    print(f"  albedo within 2e-3: {(da < 2e-3).mean() * 100:.2f} %  emitted max diff {de.max():.2e}")
    assert (da < 2e-3).mean() >= 0.98 and de.max() < 1e-3
    m = same & ~ref_hit
    if m.any():
        bg_ref = ref[m, 12:15]
        db = np.abs(r["albedo"][m] - bg_ref).max(axis=1) / np.maximum(bg_ref.max(axis=1), 1e-3)
        print(f"  background within 1e-3 rel: {(db < 1e-3).mean() * 100:.2f} % of {m.sum()} misses")
        assert (db < 1e-3).mean() >= 0.98


# ------------------------------------------------------------------ live oracle on fresh seeded rays
ZORACLE = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_build", "zoracle")


@pytest.mark.parametrize("name,seed", [("weekend", 101), ("quirks", 102), ("exterior", 103)])
def test_probe_matches_live_oracle_on_fresh_rays(name, seed, gpu_ctx, tmp_path):
    """The CUDA closest-hit + hit-record path against the CPU restatement (oracle/zoracle.cpp, itself pinned on the
    reference's fixtures) run HERE on 20 000 rays nobody has seen before: origins in and around the scene,
    random directions of random length (directions are never normalised, Appendix A q7)."""
    import subprocess
    zo = os.path.abspath(ZORACLE)
    if not os.path.exists(zo):
        subprocess.run(["make", "-C", os.path.dirname(os.path.dirname(zo)), "port"], check=True, capture_output=True)
    if name == "weekend":
        txt, cfg = scenes.weekend(seed=1, width=64, height=36)
        files = {}
    else:
        gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
        txt, cfg, files = getattr(scenes, gen)(width=64, height=36, spp=1, **{**kw, **ids_kw})
    path = scenes.materialize(str(tmp_path), txt, cfg, files)
    os.makedirs(os.path.join(str(tmp_path), "assets", "hdr_maps"), exist_ok=True)
    rng = np.random.default_rng(seed)
    n = 20000
    o = (rng.normal(size=(n, 3)) * np.array([6.0, 1.5, 6.0]) + np.array([0.0, 1.5, 0.0])).astype(np.float32)
    d = (rng.normal(size=(n, 3)) * rng.uniform(0.1, 12.0, size=(n, 1))).astype(np.float32)
    with open(os.path.join(str(tmp_path), "rays.bin"), "wb") as f:
        f.write(np.int32(n).tobytes() + o.tobytes() + d.tobytes())
    subprocess.run([zo, "probe", os.path.basename(path), "s.cfg", "rays.bin", "probe.bin"], cwd=str(tmp_path), check=True,
                   capture_output=True)
    ref = np.frombuffer(open(os.path.join(str(tmp_path), "probe.bin"), "rb").read()[4:], np.float64).reshape(n, 19)
    flat = api.flatten_zscene(path=path)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    r = sc.probe(cfg, o, d)
    sc.close()
    ref_hit = np.isfinite(ref[:, 0])
    ref_leaf = np.where(ref_hit, ref[:, 1], -1).astype(np.int64)
    gpu_leaf = np.where(r["leaf_id"] == 0xFFFFFFFF, -1, r["leaf_id"].astype(np.int64))
    same = ref_leaf == gpu_leaf
    print(f"{name}: leaf agreement {same.mean() * 100:.3f} % of {n} fresh rays ({ref_hit.mean() * 100:.1f} % hits)")
    assert same.mean() >= 0.998
    h = same & ref_hit
    rel_t = np.abs(r["t"][h] - ref[h, 0]) / np.maximum(np.abs(ref[h, 0]), 1e-6)
    dn = np.abs(r["normal"][h] - ref[h, 7:10]).max(axis=1)
    ff = r["front_face"][h] == ref[h, 3].astype(np.uint32)
    assert np.quantile(rel_t, 0.999) < 5e-4 and np.quantile(dn, 0.99) < 5e-3 and ff.mean() >= 0.999
    m = same & ~ref_hit
    if m.any():
        bg_ref = ref[m, 12:15]
        db = np.abs(r["albedo"][m] - bg_ref).max(axis=1) / np.maximum(bg_ref.max(axis=1), 1e-3)
        assert (db < 1e-3).mean() >= 0.98


def test_light_passes_match_reference(gpu_ctx, tmp_path):
    """reflection_buffer / refraction_buffer (camera.hpp:492-517) through zb_render: 16 384 spp on the GPU against
    the reference's 2048-spp buffers (fixture); the beauty pass of the same call must be unaffected."""
    gold = np.load(os.path.join(GOLDEN, "quirks_200x112_light.npz"))
    txt, cfg, files = scenes.quirks(seed=5, width=200, height=112, spp=2048)
    cfg.use_reflection = cfg.use_refraction = True
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    cfg.samples_per_pixel = 16384
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files))
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    res = sc.render(cfg, seed=11)
    cfg.use_reflection = cfg.use_refraction = False
    plain = sc.render(cfg, seed=11)
    sc.close()
    assert np.array_equal(plain["beauty"], res["beauty"])
    for key in ("reflection", "refraction"):
        ref, img = gold[key].astype(np.float64), res[key]
        assert np.isfinite(img).all()
        rrmse = float(np.sqrt(((img - ref) ** 2).sum() / (ref ** 2).sum()))
        print(f"{key}: mean gpu {img.mean():.5f} ref {ref.mean():.5f} rRMSE {rrmse * 100:.2f} %")
        assert abs(img.mean() / ref.mean() - 1.0) < 0.01
        assert rrmse < 0.06  # the 2048-spp reference carries ~5 % noise of its own on these sparse buffers


@pytest.mark.parametrize("devices", ["", "0,0,0"])
def test_drop_in_camera_render_equals_abi_render(weekend_scene, tmp_path, devices):
    """The reference-shaped C++ entry point zenith::camera::render (host/zenith/camera.hpp; camera.hpp:236) fills its
    std::vector<color> buffers with exactly what zb_render returns for the same scene, camera and seed; the same
    program also renders a hand-built scene written like the reference's scene code and checks lines_rendered /
    render_flag behaviour."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "drop_in_render")
    subprocess.run(["g++", "-std=c++20", "-O2", "-Wno-unused-function", os.path.join(root, "tests", "cpp", "drop_in_render.cpp"), "-o", exe,
                    "-L" + os.path.join(root, "raytracer_project_b200"), "-lzenith_b200",
                    "-Wl,-rpath," + os.path.join(root, "raytracer_project_b200")], check=True)
    W, H, SPP, DEPTH, SEED = 160, 90, 16, 20, 77
    txt, cfg = scenes.weekend(seed=1, width=W, height=H, spp=SPP, max_depth=DEPTH)
    cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = cfg.use_reflection = cfg.use_refraction = True
    (tmp_path / "w.zscene").write_text(txt)
    out = tmp_path / "out.bin"
    # devices = "0,0,0": zenith::camera::devices -> zb_context_create_multi (three ranks, here on one GPU): the same buffers
    r = subprocess.run([exe, str(tmp_path / "w.zscene"), str(out), str(W), str(H), str(SPP), str(DEPTH), str(SEED)] + ([devices] if devices else []),
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    bufs = np.frombuffer(out.read_bytes(), np.float64).reshape(6, H, W, 3)
    res = weekend_scene.render(cfg, seed=SEED)
    for k, key in enumerate(["beauty", "albedo", "normal", "z_depth", "reflection", "refraction"]):
        assert np.array_equal(bufs[k], res[key]), key
    assert bufs[0].mean() > 0.1 and bufs[4].mean() > 0


# ------------------------------------------------------------------ finishing chain
def test_finishing_chain_matches_reference(gpu_ctx):
    """k_bloom / k_finish (fp32) through zb_finish against the reference's final_framebuffer (double) for every
    parameter set of tests/golden/finish_cases.py: display-referred values within 5e-4 (1/8 of an 8-bit step) on
    >= 99.7 % of the channels — the rest sit on the discontinuities of the chain (hue sector edges, the luminance
    debug bands, the bloom threshold), where an fp32 rounding picks the other side."""
    from tests.golden.finish_cases import CASES, hdr_image
    gold = np.load(os.path.join(GOLDEN, "finish_96x60.npz"))
    img = hdr_image()
    for name, (pass_kind, kw) in CASES.items():
        kw = dict(kw)
        if kw.get("use_auto_exposure"):
            kw["exposure"] = float(gold[name + "_stats"][2])  # render() stores the auto-exposure back (camera.hpp:263)
        out = gpu_ctx.finish(img, api.post_struct(**kw), pass_kind)
        ref = gold[name + "_image"]
        diff = np.abs(out - ref)
        ok = diff < 5e-4
        print(f"{name}: max diff {diff.max():.2e}, within 5e-4: {ok.mean() * 100:.3f} %")
        assert ok.mean() >= 0.997, (name, ok.mean())
        assert np.median(diff) < 2e-6


def test_analyze_and_auto_exposure_match_reference(gpu_ctx):
    """k_luma_stats + zb_auto_exposure against analyze_framebuffer / apply_auto_exposure of the reference."""
    import torch
    from tests.golden.finish_cases import CASES, hdr_image
    gold = np.load(os.path.join(GOLDEN, "finish_96x60.npz"))
    img = hdr_image()
    f4 = np.concatenate([img, np.ones(img.shape[:2] + (1,))], -1).astype(np.float32)
    dev = torch.from_numpy(f4).cuda().contiguous()
    st = gpu_ctx.analyze_device(dev.data_ptr(), img.shape[0] * img.shape[1])
    avg, mx, _ = gold["default_stats"]
    assert abs(st["average_luminance"] / avg - 1) < 1e-5 and abs(st["max_luminance"] / mx - 1) < 1e-6
    moved = np.abs(st["histogram"].astype(np.int64) - gold["histogram"]).sum()
    assert st["histogram"].sum() == img.shape[0] * img.shape[1] and moved <= 4, moved  # a pixel on a bin edge moves 2
    _, kw = CASES["hsv_auto_exposure"]
    ev = gpu_ctx.auto_exposure(api.post_struct(**kw), st)
    assert abs(ev / gold["hsv_auto_exposure_stats"][2] - 1) < 1e-5


def test_finish_device_rgba8_and_float_agree(gpu_ctx):
    """zb_finish_device on a 4K device frame: float4 and RGBA8 outputs of the same launch agree under the reference's
    8-bit rule (camera.hpp:773-775), the result is deterministic, and bloom leaves a zero frame at zero."""
    import torch
    W, H = 3840, 2160
    g = torch.Generator(device="cuda").manual_seed(5)
    src = torch.rand(H * W, 4, device="cuda", generator=g) * 3.0
    post = api.post_struct(use_bloom=1, use_sharpening=1, use_aces=1)
    out_f4 = torch.empty(H * W, 4, device="cuda")
    out_u8 = torch.empty(H * W, 4, device="cuda", dtype=torch.uint8)
    gpu_ctx.finish_device(src.data_ptr(), W, H, post, 0, out_f4.data_ptr(), out_u8.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal((out_f4[:, :3] * 255.999).to(torch.uint8), out_u8[:, :3]) and bool((out_u8[:, 3] == 255).all())
    again = torch.empty_like(out_f4)
    gpu_ctx.finish_device(src.data_ptr(), W, H, post, 0, again.data_ptr(), 0)
    torch.cuda.synchronize()
    assert torch.equal(again, out_f4)
    zero = torch.zeros(H * W, 4, device="cuda")
    gpu_ctx.finish_device(zero.data_ptr(), W, H, post, 0, again.data_ptr(), 0)
    torch.cuda.synchronize()
    assert float(again[:, :3].abs().max()) == 0.0


# ------------------------------------------------------------------ device-built BLAS
@pytest.mark.parametrize("builder", [1, 0], ids=["device_sah", "device_lbvh"])
def test_gpu_built_blas_gives_the_same_hits(gpu_ctx, tmp_path, builder):
    """Mesh BLAS built on the device (binned SAH, csrc/sah_build.cu, or the linear BVH of csrc/lbvh.cu) instead of the host SAH builder: closest-hit answers
    do not depend on the tree, so primary ids and per-ray hit records must match the reference fixtures exactly like
    the host-built scene does, and the rendered image must agree with the host-built one."""
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES["mesh"]
    gold = np.load(os.path.join(GOLDEN, f"mesh_{wh[0]}x{wh[1]}_ids.npz"))
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    path = scenes.materialize(str(tmp_path), txt, cfg, files)
    built = {}
    gpu_ctx.set_tuning("gpu_builder", builder)
    try:
        for mode in (False, True):
            flat = api.flatten_zscene(path=path, gpu_blas=mode)
            if mode:
                assert all(flat.contents.meshes_array()[m].node_count == 0 for m in range(flat.contents.n_meshes))
            built[mode] = gpu_ctx.upload(flat)
            api.free_flat(flat)
    finally:
        gpu_ctx.set_tuning("gpu_builder", 1)
    leaf, prim, t = built[True].primary_ids(cfg)
    assert (leaf == gold["leaf"]).mean() >= 0.9999
    both = (leaf == gold["leaf"]) & (leaf != 0xFFFFFFFF)
    assert (prim[both] == gold["tri"][both]).mean() >= 0.999  # shared edges hit both triangles at the same t
    pg = np.load(os.path.join(GOLDEN, "mesh_probe.npz"))
    r0, r1 = built[False].probe(cfg, pg["o"], pg["d"]), built[True].probe(cfg, pg["o"], pg["d"])
    assert (r0["leaf_id"] == r1["leaf_id"]).mean() >= 0.9995
    same = r0["leaf_id"] == r1["leaf_id"]
    assert np.allclose(r0["t"][same], r1["t"][same], rtol=1e-5, atol=1e-6, equal_nan=True)
    # tree quality through the renderer's own counters: node visits + triangle tests per ray against the host SAH tree
    cfg.samples_per_pixel = 4
    work = {}
    for mode in (False, True):
        built[mode].render(cfg, seed=3, stats=True)
        st = gpu_ctx.stats()
        assert st["stack_overflows"] == 0
        work[mode] = (st["node_visits"] + st["prim_tests"]) / max(st["rays"], 1)
    print(f"traversal work per ray: host SAH {work[False]:.2f}, device builder {builder}: {work[True]:.2f}")
    assert work[True] <= work[False] * (1.15 if builder == 1 else 2.5)
    cfg.samples_per_pixel = 256
    a = built[False].render(cfg, seed=3)["beauty"]
    b = built[True].render(cfg, seed=3)["beauty"]
    for sc in built.values():
        sc.close()
    rrmse = float(np.sqrt(((a - b) ** 2).sum() / (a ** 2).sum()))
    print(f"host-SAH vs device-built image, same seed: rRMSE {rrmse * 100:.3f} %, means {a.mean():.5f} {b.mean():.5f}")
    assert abs(a.mean() / b.mean() - 1) < 0.005 and rrmse < 0.05


@pytest.mark.parametrize("name", ["mesh_full", "big"])
def test_device_sah_builder_at_full_size_matches_reference_ids(name, gpu_ctx, tmp_path):
    """The 1 M triangle (config C3) and 10 M triangle (config C5) fixtures with the BLAS built on the device: the same
    leaf and face ids as the reference rendered, and no more traversal work than the host-built SAH tree does."""
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
    gold = np.load(os.path.join(GOLDEN, f"{name}_{wh[0]}x{wh[1]}_ids.npz"))
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    path = scenes.materialize(str(tmp_path), txt, cfg, files)
    flat = api.flatten_zscene(path=path, gpu_blas=True)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    leaf, prim, t = sc.primary_ids(cfg)
    cfg.samples_per_pixel = 2
    sc.render(cfg, seed=3, stats=True)
    st = gpu_ctx.stats()
    sc.close()
    assert st["stack_overflows"] == 0
    same = leaf == gold["leaf"]
    print(f"{name} (device SAH): leaf-id agreement {same.mean() * 100:.5f} %, {st['node_visits'] / st['rays']:.2f} node visits and "
          f"{st['prim_tests'] / st['rays']:.2f} triangle tests per ray")
    assert same.mean() >= 0.9999
    hit = same & (gold["leaf"] != 0xFFFFFFFF)
    rel = np.abs(t[hit] - gold["t"][hit]) / np.maximum(gold["t"][hit], 1e-6)
    assert np.quantile(rel, 0.999) < 2e-4 and rel.max() < 0.1
    tie = ((prim == gold["tri"])[hit]) | (rel < 1e-5)
    assert tie.mean() >= 0.999


def test_device_sah_builder_on_small_and_shipped_meshes(gpu_ctx, tmp_path):
    """sah_build.cu on the shipped OBJ models (tens to thousands of triangles each, several meshes in one scene: the
    warp phase alone finishes the small ones): primary ids and hit distances equal the host-built scene's, twice in a row
    (the build is deterministic: same tree, same frame bits)."""
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES["models"]
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=1, **{**kw, **ids_kw})
    path = scenes.materialize(str(tmp_path), txt, cfg, files)
    flat = api.flatten_zscene(path=path)
    host = gpu_ctx.upload(flat)
    api.free_flat(flat)
    leaf0, prim0, t0 = host.primary_ids(cfg)
    cfg.samples_per_pixel = 8
    frames = []
    for _ in range(2):
        flat = api.flatten_zscene(path=path, gpu_blas=True)
        dev = gpu_ctx.upload(flat)
        api.free_flat(flat)
        leaf1, prim1, t1 = dev.primary_ids(cfg)
        assert (leaf1 == leaf0).mean() >= 0.9999
        same = (leaf1 == leaf0) & (leaf0 != 0xFFFFFFFF)
        assert np.allclose(t0[same], t1[same], rtol=1e-5, atol=1e-6)
        assert (prim1[same] == prim0[same]).mean() >= 0.995  # coincident faces of bowl.obj tie at equal t
        frames.append(dev.render(cfg, seed=11)["beauty"].copy())
        dev.close()
    host.close()
    assert np.array_equal(frames[0], frames[1])


# ------------------------------------------------------------------ on-disk formats
def test_checkpoint_resume_and_scene_file_reproduce_the_frame(gpu_ctx, tmp_path):
    """Progressive render stopped after 32 of 64 samples, checkpointed (ZBCK), resumed in a scene re-loaded from its
    ZBFS file: the sums equal those of an uninterrupted 2 x 32-sample render bit for bit."""
    import torch
    txt, cfg = scenes.weekend(seed=1, width=320, height=180, spp=64)
    flat = api.flatten_zscene(txt)
    api.save_flat(flat, str(tmp_path / "w.zbfs"))
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    n = 320 * 180
    whole = torch.zeros(n, 4, device="cuda")
    sc.render_device(cfg, whole.data_ptr(), seed=9, samples_per_launch=32)
    part = torch.zeros(n, 4, device="cuda")
    sc.render_device(cfg, part.data_ptr(), seed=9, sample_range=(0, 32))
    gpu_ctx.checkpoint_save(str(tmp_path / "half.zbck"), part.data_ptr(), 320, 180, 32, 9)
    sc.close()
    flat2 = api.load_flat(str(tmp_path / "w.zbfs"))
    sc2 = gpu_ctx.upload(flat2)
    api.free_flat(flat2)
    resumed = torch.full((n, 4), 7.0, device="cuda")
    done, seed = gpu_ctx.checkpoint_load(str(tmp_path / "half.zbck"), resumed.data_ptr(), 320, 180)
    assert (done, seed) == (32, 9)
    sc2.render_device(cfg, resumed.data_ptr(), seed=seed, sample_range=(done, 64), accumulate=True)
    torch.cuda.synchronize()
    sc2.close()
    assert torch.equal(resumed, whole)
    assert float(whole[:, 3].min()) == 64.0 and float(whole[:, :3].mean()) > 1.0


@pytest.mark.parametrize("name", ["weekend"] + SMALL_SCENES)
def test_wavefront_render_equals_megakernel(name, gpu_ctx, tmp_path, weekend_scene, monkeypatch):
    """(experiments build only) The wavefront kernels (wavefront.cu, tuning key "wavefront") draw the same random numbers and run the same shading
    code as k_render, and both sum paths in 64-bit fixed point: per-path radiance may differ only by the
    compiler's contraction choices, so the frames agree to fp32 rounding."""
    if not api.has_experiments():
        pytest.skip("library built without -DZB_EXPERIMENTS (the wavefront form is a rejected experiment)")
    if name == "weekend":
        sc, cfg = weekend_scene, scenes.weekend(seed=1, width=203, height=90, spp=24)[1]  # 203: ragged tile edge
    else:
        wh = GOLDEN_SCENES[name][3] or (200, 112)
        sc, cfg, _ = _load_scene(gpu_ctx, tmp_path, name, wh, 24)
    gpu_ctx.set_tuning("wavefront", 0)
    a = sc.render(cfg, seed=5)["beauty"]
    gpu_ctx.set_tuning("wavefront", 1)
    gpu_ctx.set_tuning("wf_slots", 4096)  # far fewer slots than paths: regeneration and many iterations
    b = sc.render(cfg, seed=5)["beauty"]
    c = sc.render(cfg, seed=5)["beauty"]
    gpu_ctx.set_tuning("wf_slots", 0)
    d = sc.render(cfg, seed=5)["beauty"]
    gpu_ctx.set_tuning("wavefront", 0)
    if name != "weekend":
        sc.close()
    assert np.array_equal(b, c) and np.array_equal(b, d), "wavefront frame depends on scheduling"
    same = np.mean(a == b)
    print(f"{name}: wavefront == megakernel on {same * 100:.2f} % of pixel channels, max |diff| {np.abs(a - b).max():.3g}")
    # a contraction difference can flip a roulette or scatter decision of single paths: allow a few pixels
    bad = np.abs(a - b) > 1e-5 + 1e-5 * np.abs(a)
    assert bad.mean() < 2e-3, bad.mean()
    assert abs(a.mean() - b.mean()) < 1e-3 * a.mean()


def test_lean_instantiation_equals_general_queue_kernel(weekend_scene, gpu_ctx):
    """Plain scenes (every item an inline sphere, untextured materials, solid-colour environment) run the LEAN
    instantiation of k_render_pq; ZB_NO_LEAN=1 forces the general one.  Same arithmetic on the paths a plain scene
    takes, so the frames are equal bit for bit, for every resident-blocks variant."""
    cfg = scenes.weekend(seed=1, width=203, height=90, spp=24)[1]
    gpu_ctx.set_tuning("no_lean", 1)
    a = weekend_scene.render(cfg, seed=9)["beauty"]
    gpu_ctx.set_tuning("no_lean", 0)
    b = weekend_scene.render(cfg, seed=9)["beauty"]
    assert np.array_equal(a, b)
    for blocks in (6, 7):
        gpu_ctx.set_tuning("pq_blocks", blocks)
        assert np.array_equal(a, weekend_scene.render(cfg, seed=9)["beauty"]), blocks
    gpu_ctx.set_tuning("pq_blocks", 0)
    if api.has_experiments():
        # resumable traversal (k_render_pqr, a rejected experiment): other scheduling of the same paths
        for thresh in (8, 24):
            gpu_ctx.set_tuning("pq_resume", thresh)
            assert np.array_equal(a, weekend_scene.render(cfg, seed=9)["beauty"]), thresh
        gpu_ctx.set_tuning("pq_resume", 0)
    # the same spheres under the procedural sky: the "spheres + textures / environment" set against the general kernel
    import copy
    sky = copy.copy(cfg)
    sky.env_mode = "sun"
    full = copy.copy(sky)
    full.use_albedo = full.use_normal = full.use_z_depth = full.use_reflection = full.use_refraction = True
    b = weekend_scene.render(full, seed=9)
    gpu_ctx.set_tuning("no_lean", 1)
    a = weekend_scene.render(full, seed=9)
    gpu_ctx.set_tuning("no_lean", 0)
    for key in ("beauty", "albedo", "normal", "z_depth", "reflection", "refraction"):
        assert np.array_equal(a[key], b[key]), key
    assert not np.array_equal(a["beauty"], weekend_scene.render(cfg, seed=9)["beauty"])  # the sky is not the solid colour


# ------------------------------------------------------------------ image export
def test_save_pass_png_matches_reference_export(gpu_ctx, tmp_path):
    """zb_save_pass (finishing chain of the pass on the device -> 8-bit PNG) against the pixels of the PNGs the
    reference's save_render_pass wrote for the same frame (tests/golden/export_96x60.npz): within one 8-bit step,
    equal on >= 97 % of the channels (fp32 chain vs double; 255.999 * v truncates, so a rounding difference next to
    an integer moves a whole step)."""
    import torch
    from tests.golden.finish_cases import EXPORT_CASES, hdr_image
    gold = np.load(os.path.join(GOLDEN, "export_96x60.npz"))
    img = hdr_image()
    H, W = img.shape[:2]
    f4 = np.concatenate([img, np.ones((H, W, 1))], -1).astype(np.float32)
    dev = torch.from_numpy(f4).cuda().contiguous()
    for name, (rp, kw) in EXPORT_CASES.items():
        fname = api.pass_file_name(rp)
        assert fname == str(gold[name + "_file"])
        path = str(tmp_path / fname)
        gpu_ctx.save_pass(path, dev.data_ptr(), W, H, api.post_struct(**kw), rp)
        out = api.read_png(path).astype(np.int32)
        ref = gold[name].astype(np.int32)
        diff = np.abs(out - ref)
        print(f"{name}: equal {np.mean(diff == 0) * 100:.2f} %, max step {diff.max()}")
        assert out.shape == ref.shape
        assert (diff <= 1).mean() >= 0.999 and (diff == 0).mean() >= 0.97, (name, diff.max(), (diff == 0).mean())
    # linear radiance as Radiance RGBE: what the reference's stbi_loadf would read back as an environment
    hdr_path = str(tmp_path / "frame.hdr")
    gpu_ctx.save_hdr(hdr_path, dev.data_ptr(), W, H)
    got = api.read_hdr(hdr_path)
    assert np.all(np.abs(got - f4[..., :3]) <= f4[..., :3].max(-1, keepdims=True) / 128.0 + 1e-30)


# ------------------------------------------------------------------ BVH debug view
def _wire_mean(sc, cfg, level, thickness, seed=3):
    import torch
    n = cfg.image_width * cfg.image_height
    acc = torch.empty(n, 4, device="cuda")
    sc.render_bvh_debug_device(cfg, acc.data_ptr(), level=level, thickness=thickness, seed=seed)
    torch.cuda.synchronize()
    a = acc.cpu().numpy().reshape(cfg.image_height, cfg.image_width, 4)
    assert np.all(a[..., 3] == cfg.samples_per_pixel)
    return a[..., :3] / cfg.samples_per_pixel


def test_bvh_debug_view_level0_matches_reference(gpu_ctx):
    """k_wireframe at level 0 (the root box: the one level that does not depend on how the tree was built) against the
    reference rendered with global_settings::bvh_debug_mode (tests/golden/wire_240x135.npz, 256 spp): frame, volume
    and background coverage per pixel agree within the sampling noise of the pixel jitter."""
    gold = np.load(os.path.join(GOLDEN, "wire_240x135.npz"))
    txt, cfg, files = scenes.wire_demo(spp=256)
    assert scene_digest(txt, cfg, files) == str(gold["digest"])
    flat = api.flatten_zscene(txt)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    cfg.samples_per_pixel = 1024
    for name in ("level0", "level0_thin"):
        level, thick = gold[name + "_params"]
        img = _wire_mean(sc, cfg, int(level), float(thick))
        ref = gold[name].astype(np.float64)
        diff = np.abs(img - ref)
        frame_ref, frame_gpu = (ref[..., 2] > 2.0).mean(), (img[..., 2] > 2.0).mean()
        print(f"{name}: mean |diff| {diff.mean():.4f}, > 0.5: {(diff > 0.5).mean() * 100:.3f} %, frame pixels ref {frame_ref:.4f} gpu {frame_gpu:.4f}")
        assert diff.mean() < 0.01 and (diff > 0.5).mean() < 2e-3
        assert abs(frame_gpu - frame_ref) < 0.1 * frame_ref + 1e-3
        assert np.allclose(img.mean(axis=(0, 1)), ref.mean(axis=(0, 1)), atol=0.01)
    sc.close()


def test_bvh_debug_view_levels_use_the_reference_palette(gpu_ctx, tmp_path):
    """Levels below the root depend on the tree (the reference's own is random, bvh.hpp:17), so they are checked
    through what the reference fixes: with one sample per pixel every pixel is the background, a frame colour
    4 (0.4, g, 1 - g) or a volume colour 0.1 (0.4, g, 1 - g) with g = min(0.15 depth, 1) of the requested level;
    deeper levels draw more, smaller frames; a mesh's own tree starts again at depth 0."""
    txt, cfg, files = scenes.wire_demo(spp=1)
    flat = api.flatten_zscene(txt)
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    bg = np.array(cfg.background, np.float32)

    def classes(img, depths):
        px = img.reshape(-1, 3)
        kind = np.full(len(px), -1)
        kind[np.all(np.abs(px - bg) < 1e-5, axis=1)] = 0
        for dd in depths:
            g = min(dd * 0.15, 1.0)
            base = np.array([0.4, g, 1.0 - g], np.float32)
            kind[np.all(np.abs(px - 4.0 * base) < 1e-5, axis=1)] = 1
            kind[np.all(np.abs(px - 0.1 * base) < 1e-6, axis=1)] = 2
        return kind

    frames = {}
    for level in (0, 1, 2, 3):
        k = classes(_wire_mean(sc, cfg, level, 0.05), [level])
        # below level 1 every primitive sits under a level node; deeper, a shallow leaf may not (it then shades normally)
        if level <= 1:
            assert (k >= 0).all(), (level, (k < 0).mean())
        frames[level] = (k == 1).mean()
        print(f"level {level}: frame {frames[level]:.4f} volume {(k == 2).mean():.4f} background {(k == 0).mean():.4f} other {(k < 0).mean():.4f}")
        assert frames[level] > 0.001 and (k == 0).mean() > 0.3
    leaves = classes(_wire_mean(sc, cfg, -1, 0.05), range(0, 12))
    print(f"leaves: frame {(leaves == 1).mean():.4f} volume {(leaves == 2).mean():.4f}")
    assert (leaves >= 0).all() and (leaves == 1).mean() > 0.001 and (leaves == 2).mean() > 0.0
    assert np.array_equal(_wire_mean(sc, cfg, 2, 0.02), _wire_mean(sc, cfg, 2, 0.02))
    assert (classes(_wire_mean(sc, cfg, 2, 0.05), [2]) == 1).mean() > (classes(_wire_mean(sc, cfg, 2, 0.01), [2]) == 1).mean()
    sc.close()
    # a scene with a mesh: the mesh's tree shows its own level-0 box in object space
    for name in ("mesh", "quirks"):
        sc2, cfg2, _ = _load_scene(gpu_ctx, tmp_path, name, (160, 90), 1)
        for level in (0, 4, -1):
            img = _wire_mean(sc2, cfg2, level, 0.05)
            assert np.isfinite(img).all() and (img[..., 2] > 1.0).mean() > 0.001, (name, level)
        sc2.close()


@pytest.mark.parametrize("name", ["weekend", "big_mesh"] + SMALL_SCENES)
def test_path_queue_kernel_equals_tile_kernel(name, gpu_ctx, tmp_path, weekend_scene, monkeypatch):
    """k_render_pq (frame-wide path queue, the default beauty kernel) against k_render (tile-bound warps, ZB_PQ=0): the
    same Philox keys, shading code and 64-bit fixed-point sums, so the frames are equal bit for bit — also under
    tile stripes, sample slices with accumulation, a ragged image edge and a tree big enough for the 7-blocks-per-SM
    instantiation."""
    if name == "weekend":
        sc, cfg = weekend_scene, scenes.weekend(seed=1, width=203, height=90, spp=24)[1]
    elif name == "big_mesh":
        txt, cfg, files = scenes.mesh(width=160, height=90, spp=8, grid=400, knots=2)
        flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files))
        assert flat.contents.n_nodes >= 65536
        sc = gpu_ctx.upload(flat)
        api.free_flat(flat)
    else:
        wh = GOLDEN_SCENES[name][3] or (200, 112)
        sc, cfg, _ = _load_scene(gpu_ctx, tmp_path, name, wh, 24)
    import copy
    full = copy.copy(cfg)  # AOV and light-pass modes of the queue kernel
    full.use_albedo = full.use_normal = full.use_z_depth = full.use_reflection = full.use_refraction = True
    gpu_ctx.set_tuning("pq", 0)
    a = sc.render(cfg, seed=5)["beauty"]
    a_tiles = [sc.render(cfg, seed=5, tile=(r, 3))["beauty"] for r in range(3)]
    a_full = sc.render(full, seed=5)
    gpu_ctx.set_tuning("pq", 1)
    b = sc.render(cfg, seed=5)["beauty"]
    b_tiles = [sc.render(cfg, seed=5, tile=(r, 3))["beauty"] for r in range(3)]
    c = sc.render(cfg, seed=5, samples_per_launch=5)["beauty"]
    d_full = sc.render(full, seed=5)  # default: queue kernel + AOV primary-ray pass; light passes by k_render
    d_full_batched = sc.render(full, seed=5, samples_per_launch=7)
    d_tiles = [sc.render(full, seed=5, tile=(r, 2)) for r in range(2)]
    for key in ("beauty", "albedo", "normal", "z_depth", "reflection", "refraction"):
        assert np.array_equal(a_full[key], d_full[key]), key
        assert np.allclose(d_full[key], d_full_batched[key], rtol=2e-6, atol=1e-7), key
        assert np.array_equal(d_tiles[0][key] + d_tiles[1][key], d_full[key]), key
    gpu_ctx.set_tuning("pq", 2)  # the queue kernel's AOV and light-pass modes (not the default: slower, see abi_render.cu)
    b_full = sc.render(full, seed=5)
    b_full_batched = sc.render(full, seed=5, samples_per_launch=7)
    gpu_ctx.set_tuning("pq", 1)
    for key in ("beauty", "albedo", "normal", "z_depth", "reflection", "refraction"):
        assert np.array_equal(a_full[key], b_full[key]), key
        assert np.allclose(b_full[key], b_full_batched[key], rtol=2e-6, atol=1e-7), key
    if name != "weekend":
        sc.close()
    assert np.array_equal(a, b), np.abs(a - b).max()
    for r in range(3):
        assert np.array_equal(a_tiles[r], b_tiles[r]), r
    assert np.allclose(b, c, rtol=2e-6, atol=1e-7)


# ------------------------------------------------------------------ 8-wide compressed tree (host/wide_bvh.hpp, trace_wide)
@pytest.mark.parametrize("name", ["mesh", "models", "exterior"])
def test_wide_tree_traversal_equals_binary_tree(name, gpu_ctx, tmp_path):
    """Closest hits do not depend on the tree: primary ids, hit distances and rendered frames through the 8-wide compressed
    tree (an opt-in: measured slower, profiles/r02_wide_bvh.txt) equal those through the binary tree (wide = 0), except where
    two primitives are hit at the same t (shared triangle edges: the traversal order decides the tie).  `exterior` has no
    mesh and therefore no wide tree: both settings run the same code."""
    gen, kw, wh, _, ids_kw = GOLDEN_SCENES[name]
    txt, cfg, files = getattr(scenes, gen)(width=wh[0], height=wh[1], spp=8, **{**kw, **ids_kw})
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files), wide=1)
    assert (flat.contents.n_wide_nodes > 0) == (name != "exterior")
    sc = gpu_ctx.upload(flat)
    api.free_flat(flat)
    try:
        gpu_ctx.set_tuning("wide", 0)
        leaf0, prim0, t0 = sc.primary_ids(cfg)
        img0 = sc.render(cfg, seed=3)["beauty"]
        st0 = sc.render(cfg, seed=3, stats=True) and gpu_ctx.stats()
        gpu_ctx.set_tuning("wide", 1)
        leaf1, prim1, t1 = sc.primary_ids(cfg)
        img1 = sc.render(cfg, seed=3)["beauty"]
        st1 = sc.render(cfg, seed=3, stats=True) and gpu_ctx.stats()
    finally:
        gpu_ctx.set_tuning("wide", 1)
        sc.close()
    assert np.array_equal(leaf0, leaf1)
    assert np.array_equal(t0, t1)
    same = prim0 == prim1
    print(f"{name}: face ids equal on {same.mean() * 100:.4f} %; node visits / ray {st0['node_visits'] / st0['rays']:.2f} (binary) "
          f"-> {st1['node_visits'] / st1['rays']:.2f} (8-wide); primitive tests / ray {st0['prim_tests'] / st0['rays']:.2f} -> {st1['prim_tests'] / st1['rays']:.2f}")
    assert same.mean() > 0.999
    assert st0["stack_overflows"] == 0 and st1["stack_overflows"] == 0
    close = np.isclose(img0, img1, rtol=1e-5, atol=1e-6)
    assert close.mean() > 0.995, close.mean()  # an edge tie can pick the neighbouring facet's normal for a path
    assert abs(img0.mean() - img1.mean()) < 2e-3 * img0.mean()
    if name != "exterior":
        assert st1["node_visits"] < 0.8 * st0["node_visits"]


# ------------------------------------------------------------------ transform-only edits
def test_scene_refit_equals_fresh_flatten(gpu_ctx, tmp_path):
    """zb_scene_refit: the scene uploaded at move = 0, handed the instance table of the same graph flattened at move = 1
    (every wrapper's translation / rotation / scale changed, the fog's boundary too), must answer every ray like a fresh
    flatten + upload of the moved scene — same primary ids and distances, same frame — while only the TLAS boxes were
    refitted on the device (topology of move = 0, BLASes untouched); and back again."""
    import time
    flats = {}
    for m in (0.0, 1.0):
        txt, cfg, files = scenes.refit_demo(move=m)
        d = tmp_path / f"m{int(m)}"
        flats[m] = api.flatten_zscene(path=scenes.materialize(str(d), txt, cfg, files))
    f0, f1 = flats[0.0].contents, flats[1.0].contents
    assert f0.n_instances == f1.n_instances >= 7 and f0.n_items == f1.n_items and f0.item_bounds
    sc = gpu_ctx.upload(flats[0.0])
    fresh = {m: gpu_ctx.upload(flats[m]) for m in (0.0, 1.0)}
    ids0 = sc.primary_ids(cfg)
    for m in (1.0, 0.0, 1.0):
        t0 = time.perf_counter()
        sc.refit(flats[m])
        dt = (time.perf_counter() - t0) * 1e3
        leaf, prim, t = sc.primary_ids(cfg)
        leaf_f, prim_f, t_f = fresh[m].primary_ids(cfg)
        print(f"move {m}: refit {dt:.3f} ms (upload of the instance table + k_refit_tlas + sync); hit fraction {(leaf != 0xFFFFFFFF).mean():.3f}")
        assert np.array_equal(leaf, leaf_f) and np.array_equal(prim, prim_f) and np.array_equal(t, t_f), m
        a, b = sc.render(cfg, seed=4)["beauty"], fresh[m].render(cfg, seed=4)["beauty"]
        assert np.array_equal(a, b), (m, np.abs(a - b).max())
        assert dt < 5.0
    assert not np.array_equal(ids0[0], leaf)  # the edit is visible
    assert gpu_ctx.stats()["stack_overflows"] == 0
    # anything but a transform is refused
    other, _, files2 = scenes.quirks()
    (tmp_path / "q").mkdir()
    q = api.flatten_zscene(path=scenes.materialize(str(tmp_path / "q"), other, cfg, files2))
    with pytest.raises(api._lib.ZbError):
        sc.refit(q)
    for s in [sc] + list(fresh.values()):
        s.close()
    for f in list(flats.values()) + [q]:
        api.free_flat(f)


@pytest.mark.parametrize("gpu_blas", [False, True], ids=["host_blas", "device_blas"])
def test_scene_rebuild_tlas_on_device(gpu_ctx, tmp_path, gpu_blas):
    """zb_scene_rebuild_tlas: the scene uploaded at move = 0 gets the instance table of move = 1 and a NEW TLAS built on the
    device over the moved items (sah_build.cu).  Every ray is answered like a fresh flatten + upload of the moved scene
    answers it; a refit of the rebuilt tree back to move = 0 works too (its levels were re-derived); with meshes whose
    BLAS was built on the device the same holds (their root boxes are read back at upload)."""
    flats = {}
    for m in (0.0, 1.0):
        txt, cfg, files = scenes.refit_demo(move=m)
        d = tmp_path / f"m{int(m)}"
        flats[m] = api.flatten_zscene(path=scenes.materialize(str(d), txt, cfg, files), gpu_blas=gpu_blas)
    sc = gpu_ctx.upload(flats[0.0])
    fresh = {m: gpu_ctx.upload(flats[m]) for m in (0.0, 1.0)}

    def same_answers(a, b, what):
        la, pa, ta = a.primary_ids(cfg)
        lb, pb, tb = b.primary_ids(cfg)
        assert (la == lb).mean() >= 0.9999 and np.allclose(ta[la == lb], tb[la == lb], rtol=1e-6, atol=0, equal_nan=True), what
        fa, fb = a.render(cfg, seed=4)["beauty"], b.render(cfg, seed=4)["beauty"]
        differ = (np.abs(fa - fb).max(axis=2) > 0).mean()
        print(f"{what}: ids equal on {(la == lb).mean() * 100:.4f} %, frames differ on {differ * 100:.4f} % of the pixels")
        assert differ < 1e-3, what  # items were re-ordered: exact ties between two items may resolve the other way

    sc.rebuild_tlas(flats[1.0])
    same_answers(sc, fresh[1.0], "rebuilt at move 1")
    sc.refit(flats[0.0])
    same_answers(sc, fresh[0.0], "rebuilt tree refitted back to move 0")
    sc.rebuild_tlas(flats[0.0])
    same_answers(sc, fresh[0.0], "rebuilt at move 0")
    # traversal work of the rebuilt tree against the flattener's tree for the same placement
    work = {}
    for key, s in (("rebuilt", sc), ("flattened", fresh[0.0])):
        s.render(cfg, seed=4, stats=True)
        st = gpu_ctx.stats()
        assert st["stack_overflows"] == 0
        work[key] = (st["node_visits"] + st["prim_tests"]) / st["rays"]
    print(f"TLAS work per ray: device-rebuilt {work['rebuilt']:.2f}, flattener {work['flattened']:.2f}")
    assert work["rebuilt"] <= 1.15 * work["flattened"]
    for s in [sc] + list(fresh.values()):
        s.close()
    for f in flats.values():
        api.free_flat(f)


def test_device_tlas_build_over_a_thousand_items(gpu_ctx):
    """The Weekend scene (485 spheres: phase A of the builder runs, not only the warp phase): TLAS rebuilt on the device
    from the uploaded boxes answers every primary ray like the flattener's TLAS, with no more traversal work."""
    txt, cfg = scenes.weekend(seed=1, width=320, height=180, spp=4)
    flat = api.flatten_zscene(txt)
    a, b = gpu_ctx.upload(flat), gpu_ctx.upload(flat)
    b.rebuild_tlas(flat)
    la, pa, ta = a.primary_ids(cfg)
    lb, pb, tb = b.primary_ids(cfg)
    assert np.array_equal(la, lb) and np.array_equal(ta, tb)
    work = []
    for s in (a, b):
        s.render(cfg, seed=2, stats=True)
        st = gpu_ctx.stats()
        assert st["stack_overflows"] == 0
        work.append((st["node_visits"] + st["prim_tests"]) / st["rays"])
    print(f"Weekend TLAS work per ray: flattener {work[0]:.2f}, device-rebuilt {work[1]:.2f}")
    assert work[1] <= 1.1 * work[0]
    fa, fb = a.render(cfg, seed=2)["beauty"], b.render(cfg, seed=2)["beauty"]
    assert (np.abs(fa - fb).max(axis=2) > 0).mean() < 1e-3
    a.close(); b.close()
    api.free_flat(flat)


def test_fused_finishing_kernel_equals_three_kernel_chain(gpu_ctx):
    """k_finish_fused (bloom + sharpening + process() in one pass over shared-memory tiles) against k_bloom_h + k_bloom_v +
    k_finish on a frame whose size is not a multiple of the tile (ragged edges, image borders inside the halo), for several
    radii and parameter sets: the same arithmetic in the same order, so the float4 results agree to rounding (5e-6) and the
    RGBA8 results to one step on a vanishing fraction of the channels."""
    import torch
    W, H = 203, 117
    g = torch.Generator(device="cuda").manual_seed(7)
    src = (torch.rand(W * H, 4, device="cuda", generator=g) ** 3) * 6.0
    for kw in (dict(use_bloom=1), dict(use_bloom=1, bloom_radius=1, use_sharpening=1), dict(use_bloom=1, bloom_radius=6, use_sharpening=1, use_aces=1),
               dict(use_bloom=1, bloom_radius=3, bloom_threshold=0.5, bloom_intensity=0.8, use_sharpening=1, saturation=1.3, hue_shift=25.0, contrast=1.1)):
        outs = {}
        for fuse in (1, 0):
            gpu_ctx.set_tuning("fuse_finish", fuse)
            f4 = torch.zeros(W * H, 4, device="cuda")
            u8 = torch.zeros(W * H, 4, device="cuda", dtype=torch.uint8)
            gpu_ctx.finish_device(src.data_ptr(), W, H, api.post_struct(**kw), 0, f4.data_ptr(), u8.data_ptr())
            torch.cuda.synchronize()
            outs[fuse] = (f4.cpu().numpy(), u8.cpu().numpy().astype(np.int32))
        gpu_ctx.set_tuning("fuse_finish", 1)
        df = np.abs(outs[1][0] - outs[0][0]).max()
        du = np.abs(outs[1][1] - outs[0][1])
        print(kw, "max |float4 diff|", df, "RGBA8 steps differing", float((du != 0).mean()))
        assert df < 5e-6 and du.max() <= 1 and (du != 0).mean() < 1e-3, (kw, df, du.max())
        assert outs[1][0][:, :3].mean() > 0.05
