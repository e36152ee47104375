"""CPU-side checks: the C-ABI library loads, exports every declared symbol, the host scene API
flattens scenes correctly, malformed flat scenes are rejected, and compute fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from raytracer_project_b200 import _lib, api, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_all_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "zenith_b200.h")).read()
    declared = set(re.findall(r"^(?:int|double|size_t|const char\*)\s+(zb_\w+)\s*\(", hdr, flags=re.M))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.zb_abi_version() == 4


def test_struct_sizes_match_header():
    # the PODs of include/zenith_b200.h as laid out by the C compiler
    assert C.sizeof(_lib.Camera) == 8 + 8 + 72 + 16
    assert C.sizeof(_lib.Environment) == 8 + 24 + 8 + 24 + 48 + 16
    assert C.sizeof(_lib.Stats) == 6 * 8 + 4 * 8 + 16
    assert C.sizeof(_lib.RenderOpts) == 88 and C.sizeof(_lib.GroupHandle) == 256


def _views(f):
    nodes = np.ctypeslib.as_array(C.cast(f.nodes, C.POINTER(C.c_float)), (f.n_nodes, 16))
    items = np.ctypeslib.as_array(C.cast(f.items, C.POINTER(C.c_float)), (f.n_items, 8))
    return nodes, items


def test_weekend_flatten_counts_and_bvh_invariants():
    txt, cfg = scenes.weekend(seed=1)
    flat = api.flatten_zscene(txt)
    f = flat.contents
    assert f.n_items == 484 and f.n_leaves == 484 and f.n_materials == 484 and f.n_instances == 0
    nodes, items = _views(f)
    child = nodes[:, 12:14].copy().view(np.int32)
    meta = items[:, 4:8].copy().view(np.uint32)
    # every leaf id appears exactly once; every item is referenced by exactly one leaf
    assert sorted(meta[:, 2].tolist()) == list(range(484))
    seen = np.zeros(f.n_items, int)
    for node in range(f.n_nodes):
        boxes = {0: (nodes[node, [0, 2, 8]], nodes[node, [1, 3, 9]]), 1: (nodes[node, [4, 6, 10]], nodes[node, [5, 7, 11]])}
        for side in (0, 1):
            c = int(child[node, side])
            if c == -2**31:
                continue
            lo, hi = boxes[side]
            if c >= 0:
                # child's own two boxes are inside the box stored for it in the parent
                clo = np.minimum(nodes[c, [0, 2, 8]], nodes[c, [4, 6, 10]])
                chi = np.maximum(nodes[c, [1, 3, 9]], nodes[c, [5, 7, 11]])
                assert np.all(clo >= lo) and np.all(chi <= hi)
            else:
                code = (~c) & 0xFFFFFFFF
                first, count = code >> 3, (code & 7) + 1
                for i in range(first, first + count):
                    seen[i] += 1
                    cx, cy, cz, r = items[i, :4]
                    assert np.all(np.array([cx - r, cy - r, cz - r]) >= lo - 1e-3)
                    assert np.all(np.array([cx + r, cy + r, cz + r]) <= hi + 1e-3)
    assert np.all(seen == 1)
    api.free_flat(flat)


def test_tlas_region_has_room_for_a_device_rebuild():
    """The flattener pads the TLAS node region to n_items - 1 nodes (zb_scene_rebuild_tlas builds any tree over the items in
    place); the spare nodes are unreachable and empty, and mesh BLAS roots lie behind the region."""
    txt, cfg = scenes.weekend(seed=1)
    flat = api.flatten_zscene(txt)
    f = flat.contents
    nodes, _ = _views(f)
    child = nodes[:, 12:14].copy().view(np.int32)
    assert f.n_tlas_nodes <= f.n_items - 1 <= f.n_nodes
    assert np.all(child[f.n_tlas_nodes:f.n_items - 1] == -2**31)
    reachable = {0}
    todo = [0]
    while todo:
        for c in child[todo.pop()]:
            if c >= 0 and int(c) not in reachable:
                reachable.add(int(c))
                todo.append(int(c))
    assert reachable == set(range(f.n_tlas_nodes))
    api.free_flat(flat)
    txt, cfg, files = scenes.refit_demo()
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        flat = api.flatten_zscene(path=scenes.materialize(d, txt, cfg, files))
    f = flat.contents
    roots = [f.meshes_array()[m].node_root for m in range(f.n_meshes)]
    assert roots and min(roots) >= max(f.n_items - 1, f.n_tlas_nodes)
    api.free_flat(flat)


def test_flatten_wrappers_and_flags():
    w = scenes.SceneWriter()
    glass = w.dielectric(1.5)
    red = w.lambertian((0.8, 0.1, 0.1))
    s = w.sphere((0, 0, 0), 1.0, glass)
    objs = [
        w.translate(s, (1, 2, 3)),                       # baked sphere + force-front flag (SURVEY A.q1)
        w.scale(s, (1, 2, 1)),                           # instance, keeps front_face (A.q2)
        w.translate(w.matinst(w.rotate("y", w.scale(s, (2, 2, 2)), 0.5), red), (5, 0, 0)),
        w.cube((-1, -1, -1), (1, 1, 1), red),
        w.matinst(s, None),                              # null material -> magenta lambertian
    ]
    w.world(w.container("list", objs))
    flat = api.flatten_zscene(w.text())
    f = flat.contents
    assert (f.n_items, f.n_instances, f.n_cubes, f.n_spheres_obj) == (5, 2, 1, 1)   # master sphere shared
    _, items = _views(f)
    meta = items[:, 4:8].copy().view(np.uint32)
    by_leaf = {int(m[2]): (items[i, :4], m) for i, m in enumerate(meta)}
    g0, m0 = by_leaf[0]
    assert m0[0] == (0 | 0x100) and np.allclose(g0, [1, 2, 3, 1])
    assert by_leaf[1][1][0] & 0xFF == 3 and by_leaf[2][1][0] & 0xFF == 3 and by_leaf[3][1][0] & 0xFF == 1
    inst = np.ctypeslib.as_array(C.cast(f.instances, C.POINTER(C.c_float)), (f.n_instances, 32))
    flags = inst[:, 24:28].copy().view(np.uint32)[:, 2]
    i1, i2 = int(by_leaf[1][1][3]), int(by_leaf[2][1][3])
    assert flags[i1] == 0 and flags[i2] == 0x100
    # o2w of leaf 2 = T(5,0,0) * Ry(0.5) * S(2): column 3 is the translation, w2o is its inverse
    o2w = np.vstack([inst[i2, 12:24].reshape(3, 4), [0, 0, 0, 1]])
    w2o = np.vstack([inst[i2, 0:12].reshape(3, 4), [0, 0, 0, 1]])
    assert np.allclose(o2w[:3, 3], [5, 0, 0]) and np.allclose(o2w @ w2o, np.eye(4), atol=1e-6)
    c, s_ = np.cos(0.5), np.sin(0.5)
    assert np.allclose(o2w[:3, :3], 2 * np.array([[c, 0, -s_], [0, 1, 0], [s_, 0, c]]), atol=1e-6)  # rotate_y.hpp:56-62
    mats = np.ctypeslib.as_array(C.cast(f.materials, C.POINTER(C.c_float)), (f.n_materials, 8))
    magenta = mats[int(by_leaf[4][1][1])]
    assert np.allclose(magenta[4:7], [1, 0, 1])
    api.free_flat(flat)


def test_scene_feature_bits(tmp_path):
    """zb_flat_scene_features: the ZB_F_* bits that select the render-kernel instantiation (smallest compiled superset).
    Weekend (BASELINE C1 / headline) must be PLAIN = 0: inline spheres, untextured materials."""
    F = api.FEATURES
    cases = {
        "weekend": (scenes.weekend(seed=1, width=64, height=36, spp=1), 0),
        "cornell": (scenes.cornell(width=64, height=36, spp=1), F["INST"] | F["CUBE"] | F["TRI"] | F["MEDIA"]),
        "mesh": (scenes.mesh(width=64, height=36, spp=1, grid=24, knots=2), F["INST"] | F["MESH"] | F["TRI"]),
    }
    for name, (out, want) in cases.items():
        txt, cfg = out[0], out[1]
        if len(out) > 2 and out[2]:
            flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path / name), txt, cfg, out[2]))
        else:
            flat = api.flatten_zscene(txt)
        got = api.flat_features(flat)
        api.free_flat(flat)
        assert got == want, (name, got, want)
    txt, cfg, files = scenes.exterior(width=64, height=36, spp=1, small=20)
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path / "exterior"), txt, cfg, files))
    got = api.flat_features(flat)
    api.free_flat(flat)
    assert got & F["TEX"] and not got & (F["MESH"] | F["ENV"]), got  # ENV is per render, never a scene bit


def test_zscene_errors_are_reported():
    with pytest.raises(_lib.ZbError) as e:
        api.flatten_zscene("obj a sphere 0 0 0 1 nomat\nworld a\n")
    assert "unknown material" in str(e.value)
    with pytest.raises(_lib.ZbError):
        api.flatten_zscene("obj a sphere 0 0 0 1\n")  # no world


def test_empty_world_flattens():
    flat = api.flatten_zscene("obj w list\nworld w\n")
    assert flat.contents.n_items == 0 and flat.contents.n_nodes == 0
    api.free_flat(flat)


def test_upload_rejects_malformed_scene_before_touching_cuda(lib):
    txt, _ = scenes.weekend(seed=1)
    flat = api.flatten_zscene(txt)
    bad = _lib.FlatScene.from_buffer_copy(flat.contents)
    bad.n_materials = 3  # items now reference materials out of range
    out = C.c_void_p()
    fake_ctx = C.c_void_p(1)
    rc = lib.zb_scene_upload(fake_ctx, C.byref(bad), C.byref(out))
    assert rc == -1 and b"material out of range" in lib.zb_last_error()
    api.free_flat(flat)


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = lib.zb_context_create(0, C.byref(h))
    assert rc == -2 and b"no CPU fallback" in lib.zb_last_error()


def test_drop_in_camera_adapter_compiles_and_links(tmp_path):
    """zenith::camera::render (host/zenith/camera.hpp) with reference-style scene code: C++20, header-only over
    the C ABI; must compile and link against the in-tree library without a GPU."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "drop_in_render")
    subprocess.run(["g++", "-std=c++20", "-O1", "-Wall", "-Wno-unused-function", os.path.join(root, "tests", "cpp", "drop_in_render.cpp"),
                    "-o", exe, "-L" + os.path.join(root, "raytracer_project_b200"), "-lzenith_b200",
                    "-Wl,-rpath," + os.path.join(root, "raytracer_project_b200")], check=True)
    assert subprocess.run([exe]).returncode == 2  # usage error: no compute without arguments


def test_flat_scene_file_round_trip(tmp_path):
    """ZBFS container: save -> load reproduces every array byte for byte; garbage and truncated files
    are rejected with an error, not a crash."""
    # the exterior scene fills every array but media: meshes-free items, instances, cubes, u8 + f32 texel pools
    txt, cfg, files = scenes.exterior(seed=4, width=64, height=36, spp=1, small=20, hdr_size=64)
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files))
    p = str(tmp_path / "scene.zbfs")
    api.save_flat(flat, p)
    back = api.load_flat(p)
    a, b = flat.contents, back.contents
    for name, _ in a._fields_:
        if not isinstance(getattr(a, name), int):
            continue
        if name.startswith("n_") or name in ("abi_version", "pad"):
            assert getattr(a, name) == getattr(b, name), name
    sizes = {"nodes": a.n_nodes * 64, "items": a.n_items * 32, "cubes": a.n_cubes * 32, "tris": a.n_tris * 48,
             "tri_normals": a.n_tris * 48, "instances": a.n_instances * 128, "meshes": a.n_meshes * 16, "media": a.n_media * 32,
             "materials": a.n_materials * 32, "textures": a.n_textures * 32, "spheres_obj": a.n_spheres_obj * 16,
             "texels_u8": a.n_texels_u8 * 4, "texels_f32": a.n_texels_f32 * 16}
    assert sizes["texels_u8"] and sizes["texels_f32"] and sizes["instances"] and sizes["cubes"]
    for name, nbytes in sizes.items():
        if nbytes:
            assert C.string_at(getattr(a, name), nbytes) == C.string_at(getattr(b, name), nbytes), name
    api.free_flat(back)
    api.free_flat(flat)
    raw = open(p, "rb").read()
    open(p, "wb").write(raw[: len(raw) // 2])
    with pytest.raises(_lib.ZbError):
        api.load_flat(p)
    open(p, "wb").write(b"not a scene at all" * 10)
    with pytest.raises(_lib.ZbError):
        api.load_flat(p)


def test_png_and_hdr_writers_round_trip(tmp_path):
    """zb_write_png / zb_write_hdr (host-side encoders of the export): the PNG decodes to the same
    bytes, the RGBE file to the same radiance within the format's 8-bit mantissa."""
    rng = np.random.default_rng(3)
    rgb = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    rgb[5:9, :, :] = 255
    rgb[20:, 10:30, :] = 0
    p = str(tmp_path / "a.png")
    api.write_png(p, rgb)
    back = api.read_png(p)
    assert back.shape == rgb.shape and np.array_equal(back, rgb)
    rgba = np.concatenate([rgb, rng.integers(0, 256, (37, 53, 1), dtype=np.uint8)], -1)
    api.write_png(p, rgba)
    assert np.array_equal(api.read_png(p), rgb), "alpha must be dropped"
    hdr = (rng.lognormal(0.0, 2.5, (19, 31, 3))).astype(np.float32)
    hdr[3, 4] = 0.0
    hdr[5, 6] = [5.0e4, 1.0, 1e-3]
    q = str(tmp_path / "a.hdr")
    api.write_hdr(q, hdr)
    got = api.read_hdr(q)
    m = hdr.max(-1, keepdims=True)
    assert np.all(np.abs(got - hdr) <= m / 128.0 + 1e-30), np.abs(got - hdr).max()
    assert np.all(got[3, 4] == 0)
    assert [api.pass_file_name(i) for i in (0, 2, 6)] == ["render_RGB.png", "render_Albedo.png", "render_Z-Depth.png"]
    with pytest.raises(_lib.ZbError):
        api.write_png(str(tmp_path / "missing_dir" / "x.png"), rgb)


def test_image_texture_decodes_shipped_jpeg_png_like_the_reference(tmp_path):
    """image_texture("x.jpg") with no side-car file: the host decodes the reference's shipped textures (baseline and
    progressive JPEG, 4:2:0 and 4:4:4, PNG) to exactly the texels the reference's stbi_load produced
    (tests/golden/shipped_texels.json, made by oracle/_ref/ref_harness decode; texture.hpp:31)."""
    import hashlib
    import json
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "shipped_texels.json")))
    for rel, g in gold.items():
        src = os.path.join(scenes.ASSET_DIR, rel)
        txt = f"tex t image {src} 0\nmat m lambertian tex t\nobj s sphere 0 0 0 1 m\nobj w bvh s\nworld w\n"
        flat = api.flatten_zscene(txt)
        f = flat.contents
        tex = np.ctypeslib.as_array(C.cast(f.textures, C.POINTER(C.c_uint32)), (f.n_textures, 8))
        img = [t for t in tex if t[0] == 2]
        assert len(img) == 1, (rel, tex[:, 0])
        w, h, first = int(img[0][1]), int(img[0][2]), int(img[0][3])
        assert (w, h) == (g["width"], g["height"]), rel
        px = np.ctypeslib.as_array(C.cast(f.texels_u8, C.POINTER(C.c_uint8)), (f.n_texels_u8, 4))[first:first + w * h, :3].copy()
        api.free_flat(flat)
        assert hashlib.sha256(px.tobytes()).hexdigest() == g["sha256"], rel


def _tree_depth(nodes_i32, root, n_nodes):
    depth, todo = 0, [(root, 1)]
    while todo:
        n, d = todo.pop()
        depth = max(depth, d)
        assert d <= 200, "cycle"
        for c in nodes_i32[n, 12:14]:
            if c >= 0:
                assert c < n_nodes
                todo.append((int(c), d + 1))
    return depth


def test_builders_bound_tree_depth_on_degenerate_scenes(tmp_path):
    """The traversal stack holds one deferred child per level (csrc/trace.cuh): the host builder must keep the TLAS within
    ZB_MAX_TLAS_DEPTH and every BLAS within ZB_MAX_BLAS_DEPTH whatever SAH would like — geometrically spaced objects (SAH
    peels one object per level: a 3000-deep comb without the cap) and a mesh whose triangles share one centroid."""
    w = scenes.SceneWriter()
    m = w.lambertian((0.5, 0.5, 0.5))
    objs = [w.sphere((1.02 ** k, 0.0, 0.0), 0.004 * 1.02 ** k, m) for k in range(3000)]
    # 4096 coincident-centroid triangles (a fan of thin slivers around the origin) + geometrically growing ones
    obj = ["v 0 0 0"]
    import math
    for k in range(4097):
        a = 2 * math.pi * k / 4097
        obj.append("v %.6f %.6f 0" % (math.cos(a), math.sin(a)))
        obj.append("v %.6f %.6f 0" % (-math.cos(a), -math.sin(a)))
    for k in range(4096):
        obj.append(f"f {2 * k + 2} {2 * k + 4} {2 * k + 3}")  # (c, c', -c): centroid ~ 0 for all
    for k in range(600):
        s = 1.05 ** k
        obj.append("v %.6f 0 5\nv %.6f %.6f 5\nv %.6f 0 5" % (s, s, 0.01 * s, 1.02 * s))
        n = 1 + 2 * 4097 + 3 * k
        obj.append(f"f {n + 1} {n + 2} {n + 3}")
    (tmp_path / "fan.obj").write_text("\n".join(obj) + "\n")
    objs.append(w.model("fan.obj", m, 1.0))
    w.world(w.container("bvh", objs, name="world"))
    (tmp_path / "s.zscene").write_text(w.text())
    flat = api.flatten_zscene(path=str(tmp_path / "s.zscene"))
    f = flat.contents
    nodes = np.ctypeslib.as_array(C.cast(f.nodes, C.POINTER(C.c_int32)), (f.n_nodes, 16))
    tlas = _tree_depth(nodes, 0, f.n_nodes)
    meshes = np.ctypeslib.as_array(C.cast(f.meshes, C.POINTER(C.c_uint32)), (f.n_meshes, 4))
    blas = max(_tree_depth(nodes, int(r[0]), f.n_nodes) for r in meshes)
    print("TLAS depth", tlas, "BLAS depth", blas, "items", f.n_items, "tris", f.n_tris)
    api.free_flat(flat)
    assert f.n_items >= 3001 and tlas <= 22 and blas <= 38, (tlas, blas)  # ZB_MAX_TLAS_DEPTH / ZB_MAX_BLAS_DEPTH


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference sources (build container only)")
def test_reference_load_materials_compiles_against_host_api(tmp_path):
    """the reference's OWN material-library code — load_materials(), scene_management.hpp:49-100, 36 add()
    calls with image textures, bump maps, checker textures, color::random — is the compile contract of the host scene API.
    Its text is read from /root/reference at test time (nothing is copied into the repository), compiled under
    `using namespace zenith;` and run where the shipped assets are: every material must exist and the JPEG textures decode."""
    import subprocess
    src = open("/root/reference/scene_management.hpp").read().splitlines()
    start = next(i for i, l in enumerate(src) if l.startswith("void load_materials("))
    end = next(i for i in range(start, len(src)) if src[i].startswith("}"))
    body = "\n".join(src[start:end + 1])
    assert body.count("mat_lib.add(") >= 36
    host = os.path.join(ROOT, "raytracer_project_b200", "host")
    (tmp_path / "t.cpp").write_text(f'''
#include "zenith/scene.hpp"
#include "zenith/materials.hpp"
#include <cstdio>
using namespace zenith;
using std::make_shared;
{body}
int main() {{
    MaterialLibrary lib;
    load_materials(lib);
    auto wood = std::dynamic_pointer_cast<lambertian>(lib.get("wood_bumpy_texture"));
    auto img = wood ? std::dynamic_pointer_cast<image_texture>(wood->tex) : nullptr;
    std::printf("%zu %d %d %zu\\n", lib.get_material_names().size(), img ? img->width : -1, img ? img->height : -1,
                lib.get_emissive_names().size());
    return 0;
}}
''')
    r = subprocess.run(["g++", "-std=c++20", "-O1", "-I", host, "-I", os.path.join(ROOT, "include"), "t.cpp", "-o", "t"], cwd=tmp_path,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    os.symlink(os.path.join(ROOT, "tests", "golden", "assets"), tmp_path / "assets")
    out = subprocess.run(["./t"], cwd=tmp_path, capture_output=True, text=True, check=True).stdout.split()
    n, w, h, emissive = (int(x) for x in out)
    assert n >= 33 and (w, h) == (280, 280) and emissive >= 7, out


WIDE_NODE = np.dtype([("p", "f4", 3), ("e", "u1", 3), ("imask", "u1"), ("child_base", "u4"), ("prim_base", "u4"), ("meta", "u1", 8),
                      ("qlo", "u1", (3, 8)), ("qhi", "u1", (3, 8))])


def test_wide_tree_covers_every_primitive_conservatively(tmp_path):
    """The 8-wide compressed tree of a mesh scene (host/wide_bvh.hpp): 80-byte nodes; every TLAS item and every triangle is
    referenced by exactly one leaf child; a child's dequantised box contains the triangles (or child boxes) below it; inner
    children follow their parent and sit in slot order; the binary tree's leaves address the same (re-ordered) primitives."""
    assert WIDE_NODE.itemsize == 80
    txt, cfg, files = scenes.mesh(width=64, height=36, spp=1, grid=40, knots=3, shipped=True)
    flat = api.flatten_zscene(path=scenes.materialize(str(tmp_path), txt, cfg, files), wide=1)
    f = flat.contents
    assert f.n_wide_nodes > 0 and 0 < f.wide_stack_need <= 31
    wide = np.frombuffer(C.string_at(f.wide_nodes, f.n_wide_nodes * 80), WIDE_NODE)
    tris = np.ctypeslib.as_array(C.cast(f.tris, C.POINTER(C.c_float)), (f.n_tris, 3, 4))[:, :, :3].astype(np.float64)
    inst = np.ctypeslib.as_array(C.cast(f.instances, C.POINTER(C.c_uint32)), (f.n_instances, 32))
    seen_items, seen_tris = np.zeros(f.n_items, int), np.zeros(f.n_tris, int)

    def walk(ni, in_blas, depth=1):
        """returns the TRUE bounds of the triangles below node ni (a child's own quantised boxes may stick out of the box its
        parent holds for it — both contain the geometry, which is what closest hits need)"""
        n = wide[ni]
        scale = np.ldexp(1.0, n["e"].astype(int) - 127)
        lo_all, hi_all = [np.full(3, np.inf)], [np.full(3, -np.inf)]
        inner = 0
        for s in range(8):
            cnt, idx = int(n["meta"][s]) >> 5, int(n["meta"][s]) & 31
            if cnt == 0:
                continue
            lo = n["p"].astype(np.float64) + n["qlo"][:, s] * scale
            hi = n["p"].astype(np.float64) + n["qhi"][:, s] * scale
            if idx >= 24:
                assert idx == 24 + s and (int(n["imask"]) >> s) & 1
                child = int(n["child_base"]) + bin(int(n["imask"]) & ((1 << s) - 1)).count("1")
                assert child > ni
                clo, chi = walk(child, in_blas, depth + 1)
                if in_blas:
                    assert np.all(lo <= clo + 1e-9) and np.all(hi >= chi - 1e-9), (ni, s)
                inner += 1
            else:
                clo, chi = np.full(3, np.inf), np.full(3, -np.inf)
                for j in range(cnt):
                    prim = int(n["prim_base"]) + idx + j
                    if in_blas:
                        seen_tris[prim] += 1
                        clo, chi = np.minimum(clo, tris[prim].min(0)), np.maximum(chi, tris[prim].max(0))
                    else:
                        seen_items[prim] += 1
                if in_blas:
                    assert np.all(lo <= clo + 1e-9) and np.all(hi >= chi - 1e-9), (ni, s)
            lo_all.append(clo)
            hi_all.append(chi)
        assert inner == bin(int(n["imask"])).count("1")
        return np.min(lo_all, 0), np.max(hi_all, 0)

    walk(0, False)
    for root in sorted({int(r[28]) for r in inst if r[24] == 2}):  # zb_instance: geom_kind == MESH -> pad[1]
        walk(root, True)
    assert np.all(seen_items == 1) and np.all(seen_tris == 1)
    # the binary tree addresses the same primitive order: its leaves still cover every triangle once
    nodes = np.ctypeslib.as_array(C.cast(f.nodes, C.POINTER(C.c_int32)), (f.n_nodes, 16))
    cover = np.zeros(f.n_tris, int)
    for n in range(f.n_tlas_nodes, f.n_nodes):
        for c in nodes[n, 12:14]:
            if c < 0 and c != -2**31:
                code = (~int(c)) & 0xFFFFFFFF
                cover[(code >> 3):(code >> 3) + (code & 7) + 1] += 1
    assert np.all(cover == 1)
    api.free_flat(flat)
