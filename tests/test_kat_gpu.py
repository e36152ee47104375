"""Per-function known answers: single device functions of the shading code (zb_kat_eval) against the values
the REFERENCE's own functions returned for the same float inputs (tests/golden/kat.json, made by oracle/ref_kat.inc through
oracle/_ref/ref_harness): refract, reflect, Schlick reflectance, texture::value (image wrap / clamp on the shipped JPEGs,
checker parity at negative coordinates, solids), get_bumped_normal, get_sphere_uv and get_background_color in its three
modes.  Tolerance 1e-5 (fp32 device arithmetic against the reference's doubles); texel look-ups are exact 8-bit values."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from raytracer_project_b200 import _lib, api, scenes
from tests.golden.make_golden import scene_digest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
OPS = {"refract": 0, "reflect": 1, "reflectance": 2, "texture": 3, "bump": 4, "sphere_uv": 5, "background": 6}
N_OUT = {"refract": 3, "reflect": 3, "reflectance": 1, "texture": 3, "bump": 3, "sphere_uv": 2, "background": 3}
KAT_CASE = np.dtype([("op", "u4"), ("index", "u4"), ("a", "f4", 14)])
IMAGE_WIDTH = {"t_wood": 280, "t_woodbump": 1024, "t_scratches": 600, "t_concrete": 626}


def _indices(flat):
    """zscene names of the KAT scene -> textures[] / materials[] indices of the flattened scene."""
    f = flat.contents
    tex = np.ctypeslib.as_array(C.cast(f.textures, C.POINTER(C.c_uint32)), (f.n_textures, 8)).copy()
    texf = tex.view(np.float32)
    t = {}
    for i, row in enumerate(tex):
        if row[0] == 1:
            t["t_checker"] = i
        elif row[0] == 2:
            for name, w in IMAGE_WIDTH.items():
                if row[1] == w:
                    t[name] = i
        elif row[0] == 0:
            if abs(texf[i, 4] - 0.15) < 1e-6:
                t["t_odd"] = i
            elif abs(texf[i, 4] - 0.9) < 1e-6:
                t["t_even"] = i
    mat = np.ctypeslib.as_array(C.cast(f.materials, C.POINTER(C.c_uint32)), (f.n_materials, 8)).copy()
    m = {}
    for i, row in enumerate(mat):
        if row[2] == t.get("t_woodbump"):
            m["m_wood"] = i
        elif row[2] == t.get("t_scratches"):
            m["m_gold"] = i
        elif row[2] == t.get("t_concrete"):
            m["m_glass"] = i
    return t, m


def test_device_functions_match_reference_known_answers(gpu_ctx, tmp_path):
    gold = json.load(open(os.path.join(GOLDEN, "kat.json")))
    seen = set()
    for env in ("solid", "hdr", "sun"):
        txt, cfg, files = scenes.kat_scene(env)
        assert scene_digest(txt, cfg, files) == gold["digests"][env]
        d = tmp_path / env
        flat = api.flatten_zscene(path=scenes.materialize(str(d), txt, cfg, files))
        tex, mat = _indices(flat)
        assert len(tex) == 7 and len(mat) == 3, (tex, mat)
        sc = gpu_ctx.upload(flat)
        api.free_flat(flat)
        cases = [c for c in gold["cases"] if (c["op"] == "background") == (env != "solid") or (env == "solid")]
        cases = [c for c in cases if c["op"] != "background" or c["name"] == env]
        arr = np.zeros(len(cases), KAT_CASE)
        for k, c in enumerate(cases):
            arr["op"][k] = OPS[c["op"]]
            arr["index"][k] = tex[c["name"]] if c["op"] == "texture" else mat[c["name"]] if c["op"] == "bump" else 0
            arr["a"][k, :len(c["in"])] = c["in"]
        assert arr.dtype.itemsize == 64
        out = np.zeros((len(cases), 4), np.float32)
        e = api.env_struct(cfg)
        api.check(gpu_ctx.lib.zb_kat_eval(gpu_ctx.h, sc.h, C.byref(e), len(cases), arr.ctypes.data, out.ctypes.data))
        sc.close()
        worst = {}
        for k, c in enumerate(cases):
            want = np.array(c["out"])
            got = out[k, :N_OUT[c["op"]]].astype(np.float64)
            err = np.abs(got - want).max() / max(1.0, np.abs(want).max())
            key = (c["op"], c["name"])
            worst[key] = max(worst.get(key, 0.0), err)
            seen.add(key)
            assert np.all(out[k, 3] >= 0)
        for key, err in sorted(worst.items()):
            print(env, key, f"max err {err:.2e}")
            assert err < 1e-5, (env, key, err)
    assert {k[0] for k in seen} == set(OPS)
