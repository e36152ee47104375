"""Seeded scene generators for the BASELINE.json configs, emitted as neutral "zscene" text.

The same text is consumed by (a) the product's C++20 host (scene API -> flatten) and (b) the oracle
harness that instantiates the reference's own classes (oracle/ref_harness.cpp), so both sides see
the same primitive list.  The reference has no scene files and no Weekend /
Cornell scene of its own (scene_management.hpp:103-235 is its only, hard-coded, scene).

Scene randomness comes from random.Random(seed) — never from the reference's unseedable
random_double() (common.hpp:29-34).
"""
from __future__ import annotations

import math
import random
from dataclasses import dataclass, field


def _f(x: float) -> str:
    return repr(float(x))


@dataclass
class RenderConfig:
    """The reference's camera fields (camera.hpp:26-57) + EnvironmentSettings (environment.hpp:8-77)."""

    image_width: int = 400
    image_height: int = 225
    samples_per_pixel: int = 10
    max_depth: int = 50
    vfov: float = 30.0
    lookfrom: tuple = (10.0, 1.5, 0.0)
    lookat: tuple = (0.0, 0.0, 0.0)
    vup: tuple = (0.0, 1.0, 0.0)
    defocus_angle: float = 0.0
    focus_dist: float = 10.0
    env_mode: str = "solid"  # solid | hdr | sun
    background: tuple = (0.0, 0.0, 0.0)
    env_intensity: float = 1.0
    hdr_path: str = ""
    hdri_rotation: float = 0.0
    hdri_tilt: float = 0.0
    hdri_roll: float = 0.0
    sun_direction: tuple = (1.0, 0.5, -0.5)
    sun_color: tuple = (1.0, 1.0, 1.0)
    sun_intensity: float = 1.0
    sun_size: float = 1.0
    use_albedo: bool = False
    use_normal: bool = False
    use_z_depth: bool = False
    z_depth_max_dist: float = 20.0
    use_reflection: bool = False  # camera.hpp:56-57: the two light passes
    use_refraction: bool = False
    extra: dict = field(default_factory=dict)

    def to_text(self) -> str:
        v3 = lambda t: " ".join(_f(x) for x in t)
        lines = [
            f"image_width {self.image_width}",
            f"image_height {self.image_height}",
            f"samples_per_pixel {self.samples_per_pixel}",
            f"max_depth {self.max_depth}",
            f"vfov {_f(self.vfov)}",
            f"lookfrom {v3(self.lookfrom)}",
            f"lookat {v3(self.lookat)}",
            f"vup {v3(self.vup)}",
            f"defocus_angle {_f(self.defocus_angle)}",
            f"focus_dist {_f(self.focus_dist)}",
            f"env_mode {self.env_mode}",
            f"background {v3(self.background)}",
            f"env_intensity {_f(self.env_intensity)}",
            f"hdri_rotation {_f(self.hdri_rotation)}",
            f"hdri_tilt {_f(self.hdri_tilt)}",
            f"hdri_roll {_f(self.hdri_roll)}",
            f"sun_direction {v3(self.sun_direction)}",
            f"sun_color {v3(self.sun_color)}",
            f"sun_intensity {_f(self.sun_intensity)}",
            f"sun_size {_f(self.sun_size)}",
            f"use_albedo {int(self.use_albedo)}",
            f"use_normal {int(self.use_normal)}",
            f"use_z_depth {int(self.use_z_depth)}",
            f"z_depth_max_dist {_f(self.z_depth_max_dist)}",
        ]
        if self.hdr_path:
            lines.append(f"hdr_path {self.hdr_path}")
        if self.use_reflection or self.use_refraction:  # written only when set: fixture digests cover this text
            lines += [f"use_reflection {int(self.use_reflection)}", f"use_refraction {int(self.use_refraction)}"]
        return "\n".join(lines) + "\n"


class SceneWriter:
    """Tiny builder for zscene text (format: raytracer_project_b200/host/zscene.hpp)."""

    def __init__(self):
        self.lines: list[str] = []
        self._n = 0

    def _name(self, prefix):
        self._n += 1
        return f"{prefix}{self._n}"

    def tex_solid(self, rgb, name=None):
        name = name or self._name("t")
        self.lines.append(f"tex {name} solid {_f(rgb[0])} {_f(rgb[1])} {_f(rgb[2])}")
        return name

    def tex_checker(self, scale, odd, even, name=None):
        name = name or self._name("t")
        self.lines.append(f"tex {name} checker {_f(scale)} {odd} {even}")
        return name

    def tex_image(self, path, is_hdr=False, name=None):
        name = name or self._name("t")
        self.lines.append(f"tex {name} image {path} {int(is_hdr)}")
        return name

    @staticmethod
    def _albedo(rgb=None, tex=None):
        return f"tex {tex}" if tex else f"color {_f(rgb[0])} {_f(rgb[1])} {_f(rgb[2])}"

    @staticmethod
    def _bump(bump, strength):
        return f" bump {bump} {_f(strength)}" if bump else ""

    def lambertian(self, rgb=None, tex=None, bump=None, strength=1.0, name=None):
        name = name or self._name("m")
        self.lines.append(f"mat {name} lambertian {self._albedo(rgb, tex)}{self._bump(bump, strength)}")
        return name

    def metal(self, rgb=None, fuzz=0.0, tex=None, bump=None, strength=1.0, name=None):
        name = name or self._name("m")
        self.lines.append(f"mat {name} metal {self._albedo(rgb, tex)} {_f(fuzz)}{self._bump(bump, strength)}")
        return name

    def dielectric(self, ior, tint=None, bump=None, strength=1.0, name=None):
        name = name or self._name("m")
        t = f" tint {_f(tint[0])} {_f(tint[1])} {_f(tint[2])}" if tint else ""
        self.lines.append(f"mat {name} dielectric {_f(ior)}{t}{self._bump(bump, strength)}")
        return name

    def light(self, rgb=None, tex=None, name=None):
        name = name or self._name("m")
        self.lines.append(f"mat {name} light {self._albedo(rgb, tex)}")
        return name

    def sphere(self, c, r, mat, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} sphere {_f(c[0])} {_f(c[1])} {_f(c[2])} {_f(r)} {mat or 'null'}")
        return name

    def cube(self, lo, hi, mat, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} cube {' '.join(_f(x) for x in lo)} {' '.join(_f(x) for x in hi)} {mat or 'null'}")
        return name

    def cube1(self, c, mat, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} cube1 {' '.join(_f(x) for x in c)} {mat or 'null'}")
        return name

    def triangle(self, a, b, c, n0, n1, n2, mat, name=None):
        name = name or self._name("o")
        nums = " ".join(_f(x) for v in (a, b, c, n0, n1, n2) for x in v)
        self.lines.append(f"obj {name} triangle {nums} {mat or 'null'}")
        return name

    def translate(self, child, d, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} translate {child} {_f(d[0])} {_f(d[1])} {_f(d[2])}")
        return name

    def scale(self, child, s, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} scale {child} {_f(s[0])} {_f(s[1])} {_f(s[2])}")
        return name

    def rotate(self, axis, child, angle, name=None):
        """axis 'x'/'z': degrees; 'y': radians — the reference's own inconsistency (SURVEY A.q3)."""
        name = name or self._name("o")
        self.lines.append(f"obj {name} rotate_{axis} {child} {_f(angle)}")
        return name

    def matinst(self, child, mat, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} matinst {child} {mat or 'null'}")
        return name

    def model(self, path, mat, scale=1.0, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} model {path} {mat or 'null'} {_f(scale)}")
        return name

    def medium(self, child, density, rgb=None, tex=None, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} medium {child} {_f(density)} {self._albedo(rgb, tex)}")
        return name

    def container(self, kind, children, name=None):
        name = name or self._name("g")
        self.lines.append(f"obj {name} {kind} {' '.join(children)}")
        return name

    def world(self, name):
        self.lines.append(f"world {name}")

    def text(self) -> str:
        return "\n".join(self.lines) + "\n"


def weekend(seed: int = 1, width: int = 1200, height: int = 675, spp: int = 10, max_depth: int = 50):
    """C1 / headline: 'Ray Tracing in One Weekend' final scene through the reference's API
    (SURVEY.md section 8(d) row C1): ground r=1000, 22x22 jittered small spheres (80% lambertian,
    15% metal, 5% glass), three r=1 spheres; plain spheres in one bvh_node root."""
    rng = random.Random(seed)
    w = SceneWriter()
    objs = [w.sphere((0, -1000, 0), 1000, w.lambertian((0.5, 0.5, 0.5)))]
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose = rng.random()
            c = (a + 0.9 * rng.random(), 0.2, b + 0.9 * rng.random())
            if math.dist(c, (4, 0.2, 0)) <= 0.9:
                continue
            if choose < 0.8:
                alb = tuple(rng.random() * rng.random() for _ in range(3))
                m = w.lambertian(alb)
            elif choose < 0.95:
                alb = tuple(0.5 + 0.5 * rng.random() for _ in range(3))
                m = w.metal(alb, 0.5 * rng.random())
            else:
                m = w.dielectric(1.5)
            objs.append(w.sphere(c, 0.2, m))
    objs.append(w.sphere((0, 1, 0), 1.0, w.dielectric(1.5)))
    objs.append(w.sphere((-4, 1, 0), 1.0, w.lambertian((0.4, 0.2, 0.1))))
    objs.append(w.sphere((4, 1, 0), 1.0, w.metal((0.7, 0.6, 0.5), 0.0)))
    w.world(w.container("bvh", objs, name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=20.0,
        lookfrom=(13.0, 2.0, 3.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.6, focus_dist=10.0,
        env_mode="solid", background=(0.7, 0.8, 1.0), env_intensity=1.0)
    return w.text(), cfg


def big(seed: int = 5, width: int = 3840, height: int = 2160, spp: int = 1024, max_depth: int = 50, grid: int = 2237):
    """C5: the C1 spheres + one OBJ heightfield of 2 (grid-1)^2 triangles (grid 2237 ->
    10 000 712) under translate(scale(material_instance(model))), C1 camera pulled back."""
    txt, cfg = weekend(seed=1, width=width, height=height, spp=spp, max_depth=max_depth)
    lines = txt.rstrip("\n").split("\n")
    world = [l for l in lines if l.startswith("obj world ")][0]
    body = [l for l in lines if l is not world and not l.startswith("world ")]
    body += ["mat terrain_mat lambertian color 0.45 0.55 0.35",
             "obj terrain_model model heightfield.obj null 1.0",
             "obj terrain_mi matinst terrain_model terrain_mat",
             "obj terrain_s scale terrain_mi 2.0 1.0 2.0",
             "obj terrain translate terrain_s 0.0 -0.25 0.0"]
    body.append(world + " terrain")
    body.append("world world")
    cfg.lookfrom = (19.5, 3.0, 4.5)
    cfg.focus_dist = 15.0
    return "\n".join(body) + "\n", cfg, {"heightfield.obj": _heightfield_obj(grid, seed)}


def materialize(directory: str, text: str, cfg: RenderConfig, files: dict | None = None):
    """Write `s.zscene`, `s.cfg` and the scene's asset files into `directory`; returns the zscene path."""
    import os
    os.makedirs(directory, exist_ok=True)
    for rel, data in (files or {}).items():
        path = os.path.join(directory, rel)
        os.makedirs(os.path.dirname(path) or directory, exist_ok=True)
        with open(path, "wb" if isinstance(data, (bytes, bytearray)) else "w") as f:
            f.write(data)
    with open(os.path.join(directory, "s.zscene"), "w") as f:
        f.write(text)
    with open(os.path.join(directory, "s.cfg"), "w") as f:
        f.write(cfg.to_text())
    return os.path.join(directory, "s.zscene")


ASSET_DIR = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))),
                                     "tests", "golden", "assets")


def shipped_asset(rel: str) -> bytes:
    """Bytes of one of the reference's shipped assets (assets/models/*.obj, assets/bump_maps/*.jpg,
    assets/textures/*), committed as test vectors under tests/golden/assets/."""
    import os
    with open(os.path.join(ASSET_DIR, rel), "rb") as f:
        return f.read()


def _quad(w: SceneWriter, p0, p1, p2, p3, n, mat):
    """A quad as the two triangles the reference needs (it has no quad primitive)."""
    return [w.triangle(p0, p1, p2, n, n, n, mat), w.triangle(p0, p2, p3, n, n, n, mat)]


def cornell(seed: int = 2, width: int = 1920, height: int = 1080, spp: int = 256, max_depth: int = 50, fog: bool = True):
    """C2: 555^3 Cornell room from triangle pairs, two translate(rotate_y(cube)) boxes
    (rotate_y takes RADIANS, Appendix A q3), a 130x105 diffuse_light(15) patch, and a sphere-bounded
    constant_medium (density 0.002) kept BESIDE the bvh_node in a list root (Appendix A q5, q6)."""
    w = SceneWriter()
    red, white, green = w.lambertian((0.65, 0.05, 0.05)), w.lambertian((0.73, 0.73, 0.73)), w.lambertian((0.12, 0.45, 0.15))
    light = w.light((15.0, 15.0, 15.0))
    objs = []
    objs += _quad(w, (555, 0, 0), (555, 555, 0), (555, 555, 555), (555, 0, 555), (-1, 0, 0), green)
    objs += _quad(w, (0, 0, 0), (0, 0, 555), (0, 555, 555), (0, 555, 0), (1, 0, 0), red)
    objs += _quad(w, (0, 0, 0), (555, 0, 0), (555, 0, 555), (0, 0, 555), (0, 1, 0), white)
    objs += _quad(w, (0, 555, 0), (0, 555, 555), (555, 555, 555), (555, 555, 0), (0, -1, 0), white)
    objs += _quad(w, (0, 0, 555), (555, 0, 555), (555, 555, 555), (0, 555, 555), (0, 0, -1), white)
    objs += _quad(w, (213, 554, 227), (343, 554, 227), (343, 554, 332), (213, 554, 332), (0, -1, 0), light)
    box1 = w.cube((-82.5, -82.5, -82.5), (82.5, 82.5, 82.5), white)
    objs.append(w.translate(w.rotate("y", box1, math.radians(-18.0)), (212.5, 82.5, 147.5)))
    box2 = w.cube((-82.5, -165.0, -82.5), (82.5, 165.0, 82.5), white)
    objs.append(w.translate(w.rotate("y", box2, math.radians(15.0)), (347.5, 165.0, 377.5)))
    root = [w.container("bvh", objs, name="room")]
    if fog:
        boundary = w.sphere((278, 278, 278), 400.0, white)
        root.append(w.medium(boundary, 0.002, rgb=(1.0, 1.0, 1.0)))
    w.world(w.container("list", root, name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=40.0,
        lookfrom=(278.0, 278.0, -800.0), lookat=(278.0, 278.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0,
        focus_dist=10.0, env_mode="solid", background=(0.0, 0.0, 0.0), env_intensity=1.0)
    return w.text(), cfg, {}


def _heightfield_obj(n: int, seed: int) -> str:
    """n x n vertices over [-5,5]^2, y = .3 sin(3x) cos(3z) + .05 noise, analytic vertex normals."""
    import numpy as np
    rng = np.random.default_rng(seed)
    xs = np.linspace(-5.0, 5.0, n)
    x, z = np.meshgrid(xs, xs, indexing="ij")
    y = 0.3 * np.sin(3 * x) * np.cos(3 * z) + 0.05 * rng.random((n, n))
    dydx, dydz = 0.9 * np.cos(3 * x) * np.cos(3 * z), -0.9 * np.sin(3 * x) * np.sin(3 * z)
    nrm = np.stack([-dydx, np.ones_like(y), -dydz], axis=-1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    v = np.stack([x, y, z], axis=-1).reshape(-1, 3)
    out = ["# synthetic heightfield"]
    out += ["v %.6f %.6f %.6f" % r for r in map(tuple, v.tolist())]
    out += ["vn %.6f %.6f %.6f" % r for r in map(tuple, nrm.reshape(-1, 3).tolist())]
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a = (i * n + j + 1).ravel()  # 1-based
    b, c, d = a + n, a + n + 1, a + 1
    f1 = np.stack([a, d, c], axis=1)  # counter-clockwise seen from +y
    f2 = np.stack([a, c, b], axis=1)
    faces = np.concatenate([f1, f2], axis=0)
    out += ["f %d//%d %d//%d %d//%d" % (r[0], r[0], r[1], r[1], r[2], r[2]) for r in faces.tolist()]
    return "\n".join(out) + "\n"


def _torus_obj(nu: int = 48, nv: int = 24, R: float = 1.0, r: float = 0.35) -> str:
    import numpy as np
    u = np.linspace(0, 2 * np.pi, nu, endpoint=False)
    v = np.linspace(0, 2 * np.pi, nv, endpoint=False)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    p = np.stack([(R + r * np.cos(vv)) * np.cos(uu), r * np.sin(vv), (R + r * np.cos(vv)) * np.sin(uu)], -1).reshape(-1, 3)
    nn = np.stack([np.cos(vv) * np.cos(uu), np.sin(vv), np.cos(vv) * np.sin(uu)], -1).reshape(-1, 3)
    out = ["# synthetic torus"]
    out += ["v %.6f %.6f %.6f" % tuple(x) for x in p]
    out += ["vn %.6f %.6f %.6f" % tuple(x) for x in nn]
    for i in range(nu):
        for j in range(nv):
            a, b = i * nv + j + 1, ((i + 1) % nu) * nv + j + 1
            c, d = ((i + 1) % nu) * nv + (j + 1) % nv + 1, i * nv + (j + 1) % nv + 1
            out.append(f"f {a}//{a} {d}//{d} {c}//{c}")
            out.append(f"f {a}//{a} {c}//{c} {b}//{b}")
    return "\n".join(out) + "\n"


def mesh(seed: int = 3, width: int = 1920, height: int = 1080, spp: int = 64, max_depth: int = 50, grid: int = 708,
         knots: int = 12, shipped: bool = False):
    """C3: an OBJ heightfield of 2 (grid-1)^2 triangles (grid 708 -> 999 698) loaded by
    model(path, nullptr, 1.0) under a material_instance, plus `knots` instances of one shared torus OBJ as
    translate(rotate_y(scale(model))) with library-style materials (scene_management.hpp:57-99).  shipped=True instances
    the reference's own assets/models/torus_knot.obj (4 860 triangles) instead of the synthetic torus."""
    rng = random.Random(seed)
    w = SceneWriter()
    files = {"heightfield.obj": _heightfield_obj(grid, seed),
             "torus.obj": shipped_asset("models/torus_knot.obj") if shipped else _torus_obj()}
    ground = w.lambertian((0.45, 0.55, 0.35))
    objs = [w.matinst(w.model("heightfield.obj", None, 1.0), ground)]
    checker = w.tex_checker(0.5, w.tex_solid((0.1, 0.1, 0.1)), w.tex_solid((0.9, 0.9, 0.9)))
    mats = [w.metal((1.0, 0.8, 0.4), 0.0), w.dielectric(1.5), w.metal((0.89, 0.58, 0.51), 0.2),
            w.light((0.0, 2.0, 4.0)), w.metal(tex=checker, fuzz=0.95), w.lambertian((0.8, 0.1, 0.1))]
    torus = w.model("torus.obj", None, 1.0)
    for k in range(knots):
        ang = 2 * math.pi * k / max(knots, 1)
        rad = 2.2 + 1.6 * rng.random()
        s = (0.35 + 0.4 * rng.random()) * (0.45 if shipped else 1.0)  # the shipped knot spans about +-2.3 units
        inst = w.rotate("y", w.scale(torus, (s, s * (0.7 + 0.6 * rng.random()), s)), rng.random() * math.pi)
        inst = w.matinst(inst, mats[k % len(mats)])
        objs.append(w.translate(inst, (rad * math.cos(ang), 0.9 + 0.8 * rng.random(), rad * math.sin(ang))))
    w.world(w.container("bvh", objs, name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=40.0,
        lookfrom=(8.0, 5.0, 8.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0, focus_dist=10.0,
        env_mode="solid", background=(0.7, 0.8, 1.0), env_intensity=1.0)
    return w.text(), cfg, files


def _sky_hdr(width: int = 2048, height: int = 1024, sun_radiance: float = 5.0e4):
    """Synthetic lat-long sky: vertical gradient + a 1-degree-radius sun of radiance 5e4."""
    import numpy as np
    v = (np.arange(height) + 0.5) / height           # 0 = +y (zenith) ... 1 = -y, as texture.hpp:50-78 indexes it
    u = (np.arange(width) + 0.5) / width
    theta, phi = np.pi * v[:, None], 2 * np.pi * u[None, :] - np.pi
    d = np.stack([np.sin(theta) * np.cos(phi), np.cos(theta) * np.ones_like(phi), np.sin(theta) * np.sin(phi)], -1)
    up = np.clip(d[..., 1], 0.0, 1.0)[..., None]
    sky = (1.0 - up) * np.array([0.9, 0.95, 1.0]) + up * np.array([0.25, 0.45, 0.95])
    sky = np.where(d[..., 1:2] < 0.0, np.array([0.12, 0.11, 0.10]), sky)
    sun = np.array([0.5, 0.6, -0.62])
    sun /= np.linalg.norm(sun)
    cosang = (d * sun).sum(-1, keepdims=True)
    sky = np.where(cosang > math.cos(math.radians(1.0)), sun_radiance * np.array([1.0, 0.92, 0.8]), sky)
    return sky.astype(np.float32)


def _noise_tex(size: int, seed: int, octaves: int = 4, channels: int = 1):
    """Smooth value noise in [0, 255] as an (size, size, 3) uint8 image (bump / albedo source)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    img = np.zeros((size, size, channels))
    amp, total = 1.0, 0.0
    for o in range(octaves):
        n = 4 << o
        g = rng.random((n + 1, n + 1, channels))
        g[-1], g[:, -1] = g[0], g[:, 0]  # tile in u and v
        t = np.arange(size) * n / size
        i = t.astype(int)
        f = (t - i)[:, None]
        f = f * f * (3 - 2 * f)
        rows = g[i] * (1 - f[..., None]) + g[i + 1] * f[..., None]           # (size, n+1, c)
        fc = (t - i)[None, :, None]
        fc = fc * fc * (3 - 2 * fc)
        img += amp * (rows[:, i] * (1 - fc) + rows[:, i + 1] * fc)
        total += amp
        amp *= 0.5
    img = np.clip(img / total, 0, 1)
    if channels == 1:
        img = np.repeat(img, 3, axis=2)
    return (img * 255.0 + 0.5).astype(np.uint8)


def exterior(seed: int = 4, width: int = 1920, height: int = 1080, spp: int = 128, max_depth: int = 50, small: int = 200,
             hdr_size: int = 2048, sun_radiance: float = 5.0e4, shipped: bool = False):
    """C4: HDRI-lit exterior with thin-lens DOF (defocus_angle 3.0, focus_dist 9.0):
    r=1000 ground with a reflective checker metal, bump-mapped mirror / gold / wood spheres, a bumpy-glass
    2x2x2 cube, and `small` instanced cubes/spheres built exactly like the reference's demo grid
    translate(material_instance(rotate_y?(scale(master)))) (scene_management.hpp:139-203).  Textures are
    synthetic (the reference's JPGs cannot travel); every image exists as PNG/HDR for the reference's
    stb_image and as .zbtex for the host, with identical texels (ztex.py)."""
    from . import ztex
    rng = random.Random(seed)
    w = SceneWriter()
    files = {}
    if shipped:
        # the reference's own JPEGs under their own relative paths (scene_management.hpp:51-54, :67-68): baseline and
        # progressive, 4:2:0 and 4:4:4 — decoded by stb_image there and by host/jpeg_png.hpp here, no side-car dumps
        names = ("assets/bump_maps/scratches_bump_map.jpg", "assets/bump_maps/wood_bump_map.jpg",
                 "assets/bump_maps/concrete_bump_map.jpg", "assets/textures/fine-wood.jpg")
        for n in names:
            files[n] = shipped_asset(n[len("assets/"):])
    else:
        names = ("scratches.png", "woodbump.png", "concrete.png", "wood.png")
        files.update(ztex.ldr_pair("scratches.png", _noise_tex(512, seed * 10 + 1, octaves=6)))
        files.update(ztex.ldr_pair("woodbump.png", _noise_tex(512, seed * 10 + 2, octaves=3)))
        files.update(ztex.ldr_pair("concrete.png", _noise_tex(256, seed * 10 + 3, octaves=5)))
        files.update(ztex.ldr_pair("wood.png", _noise_tex(512, seed * 10 + 4, octaves=4, channels=3)))
    files.update(ztex.hdr_pair("sky.hdr", _sky_hdr(hdr_size, hdr_size // 2, sun_radiance)))
    scratches, woodbump, concrete = (w.tex_image(n) for n in names[:3])
    wood = w.tex_image(names[3])
    checker = w.tex_checker(0.5, w.tex_solid((0.1, 0.1, 0.1)), w.tex_solid((0.9, 0.9, 0.9)))
    objs = [w.matinst(w.sphere((0, -1000, 0), 1000.0, None), w.metal(tex=checker, fuzz=0.02))]
    objs.append(w.translate(w.matinst(w.sphere((0, 0, 0), 1.0, None), w.metal((1, 1, 1), 0.0, bump=scratches, strength=1.0)), (0, 1, 0)))
    objs.append(w.translate(w.matinst(w.sphere((0, 0, 0), 0.5, None), w.metal((1.0, 0.8, 0.4), 0.0, bump=scratches, strength=-1.0)), (2.0, 0.5, 1.5)))
    objs.append(w.translate(w.matinst(w.sphere((0, 0, 0), 0.5, None), w.lambertian(tex=wood, bump=woodbump, strength=8.0)), (2.0, 0.5, -1.5)))
    objs.append(w.translate(w.matinst(w.cube1((0, 0, 0), None), w.dielectric(1.5, bump=concrete, strength=0.02)), (-3.0, 1.0, 2.5)))
    master_cube = w.cube((-0.2, -0.2, -0.2), (0.2, 0.2, 0.2), None)
    master_sphere = w.sphere((0, 0, 0), 0.2, None)
    neon = [w.light((3.0, 0.0, 1.5)), w.light((0.0, 2.0, 4.0)), w.light((0.4, 4.0, 0.4)), w.light((6.0, 4.8, 0.0))]
    regular = [w.lambertian((0.8, 0.1, 0.1)), w.metal((0.95, 0.64, 0.54), 0.0), w.metal((1.0, 0.84, 0.0), 0.15),
               w.lambertian((0.1, 0.4, 0.9)), w.metal((1, 1, 1), 0.25), w.lambertian(tex=checker),
               w.metal((0.9, 0.9, 0.9), 0.6, bump=concrete, strength=2.0), w.lambertian(tex=wood)]
    glass, bubble = w.dielectric(1.5), w.dielectric(1.0 / 1.5)
    side = max(1, int(math.ceil(math.sqrt(small))))
    placed = 0
    for a in range(-side // 2, side - side // 2):
        for b in range(-side // 2, side - side // 2):
            if placed >= small:
                break
            c = (a + 0.9 * rng.random(), 0.2, b + 0.9 * rng.random())
            if math.dist(c, (4.0, 0.2, 0.0)) <= 0.9 or math.dist(c, (0, 0.2, 0)) < 1.3:
                continue
            dice = rng.random()
            if dice < 0.25:
                m, geom, sv = rng.choice(neon), master_cube, (0.4, 1.5 + 3.0 * rng.random(), 0.4)
            elif dice < 0.55:
                m, geom = (glass if rng.random() < 0.7 else bubble), master_sphere
                s = 0.5 + 0.5 * rng.random()
                sv = (s, s, s)
            else:
                m, geom = rng.choice(regular), (master_sphere if rng.random() < 0.5 else master_cube)
                s = 0.8 + 0.4 * rng.random()
                sv = (s, s, s)
            o = w.scale(geom, sv)
            if geom == master_cube:
                o = w.rotate("y", o, 90.0 * rng.random())  # radians, as the demo passes U[0,90) (Appendix A q3)
            objs.append(w.translate(w.matinst(o, m), c))
            placed += 1
    w.world(w.container("bvh", objs, name="world"))
    w.lines.append("envmap sky.hdr")
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=30.0,
        lookfrom=(10.0, 1.5, 0.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=3.0, focus_dist=9.0,
        env_mode="hdr", env_intensity=1.0, hdr_path="sky.hdr", hdri_rotation=0.5)
    return w.text(), cfg, files


def quirks(seed: int = 5, width: int = 640, height: int = 360, spp: int = 64, max_depth: int = 50, env: str = "sun"):
    """A small scene that puts every wrapper / primitive quirk of SURVEY.md Appendix A in view: glass under
    translate / rotate_y (front_face forced, q1) and under rotate_x / rotate_z / scale (kept, q2), degrees vs
    radians (q3), glass cubes (q5), (an off-origin cube, q4, is NOT here: its phantom hit about the origin lies
    outside its own bounding box, so whether the reference reports it depends on its random BVH order), loose triangles, a null material_instance (magenta),
    checker with negative coordinates (q9), metal fuzz 0 (q10), a light, nested list in bvh; PHYSICAL_SUN sky."""
    w = SceneWriter()
    glass = w.dielectric(1.5, tint=(0.9, 1.0, 0.95))
    red, blue = w.lambertian((0.8, 0.15, 0.1)), w.lambertian((0.1, 0.3, 0.8))
    mirror, gold = w.metal((0.95, 0.95, 0.95), 0.0), w.metal((1.0, 0.8, 0.4), 0.3)
    checker = w.lambertian(tex=w.tex_checker(0.7, w.tex_solid((0.15, 0.15, 0.15)), w.tex_solid((0.85, 0.85, 0.85))))
    lamp = w.light((4.0, 3.5, 3.0))
    s1 = w.sphere((0, 0, 0), 0.5, glass)
    c1 = w.cube((-0.4, -0.4, -0.4), (0.4, 0.4, 0.4), glass)
    objs = [
        w.sphere((0, -200, 0), 200.0, checker),
        w.sphere((-3.0, 0.5, -2.0), 0.5, glass),                                   # plain: front_face real
        w.translate(s1, (-1.8, 0.5, -2.0)),                                        # q1 forced
        w.translate(w.rotate("y", w.scale(s1, (1.0, 1.6, 0.7)), 0.6), (-0.4, 0.8, -2.0)),  # ellipsoid, q1 via rotate_y
        w.scale(w.sphere((1.0, 0.5, -2.0), 0.5, glass), (1.0, 1.0, 1.0)),          # q2 kept (identity scale instance)
        w.rotate("x", w.sphere((2.4, 0.5, -2.0), 0.5, glass), 0.0),               # q2 kept (degrees)
        w.translate(w.rotate("z", w.rotate("x", c1, 30.0), 20.0), (-2.6, 0.6, 0.0)),      # degrees
        w.translate(w.rotate("y", c1, 0.7), (-1.2, 0.4, 0.2)),                     # radians
        w.translate(w.matinst(w.cube1((0, 0, 0), None), mirror), (3.2, 1.0, -0.2)),
        w.cube((-0.25, -0.35, -0.25), (0.25, 0.35, 0.25), gold),                   # origin-centred plain cube
        w.matinst(w.sphere((1.9, 0.35, 0.9), 0.35, red), None),                    # magenta
        w.translate(w.matinst(w.scale(w.sphere((0, 0, 0), 0.3, None), (1.0, 0.5, 1.0)), blue), (0.3, 0.15, 1.4)),
        w.triangle((-2.2, 0.0, 1.2), (-1.2, 0.0, 1.6), (-1.7, 1.1, 1.3), (0, 0, 1), (0, 0, 1), (0.3, 0.2, 0.9), gold),
        w.translate(w.triangle((0, 0, 0), (0.8, 0, 0.1), (0.4, 0.9, 0), (0, 0, 1), (0, 0, 1), (0, 0, 1), blue), (-0.6, 0.0, 1.9)),
        w.sphere((0.0, 3.2, 0.0), 0.6, lamp),
        w.container("list", [w.sphere((2.6, 0.3, 1.6), 0.3, mirror), w.sphere((3.1, 0.25, 1.1), 0.25, blue)]),
    ]
    w.world(w.container("bvh", objs, name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=38.0,
        lookfrom=(0.5, 2.2, 7.5), lookat=(0.2, 0.5, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0, focus_dist=8.0,
        env_mode=env, background=(0.6, 0.7, 0.9), env_intensity=1.0, sun_direction=(0.6, 0.45, 0.5),
        sun_color=(1.0, 0.95, 0.85), sun_intensity=1.0, sun_size=1.0)
    return w.text(), cfg, {}


def shipped_models(seed: int = 6, width: int = 640, height: int = 360, spp: int = 64, max_depth: int = 50):
    """The reference's five shipped OBJ models (assets/models/, loaded with the scales of sceneAssetsLoader,
    scene_management.hpp:36-43: torus 0.4, bowl 0.02, cylinder / pyramid / torus_knot 0.4) on a ground sphere, each under
    the wrapper chains the demo scene uses (material_instance, rotate_x / rotate_y, translate).  cylinder.obj carries
    quads and n-gons (TinyObj fan triangulation), bowl.obj coincident faces, pyramid.obj no normals (flat shading)."""
    rng = random.Random(seed)
    w = SceneWriter()
    names = ("torus", "bowl", "cylinder", "pyramid", "torus_knot")
    files = {f"assets/models/{n}.obj": shipped_asset(f"models/{n}.obj") for n in names}
    scales = {"torus": 0.4, "bowl": 0.02, "cylinder": 0.4, "pyramid": 0.4, "torus_knot": 0.4}
    models = {n: w.model(f"assets/models/{n}.obj", None, scales[n]) for n in names}
    mats = [w.lambertian((0.8, 0.1, 0.1)), w.metal((1.0, 0.8, 0.4), 0.0), w.dielectric(1.5), w.metal((0.89, 0.58, 0.51), 0.2),
            w.lambertian((0.1, 0.4, 0.9)), w.light((0.0, 2.0, 4.0))]
    objs = [w.sphere((0, -1000, 0), 1000.0, w.lambertian((0.5, 0.5, 0.5)))]
    k = 0
    for row, z in enumerate((-2.2, 0.0, 2.2)):
        for col, n in enumerate(names):
            o = w.matinst(models[n], mats[k % len(mats)])
            if row == 1:
                o = w.rotate("y", o, rng.random() * 3.0)               # radians (Appendix A q3)
            elif row == 2:
                o = w.rotate("y", w.rotate("x", o, -30.0 + 60.0 * rng.random()), rng.random() * 3.0)  # degrees, then radians
            objs.append(w.translate(o, (-4.4 + 2.2 * col + 0.3 * rng.random(), 0.0 if row < 2 else 0.6, z)))
            k += 1
    w.world(w.container("bvh", objs, name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=34.0,
        lookfrom=(1.0, 6.5, 10.5), lookat=(0.0, 0.4, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0, focus_dist=10.0,
        env_mode="solid", background=(0.7, 0.8, 1.0), env_intensity=1.0)
    return w.text(), cfg, files


def kat_scene(env: str = "solid"):
    """Scene of the per-function known-answer tests (tests/golden/kat.json): every texture kind — solids, a checker,
    the reference's shipped JPEGs (baseline 4:2:0 fine-wood, progressive 4:4:4 wood bump, baseline scratches, progressive
    concrete) — and one bump-mapped material of each type, on spheres so that the flattener keeps them; `env` picks the
    environment whose get_background_color is probed (solid | hdr with yaw / tilt / roll | sun)."""
    from . import ztex
    w = SceneWriter()
    names = {"wood": "assets/textures/fine-wood.jpg", "woodbump": "assets/bump_maps/wood_bump_map.jpg",
             "scratches": "assets/bump_maps/scratches_bump_map.jpg", "concrete": "assets/bump_maps/concrete_bump_map.jpg"}
    files = {p: shipped_asset(p[len("assets/"):]) for p in names.values()}
    files.update(ztex.hdr_pair("sky.hdr", _sky_hdr(256, 128, 60.0)))
    tex = {k: w.tex_image(p, name="t_" + k) for k, p in names.items()}
    checker = w.tex_checker(0.7, w.tex_solid((0.15, 0.2, 0.1), name="t_odd"), w.tex_solid((0.9, 0.85, 0.8), name="t_even"), name="t_checker")
    mats = [w.lambertian(tex=tex["wood"], bump=tex["woodbump"], strength=8.0, name="m_wood"),
            w.metal((1.0, 0.8, 0.4), 0.0, bump=tex["scratches"], strength=-1.0, name="m_gold"),
            w.dielectric(1.5, bump=tex["concrete"], strength=0.02, name="m_glass"),
            w.lambertian(tex=checker, name="m_checker")]
    objs = [w.sphere((2.0 * k - 3.0, 0.5, 0.0), 0.5, m) for k, m in enumerate(mats)]
    w.world(w.container("bvh", objs, name="world"))
    w.lines.append("envmap sky.hdr")
    cfg = RenderConfig(image_width=64, image_height=36, samples_per_pixel=1, env_mode=env, background=(0.3, 0.5, 0.9), env_intensity=1.25,
                       hdr_path="sky.hdr", hdri_rotation=0.5, hdri_tilt=-0.2, hdri_roll=0.3, sun_direction=(0.6, 0.45, 0.5),
                       sun_color=(1.0, 0.95, 0.85), sun_intensity=1.5, sun_size=2.0)
    return w.text(), cfg, files


def refit_demo(move: float = 0.0, width: int = 320, height: int = 180, spp: int = 16):
    """Scene of the zb_scene_refit test: plain spheres (their items never move), and instances of every geometry kind —
    cube under translate(rotate_y), sphere under translate(scale), two shipped OBJ meshes under translate(rotate_x(scale)),
    a fog whose boundary is a scaled sphere — whose wrapper parameters depend on `move`; the object graph (and therefore
    the instance table's geometry references) is the same for every value of `move`."""
    w = SceneWriter()
    files = {"assets/models/torus.obj": shipped_asset("models/torus.obj"), "assets/models/pyramid.obj": shipped_asset("models/pyramid.obj")}
    grey, red, gold, glass = w.lambertian((0.5, 0.5, 0.5)), w.lambertian((0.8, 0.15, 0.1)), w.metal((1.0, 0.8, 0.4), 0.1), w.dielectric(1.5)
    objs = [w.sphere((0, -300, 0), 300.0, grey), w.sphere((-3.5, 0.5, 1.0), 0.5, gold), w.sphere((3.6, 0.4, 1.4), 0.4, red),
            w.sphere((0.0, 4.0, 0.0), 0.7, w.light((5.0, 4.5, 4.0)))]
    m = move
    box = w.cube((-0.4, -0.4, -0.4), (0.4, 0.4, 0.4), red)
    objs.append(w.translate(w.rotate("y", box, 0.3 + 0.9 * m), (-2.0 + 1.5 * m, 0.4, -0.5 + 0.8 * m)))
    objs.append(w.translate(w.rotate("y", box, 1.1 - 0.5 * m), (2.2 - 0.7 * m, 0.4, -1.0)))
    ball = w.sphere((0, 0, 0), 0.5, glass)
    objs.append(w.translate(w.scale(ball, (1.0 + 0.5 * m, 0.7, 1.0)), (0.0 + 0.6 * m, 0.35 + 0.3 * m, 1.2)))
    torus = w.model("assets/models/torus.obj", None, 0.25)
    pyramid = w.model("assets/models/pyramid.obj", None, 0.4)
    objs.append(w.translate(w.rotate("x", w.scale(w.matinst(torus, gold), (1.0, 1.0 + 0.5 * m, 1.0)), 20.0 + 40.0 * m), (-0.8 - 0.9 * m, 0.5, -1.8 + 0.5 * m)))
    objs.append(w.translate(w.rotate("x", w.scale(w.matinst(torus, red), (0.8, 0.8, 0.8)), -15.0 * m), (1.0 + 0.8 * m, 0.3 + 0.4 * m, 0.2)))
    objs.append(w.translate(w.rotate("y", w.matinst(pyramid, gold), 0.4 + m), (3.0 - 1.2 * m, 0.0, -2.2 + 0.6 * m)))
    room = w.container("bvh", objs, name="room")
    fog = w.medium(w.scale(w.sphere((0, 0.5, 0), 2.5, None), (1.0 + 0.4 * m, 0.6, 1.0)), 0.05, rgb=(0.9, 0.9, 1.0))
    w.world(w.container("list", [room, fog], name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=20, vfov=38.0, lookfrom=(0.5, 2.4, 7.0),
        lookat=(0.2, 0.4, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0, focus_dist=8.0, env_mode="solid", background=(0.6, 0.7, 0.9))
    return w.text(), cfg, files


def wire_demo(seed: int = 7, width: int = 240, height: int = 135, spp: int = 64, max_depth: int = 50):
    """Scene of the BVH debug-view tests: 24 seeded spheres, two cubes and a lamp inside a region
    the camera sees whole, so the root box of the tree (level 0 — the one level that does not depend on how the
    tree was built) is in frame; one bvh root."""
    rng = random.Random(seed)
    w = SceneWriter()
    mats = [w.lambertian((0.7, 0.3, 0.2)), w.metal((0.8, 0.8, 0.9), 0.1), w.dielectric(1.5), w.lambertian((0.2, 0.5, 0.7))]
    objs = []
    for i in range(24):
        c = (rng.uniform(-2.0, 2.0), rng.uniform(0.2, 1.4), rng.uniform(-2.0, 2.0))
        objs.append(w.sphere(c, rng.uniform(0.12, 0.3), mats[i % 4]))
    objs.append(w.translate(w.cube((-0.3, -0.3, -0.3), (0.3, 0.3, 0.3), mats[0]), (0.8, 0.3, 0.4)))
    objs.append(w.translate(w.rotate("y", w.cube((-0.2, -0.4, -0.2), (0.2, 0.4, 0.2), mats[1]), 0.5), (-1.1, 0.4, -0.6)))
    objs.append(w.sphere((0.0, 1.9, 0.0), 0.25, w.light((6.0, 5.0, 4.0))))
    w.world(w.container("bvh", objs, name="world"))
    cfg = RenderConfig(
        image_width=width, image_height=height, samples_per_pixel=spp, max_depth=max_depth, vfov=36.0,
        lookfrom=(6.0, 4.5, 7.5), lookat=(0.0, 0.8, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0, focus_dist=10.0,
        env_mode="solid", background=(0.7, 0.8, 1.0), env_intensity=1.0)
    return w.text(), cfg, {}
