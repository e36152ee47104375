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
        self.lines.append(f"obj {name} sphere {_f(c[0])} {_f(c[1])} {_f(c[2])} {_f(r)} {mat}")
        return name

    def cube(self, lo, hi, mat, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} cube {' '.join(_f(x) for x in lo)} {' '.join(_f(x) for x in hi)} {mat}")
        return name

    def cube1(self, c, mat, name=None):
        name = name or self._name("o")
        self.lines.append(f"obj {name} cube1 {' '.join(_f(x) for x in c)} {mat}")
        return name

    def triangle(self, a, b, c, n0, n1, n2, mat, name=None):
        name = name or self._name("o")
        nums = " ".join(_f(x) for v in (a, b, c, n0, n1, n2) for x in v)
        self.lines.append(f"obj {name} triangle {nums} {mat}")
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
