"""Reduced-size instances of BASELINE.json configs C2-C4 (+ the Appendix-A quirk scene) used for the golden
fixtures: name -> (generator in raytracer_project_b200.scenes, generator kwargs, ids WxH, stats WxH,
extra kwargs for the primary-id fixture).  The id fixture of the Cornell room leaves the fog out: a
constant_medium hit is a random event (constant_medium.hpp:64), so it has no deterministic id."""
GOLDEN_SCENES = {
    "cornell": ("cornell", dict(seed=2), (640, 360), (200, 112), dict(fog=False)),
    "mesh": ("mesh", dict(seed=3, grid=120, knots=8), (640, 360), (200, 112), {}),
    "exterior": ("exterior", dict(seed=4, small=120, hdr_size=512), (640, 360), (200, 112), {}),
    "quirks": ("quirks", dict(seed=5), (640, 360), (200, 112), {}),
}
