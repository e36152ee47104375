"""zenith-b200: B200-native backend for the path-tracing hot path of Zenith (jarek1992/raytracer_project).

Layout: csrc/ (sm_100a CUDA kernels + C ABI), host/ (C++20 scene API, BVH builder, flattener),
scenes.py (seeded scene generators -> zscene text), api.py (ctypes driver used by tests and bench).
"""
from . import scenes  # noqa: F401

__all__ = ["scenes"]
