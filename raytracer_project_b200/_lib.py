"""ctypes binding of libzenith_b200.so (the C ABI in include/zenith_b200.h).

The library is built in-tree by raytracer_project_b200/build.sh (or __graft_entry__.build()).
There is no Python / CPU fallback: if the library is missing, loading raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# ZB_LIB selects an alternative build of the same library (kernel tuning variants under _build/)
LIB_PATH = os.environ.get("ZB_LIB") or os.path.join(_HERE, "libzenith_b200.so")

ZB_OK = 0
ZB_NO_TEXTURE = 0xFFFFFFFF
ENV_MODES = {"sun": 0, "hdr": 1, "solid": 2}


class ZbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"zenith_b200 error {code}: {msg}")
        self.code = code


class FlatScene(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32), ("n_nodes", C.c_uint32), ("n_tlas_nodes", C.c_uint32),
        ("n_items", C.c_uint32), ("n_cubes", C.c_uint32), ("n_tris", C.c_uint32), ("n_instances", C.c_uint32),
        ("n_meshes", C.c_uint32), ("n_media", C.c_uint32), ("n_materials", C.c_uint32), ("n_textures", C.c_uint32),
        ("n_spheres_obj", C.c_uint32), ("n_texels_u8", C.c_uint64), ("n_texels_f32", C.c_uint64),
        ("nodes", C.c_void_p), ("items", C.c_void_p), ("cubes", C.c_void_p), ("tris", C.c_void_p),
        ("tri_normals", C.c_void_p), ("instances", C.c_void_p), ("meshes", C.c_void_p), ("media", C.c_void_p),
        ("materials", C.c_void_p), ("textures", C.c_void_p), ("spheres_obj", C.c_void_p),
        ("texels_u8", C.c_void_p), ("texels_f32", C.c_void_p), ("n_leaves", C.c_uint32), ("pad", C.c_uint32),
        ("n_wide_nodes", C.c_uint32), ("wide_stack_need", C.c_uint32), ("wide_nodes", C.c_void_p), ("item_bounds", C.c_void_p),
    ]

    def meshes_array(self):
        """zb_mesh records as an (n_meshes, 4) uint32 view: node_root, node_count, tri_first, tri_count."""
        class Mesh(C.Structure):
            _fields_ = [("node_root", C.c_uint32), ("node_count", C.c_uint32), ("tri_first", C.c_uint32), ("tri_count", C.c_uint32)]
        return C.cast(self.meshes, C.POINTER(Mesh))

    def nbytes(self) -> int:
        """Bytes zb_scene_upload copies to the device for this scene."""
        return (self.n_nodes * 64 + self.n_wide_nodes * 80 + self.n_items * 32 + self.n_cubes * 32 + self.n_tris * 96 + self.n_instances * 128
                + self.n_media * 32 + self.n_materials * 32 + self.n_textures * 32 + self.n_spheres_obj * 16
                + self.n_texels_u8 * 4 + self.n_texels_f32 * 16)


class Camera(C.Structure):
    _fields_ = [("image_width", C.c_int32), ("image_height", C.c_int32), ("vfov", C.c_double),
                ("lookfrom", C.c_double * 3), ("lookat", C.c_double * 3), ("vup", C.c_double * 3),
                ("defocus_angle", C.c_double), ("focus_dist", C.c_double)]


class Environment(C.Structure):
    _fields_ = [("mode", C.c_int32), ("hdr_texture", C.c_uint32), ("background_color", C.c_double * 3),
                ("intensity", C.c_double), ("hdri_rotation", C.c_double), ("hdri_tilt", C.c_double),
                ("hdri_roll", C.c_double), ("sun_direction", C.c_double * 3), ("sun_color", C.c_double * 3),
                ("sun_intensity", C.c_double), ("sun_size", C.c_double)]


class Probe(C.Structure):
    _fields_ = [("t", C.c_float), ("leaf_id", C.c_uint32), ("prim_id", C.c_uint32), ("front_face", C.c_uint32),
                ("p", C.c_float * 3), ("normal", C.c_float * 3), ("u", C.c_float), ("v", C.c_float),
                ("albedo", C.c_float * 3), ("emitted", C.c_float * 3)]


class RenderOpts(C.Structure):
    _fields_ = [("samples_per_pixel", C.c_int32), ("max_depth", C.c_int32), ("seed", C.c_uint64),
                ("sample_begin", C.c_int32), ("sample_end", C.c_int32), ("tile_rank", C.c_int32),
                ("tile_ranks", C.c_int32), ("use_albedo", C.c_int32), ("use_normal", C.c_int32),
                ("use_z_depth", C.c_int32), ("z_depth_max_dist", C.c_double), ("collect_stats", C.c_int32),
                ("samples_per_launch", C.c_int32), ("use_reflection", C.c_int32), ("use_refraction", C.c_int32),
                ("light_pass", C.c_int32), ("accumulate", C.c_int32), ("partition", C.c_int32), ("reserved", C.c_int32)]


class GroupHandle(C.Structure):
    _fields_ = [("opaque", C.c_ubyte * 256)]


class Outputs(C.Structure):
    _fields_ = [("beauty", C.c_void_p), ("albedo", C.c_void_p), ("normal", C.c_void_p), ("z_depth", C.c_void_p),
                ("beauty_f4", C.c_void_p), ("reflection", C.c_void_p), ("refraction", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("node_visits", C.c_uint64), ("box_tests", C.c_uint64),
                ("prim_tests", C.c_uint64), ("instance_visits", C.c_uint64), ("kernel_ms", C.c_double),
                ("resolve_ms", C.c_double), ("h2d_ms", C.c_double), ("d2h_ms", C.c_double), ("launches", C.c_uint32),
                ("stack_overflows", C.c_uint32), ("kernel_pick", C.c_uint32), ("n_gpus", C.c_uint32)]


PROGRESS_FN = C.CFUNCTYPE(None, C.c_int, C.c_void_p)

class Post(C.Structure):
    """zb_post: the fields of post_processor (color_processing.hpp:44-74)."""
    _fields_ = [("exposure", C.c_float), ("saturation", C.c_float), ("contrast", C.c_float), ("hue_shift", C.c_float),
                ("vignette_intensity", C.c_float), ("color_balance", C.c_float * 3),
                ("exposure_compensation_stops", C.c_float), ("target_luminance", C.c_float),
                ("use_aces_tone_mapping", C.c_int32), ("use_auto_exposure", C.c_int32), ("debug_red", C.c_int32),
                ("debug_green", C.c_int32), ("debug_blue", C.c_int32), ("debug_luminance", C.c_int32),
                ("debug_bvh", C.c_int32), ("use_bloom", C.c_int32), ("bloom_threshold", C.c_float),
                ("bloom_intensity", C.c_float), ("bloom_radius", C.c_int32), ("use_sharpening", C.c_int32),
                ("sharpen_amount", C.c_double)]


class ImageStats(C.Structure):
    _fields_ = [("average_luminance", C.c_float), ("max_luminance", C.c_float), ("histogram", C.c_int32 * 256)]


# every symbol include/zenith_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "zb_abi_version": (C.c_int, []),
    "zb_last_error": (C.c_char_p, []),
    "zb_context_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "zb_context_create_multi": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]),
    "zb_context_device_count": (C.c_int, [C.c_void_p]),
    "zb_group_export": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(GroupHandle)]),
    "zb_group_join": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(GroupHandle), C.c_char_p]),
    "zb_group_barrier": (C.c_int, [C.c_void_p]),
    "zb_group_leave": (C.c_int, [C.c_void_p]),
    "zb_context_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "zb_context_destroy": (C.c_int, [C.c_void_p]),
    "zb_context_set_tuning": (C.c_int, [C.c_void_p, C.c_char_p, C.c_longlong]),
    "zb_build_flags": (C.c_int, []),
    "zb_device_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_size_t)]),
    "zb_scene_upload": (C.c_int, [C.c_void_p, C.POINTER(FlatScene), C.POINTER(C.c_void_p)]),
    "zb_scene_destroy": (C.c_int, [C.c_void_p]),
    "zb_context_register_output": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "zb_context_unregister_output": (C.c_int, [C.c_void_p, C.c_void_p]),
    "zb_scene_refit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32]),
    "zb_scene_rebuild_tlas": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32]),
    "zb_render": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.POINTER(Environment), C.POINTER(RenderOpts),
                            C.POINTER(Outputs), C.POINTER(C.c_int), PROGRESS_FN, C.c_void_p]),
    "zb_probe_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Environment), C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zb_kat_eval": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Environment), C.c_size_t, C.c_void_p, C.c_void_p]),
    "zb_read_frame": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "zb_render_device": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.POINTER(Environment),
                                   C.POINTER(RenderOpts), C.c_void_p, C.c_void_p]),
    "zb_render_device_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.POINTER(Environment),
                                           C.POINTER(RenderOpts), C.c_void_p]),
    "zb_noalloc_max_pixels": (C.c_size_t, []),
    "zb_resolve_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_void_p, C.c_void_p]),
    "zb_primary_ids": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.c_void_p, C.c_void_p, C.c_void_p]),
    "zb_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "zb_checkpoint_save": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64]),
    "zb_checkpoint_load": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_uint64)]),
    "zb_analyze_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(ImageStats)]),
    "zb_analyze_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(ImageStats)]),
    "zb_auto_exposure": (C.c_double, [C.POINTER(Post), C.POINTER(ImageStats)]),
    "zb_finish_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Post), C.c_int, C.c_void_p, C.c_void_p]),
    "zb_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Post), C.c_int, C.c_void_p]),
    "zb_render_bvh_debug_device": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.POINTER(Environment), C.POINTER(RenderOpts), C.c_int,
                                   C.c_float, C.c_void_p]),
    "zb_render_bvh_debug": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.POINTER(Environment), C.POINTER(RenderOpts), C.c_int,
                            C.c_float, C.c_void_p]),
    "zb_save_pass_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Post), C.c_int, C.c_char_p]),
    "zb_pass_file_name": (C.c_char_p, [C.c_int]),
    "zb_save_pass": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Post), C.c_int, C.c_char_p]),
    "zb_save_hdr": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_char_p]),
    "zb_write_png": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_size_t]),
    "zb_write_hdr": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_size_t]),
    "zb_flat_scene_from_zscene": (C.c_int, [C.c_char_p, C.POINTER(C.POINTER(FlatScene))]),
    "zb_flat_scene_from_zscene_text": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.POINTER(FlatScene))]),
    "zb_flat_scene_from_zscene_ex": (C.c_int, [C.c_char_p, C.c_char_p, C.c_size_t, C.c_uint32, C.POINTER(C.POINTER(FlatScene))]),
    "zb_flat_scene_features": (C.c_int, [C.POINTER(FlatScene), C.POINTER(C.c_uint32)]),
    "zb_flat_scene_save": (C.c_int, [C.POINTER(FlatScene), C.c_char_p]),
    "zb_flat_scene_load": (C.c_int, [C.c_char_p, C.POINTER(C.POINTER(FlatScene))]),
    "zb_flat_scene_free": (C.c_int, [C.POINTER(FlatScene)]),
}

_lib = None


def load():
    """Load the native library (fails loudly if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run raytracer_project_b200/build.sh (or __graft_entry__.build()); "
                          "there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != ZB_OK:
        raise ZbError(rc, load().zb_last_error().decode("utf-8", "replace"))
