"""Thin Python driver over the C ABI, used by tests/ and bench.py.

Mirrors the reference's call sequence (main.cpp:1500-1521): build scene -> (flatten + upload) ->
camera parameters -> render -> six row-major double-RGB buffers (camera.hpp:63,80-85).
All compute happens in libzenith_b200.so on the GPU; nothing here falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import Camera, Environment, FlatScene, Outputs, RenderOpts, Stats, check
from .scenes import RenderConfig


def flatten_zscene(text: str | None = None, path: str | None = None, gpu_blas: bool = False, wide: int = 0):
    """zscene text/file -> host scene API -> flatten(); returns a ctypes pointer to zb_flat_scene.
    gpu_blas: leave the mesh BLAS builds to the device at zb_scene_upload (binned SAH, csrc/sah_build.cu; tuning
    "gpu_builder" 0 = linear BVH, csrc/lbvh.cu).
    wide: 1 = also build the 8-wide compressed tree for scenes with meshes, 2 = for every scene (A/B; off by default)."""
    lib = _lib.load()
    out = C.POINTER(FlatScene)()
    if gpu_blas or wide:
        data = text.encode() if text is not None else None
        flags = (1 if gpu_blas else 0) | (2 if wide == 1 else 0) | (4 if wide >= 2 else 0)
        check(lib.zb_flat_scene_from_zscene_ex(path.encode() if path is not None else None, data, len(data) if data else 0,
                                               flags, C.byref(out)))
    elif path is not None:
        check(lib.zb_flat_scene_from_zscene(path.encode(), C.byref(out)))
    else:
        data = text.encode()
        check(lib.zb_flat_scene_from_zscene_text(data, len(data), C.byref(out)))
    return out


FEATURES = {"INST": 1, "MESH": 2, "CUBE": 4, "TRI": 8, "MEDIA": 16, "TEX": 32, "ENV": 64}


def flat_features(flat) -> int:
    """zb_flat_scene_features: the ZB_F_* bits the scene needs (the render kernel runs the smallest compiled superset)."""
    out = C.c_uint32(0)
    check(_lib.load().zb_flat_scene_features(flat, C.byref(out)))
    return out.value


def save_flat(flat, path: str):
    """zb_flat_scene_save: the flattened scene as a ZBFS file."""
    check(_lib.load().zb_flat_scene_save(flat, path.encode()))


def load_flat(path: str):
    """zb_flat_scene_load: a ZBFS file -> zb_flat_scene (free with free_flat)."""
    out = C.POINTER(FlatScene)()
    check(_lib.load().zb_flat_scene_load(path.encode(), C.byref(out)))
    return out


def free_flat(flat):
    _lib.load().zb_flat_scene_free(flat)


def camera_struct(cfg: RenderConfig) -> Camera:
    cam = Camera()
    cam.image_width, cam.image_height = cfg.image_width, cfg.image_height
    cam.vfov = cfg.vfov
    cam.lookfrom[:] = cfg.lookfrom
    cam.lookat[:] = cfg.lookat
    cam.vup[:] = cfg.vup
    cam.defocus_angle, cam.focus_dist = cfg.defocus_angle, cfg.focus_dist
    return cam


def env_struct(cfg: RenderConfig) -> Environment:
    env = Environment()
    env.mode = _lib.ENV_MODES[cfg.env_mode]
    env.hdr_texture = _lib.ZB_NO_TEXTURE
    env.background_color[:] = cfg.background
    env.intensity = cfg.env_intensity
    env.hdri_rotation, env.hdri_tilt, env.hdri_roll = cfg.hdri_rotation, cfg.hdri_tilt, cfg.hdri_roll
    env.sun_direction[:] = cfg.sun_direction
    env.sun_color[:] = cfg.sun_color
    env.sun_intensity, env.sun_size = cfg.sun_intensity, cfg.sun_size
    return env


PARTITION = {"auto": 0, "samples": 1, "tiles": 2}


def opts_struct(cfg: RenderConfig, seed=1, sample_range=None, tile=None, stats=False, samples_per_launch=0,
                partition="auto") -> RenderOpts:
    o = RenderOpts()
    o.samples_per_pixel, o.max_depth, o.seed = cfg.samples_per_pixel, cfg.max_depth, seed
    o.sample_begin, o.sample_end = sample_range if sample_range else (0, cfg.samples_per_pixel)
    if sample_range and sample_range[0] == sample_range[1]:
        # an EMPTY share (a rank left without samples): (0, 0) would read as "all samples" in the C struct
        o.sample_begin = o.sample_end = cfg.samples_per_pixel
    o.partition = PARTITION[partition]
    o.tile_rank, o.tile_ranks = tile if tile else (0, 1)
    o.use_albedo, o.use_normal, o.use_z_depth = int(cfg.use_albedo), int(cfg.use_normal), int(cfg.use_z_depth)
    o.z_depth_max_dist = cfg.z_depth_max_dist
    o.collect_stats = int(stats)
    o.samples_per_launch = samples_per_launch
    o.use_reflection, o.use_refraction = int(cfg.use_reflection), int(cfg.use_refraction)
    return o


POST_DEFAULTS = dict(exposure=0.5, saturation=1.0, contrast=1.0, hue_shift=0.0, vignette_intensity=1.0,
                     color_balance=(1.0, 1.0, 1.0), exposure_compensation_stops=0.0, target_luminance=0.12, use_aces=0,
                     use_auto_exposure=0, debug_red=0, debug_green=0, debug_blue=0, debug_luminance=0, debug_bvh=0,
                     use_bloom=0, bloom_threshold=1.0, bloom_intensity=0.3, bloom_radius=4, use_sharpening=0,
                     sharpen_amount=0.2)  # post_processor defaults (color_processing.hpp:44-74)


def post_struct(**kw) -> "_lib.Post":
    """zb_post from keyword overrides of the reference's post_processor defaults."""
    v = {**POST_DEFAULTS, **kw}
    p = _lib.Post()
    for k in ("exposure", "saturation", "contrast", "hue_shift", "vignette_intensity", "exposure_compensation_stops",
              "target_luminance", "bloom_threshold", "bloom_intensity", "sharpen_amount"):
        setattr(p, k, float(v[k]))
    for k in ("use_auto_exposure", "debug_red", "debug_green", "debug_blue", "debug_luminance", "debug_bvh", "use_bloom",
              "bloom_radius", "use_sharpening"):
        setattr(p, k, int(v[k]))
    p.use_aces_tone_mapping = int(v["use_aces"])
    p.color_balance = (C.c_float * 3)(*[float(x) for x in v["color_balance"]])
    return p


RENDER_PASS = {"RGB": 0, "Denoise": 1, "Albedo": 2, "Normals": 3, "Reflections": 4, "Refractions": 5, "Z-Depth": 6}


def pass_file_name(render_pass: int) -> str:
    """"render_<Pass>.png" as the reference's export names its files (main.cpp:1317)."""
    name = _lib.load().zb_pass_file_name(render_pass)
    if name is None:
        raise ValueError(f"render_pass {render_pass} outside the reference's enum")
    return name.decode()


def write_png(path: str, rgb) -> None:
    """(H, W, 3 | 4) uint8 -> 8-bit RGB PNG (alpha dropped)."""
    a = np.ascontiguousarray(rgb, np.uint8)
    check(_lib.load().zb_write_png(path.encode(), a.shape[1], a.shape[0], a.ctypes.data, a.shape[2], 0))


def write_hdr(path: str, rgb) -> None:
    """(H, W, >= 3) float32 linear radiance -> flat Radiance RGBE file."""
    a = np.ascontiguousarray(rgb, np.float32)
    check(_lib.load().zb_write_hdr(path.encode(), a.shape[1], a.shape[0], a.ctypes.data, a.shape[2], 0))


def read_png(path: str):
    """Decoder for the tests: 8-bit RGB / RGBA, non-interlaced PNG -> (H, W, C) uint8 (all five row filters)."""
    import struct
    import zlib
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n", "not a PNG"
    at, idat, w = 8, b"", 0
    while at < len(data):
        n, kind = struct.unpack(">I4s", data[at:at + 8])
        body = data[at + 8:at + 8 + n]
        assert zlib.crc32(kind + body) == struct.unpack(">I", data[at + 8 + n:at + 12 + n])[0], "PNG chunk CRC"
        if kind == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
            assert depth == 8 and ctype in (2, 6) and interlace == 0
            ch = 3 if ctype == 2 else 4
        elif kind == b"IDAT":
            idat += body
        at += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * ch + 1)
    out = np.zeros((h, w * ch), np.uint8)
    for y in range(h):
        f, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        up = out[y - 1].astype(np.int32) if y else np.zeros(w * ch, np.int32)
        if f == 0:
            cur = line
        elif f == 2:
            cur = line + up
        else:
            cur = np.zeros(w * ch, np.int32)
            for i in range(w * ch):
                a = cur[i - ch] if i >= ch else 0
                b = up[i]
                c = up[i - ch] if i >= ch else 0
                if f == 1:
                    pred = a
                elif f == 3:
                    pred = (a + b) // 2
                else:
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pred = a if pa <= pb and pa <= pc else (b if pb <= pc else c)
                cur[i] = (line[i] + pred) & 255
        out[y] = (cur & 255).astype(np.uint8)
    return out.reshape(h, w, ch)


def read_hdr(path: str):
    """Decoder for the tests: flat Radiance RGBE -> (H, W, 3) float32, with stb_image's conversion (ldexp(1, e - 136))."""
    data = open(path, "rb").read()
    head, _, rest = data.partition(b"\n\n")
    assert head.startswith(b"#?RADIANCE")
    dims, _, pix = rest.partition(b"\n")
    tok = dims.split()
    h, w = int(tok[1]), int(tok[3])
    q = np.frombuffer(pix[:w * h * 4], np.uint8).reshape(h, w, 4).astype(np.float32)
    scale = np.where(q[..., 3:] > 0, np.exp2(q[..., 3:] - 136.0), 0.0)
    return (q[..., :3] * scale).astype(np.float32)


def post_text(pass_kind=0, **kw) -> str:
    """The same parameters as the key-value text the oracle binaries read (`finish` command)."""
    v = {**POST_DEFAULTS, **kw}
    lines = [f"pass {int(pass_kind)}"]
    for k, x in v.items():
        lines.append(f"{k} " + (" ".join(repr(float(y)) for y in x) if isinstance(x, (tuple, list)) else repr(float(x))))
    return "\n".join(lines) + "\n"


def has_experiments() -> bool:
    """True when the library was built with -DZB_EXPERIMENTS (wavefront kernels, k_render_pqr)."""
    return bool(_lib.load().zb_build_flags() & 1)


class Context:
    def __init__(self, device: int | list = 0, stream: int | None = None):
        """device: one index, or a list of indices for a multi-GPU context (zb_context_create_multi: one process drives
        all of them; upload / render / primary_ids are used exactly as on one GPU)."""
        self.lib = _lib.load()
        self.h = C.c_void_p()
        if isinstance(device, (list, tuple)):
            arr = (C.c_int * len(device))(*device)
            check(self.lib.zb_context_create_multi(arr, len(device), C.byref(self.h)))
        else:
            check(self.lib.zb_context_create(device, C.byref(self.h)))
        if stream is not None:
            check(self.lib.zb_context_set_stream(self.h, C.c_void_p(stream)))

    def close(self):
        if self.h:
            self.lib.zb_context_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_tuning(self, key: str, value: int):
        """zb_context_set_tuning: A/B switches ("pq", "no_lean", "pq_blocks", ...); the product defaults need none."""
        check(self.lib.zb_context_set_tuning(self.h, key.encode(), int(value)))

    def register_output(self, array) -> bool:
        """zb_context_register_output for a numpy array the caller keeps alive and passes to render(beauty_out=...) /
        render_into() repeatedly: finished rows are then DMA'd straight into it.  False when it cannot be page-locked."""
        try:
            check(self.lib.zb_context_register_output(self.h, array.ctypes.data, array.nbytes))
            return True
        except _lib.ZbError:
            return False

    def unregister_output(self, array):
        check(self.lib.zb_context_unregister_output(self.h, array.ctypes.data))

    def device_count(self) -> int:
        return int(self.lib.zb_context_device_count(self.h))

    # ---- one process per GPU (torchrun): CUDA IPC handles carried by the caller's transport
    def group_export(self, max_pixels: int) -> bytes:
        h = _lib.GroupHandle()
        check(self.lib.zb_group_export(self.h, max_pixels, C.byref(h)))
        return bytes(h.opaque)

    def group_join(self, rank: int, handles: list, shm_name: str):
        arr = (_lib.GroupHandle * len(handles))()
        for i, b in enumerate(handles):
            C.memmove(C.byref(arr[i]), b, 256)
        check(self.lib.zb_group_join(self.h, rank, len(handles), arr, shm_name.encode()))

    def group_barrier(self):
        """zb_group_barrier: synchronise this context's stream, then wait for the other ranks of the process group."""
        check(self.lib.zb_group_barrier(self.h))

    def group_leave(self):
        check(self.lib.zb_group_leave(self.h))

    def device_info(self):
        sm, ma, mi, l2 = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
        check(self.lib.zb_device_info(self.h, C.byref(sm), C.byref(ma), C.byref(mi), C.byref(l2)))
        return {"sm_count": sm.value, "cc": (ma.value, mi.value), "l2_bytes": l2.value}

    def analyze_device(self, mean_ptr: int, n_pixels: int) -> dict:
        """zb_analyze_device: post_processor::analyze_framebuffer over device float4 means."""
        st = _lib.ImageStats()
        check(self.lib.zb_analyze_device(self.h, C.c_void_p(mean_ptr), n_pixels, C.byref(st)))
        return {"average_luminance": st.average_luminance, "max_luminance": st.max_luminance,
                "histogram": np.array(st.histogram, np.int32), "_struct": st}

    def auto_exposure(self, post, stats: dict) -> float:
        return float(self.lib.zb_auto_exposure(C.byref(post), C.byref(stats["_struct"])))

    def finish_device(self, mean_ptr: int, width: int, height: int, post, pass_kind=0, out_f4_ptr=0, out_u8_ptr=0):
        """zb_finish_device: update_post_processing on device buffers (asynchronous on the context stream)."""
        check(self.lib.zb_finish_device(self.h, C.c_void_p(mean_ptr), width, height, C.byref(post), pass_kind,
                                        C.c_void_p(out_f4_ptr) if out_f4_ptr else None,
                                        C.c_void_p(out_u8_ptr) if out_u8_ptr else None))

    def finish(self, image, post, pass_kind=0):
        """zb_finish: (H, W, 3) float64 linear image -> the reference's final_framebuffer (H, W, 3) float64."""
        img = np.ascontiguousarray(image, np.float64)
        out = np.empty_like(img)
        check(self.lib.zb_finish(self.h, img.ctypes.data, img.shape[1], img.shape[0], C.byref(post), pass_kind, out.ctypes.data))
        return out

    def save_pass(self, path: str, mean_ptr: int, width: int, height: int, post, render_pass=0):
        """zb_save_pass: camera::save_render_pass for device-resident means -> 8-bit RGB PNG."""
        check(self.lib.zb_save_pass(self.h, C.c_void_p(mean_ptr), width, height, C.byref(post), render_pass, path.encode()))

    def save_hdr(self, path: str, mean_ptr: int, width: int, height: int):
        check(self.lib.zb_save_hdr(self.h, C.c_void_p(mean_ptr), width, height, path.encode()))

    def checkpoint_save(self, path: str, accum_ptr: int, width: int, height: int, samples_done: int, seed: int):
        check(self.lib.zb_checkpoint_save(self.h, path.encode(), C.c_void_p(accum_ptr), width, height, samples_done, seed))

    def checkpoint_load(self, path: str, accum_ptr: int, width: int, height: int):
        done, seed = C.c_int(), C.c_uint64()
        check(self.lib.zb_checkpoint_load(self.h, path.encode(), C.c_void_p(accum_ptr), width, height, C.byref(done), C.byref(seed)))
        return done.value, seed.value

    def upload(self, flat) -> "Scene":
        s = C.c_void_p()
        check(self.lib.zb_scene_upload(self.h, flat, C.byref(s)))
        return Scene(self, s)

    def stats(self) -> dict:
        st = Stats()
        check(self.lib.zb_get_stats(self.h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in Stats._fields_}


class Scene:
    def __init__(self, ctx: Context, handle):
        self.ctx, self.h = ctx, handle

    def close(self):
        if self.h:
            self.ctx.lib.zb_scene_destroy(self.h)
            self.h = C.c_void_p()

    def render(self, cfg: RenderConfig, seed=1, stats=False, want_f4=False, sample_range=None, tile=None,
               samples_per_launch=0, progress=None, cancel=None, beauty_out=None, partition="auto", no_host=False):
        """zb_render: returns dict of (H, W, 3) float64 arrays like the reference's buffers.  beauty_out: a caller-owned
        C-contiguous float64 array of W*H*3 elements to fill instead of a new one (the reference's render_accumulator
        is sized once by reset_accumulator and reused, camera.hpp:209-233)."""
        n = cfg.image_width * cfg.image_height
        cam, env = camera_struct(cfg), env_struct(cfg)
        o = opts_struct(cfg, seed, sample_range, tile, stats, samples_per_launch, partition)
        out = Outputs()
        if no_host:
            # device-only frame (timing): the resolved, packed frame stays in the context's staging buffer on the device
            check(self.ctx.lib.zb_render(self.ctx.h, self.h, C.byref(cam), C.byref(env), C.byref(o), C.byref(out), None,
                                         C.cast(None, _lib.PROGRESS_FN), None))
            return {}
        if beauty_out is not None:
            assert beauty_out.dtype == np.float64 and beauty_out.flags.c_contiguous and beauty_out.size == 3 * n
            res = {"beauty": beauty_out.reshape(cfg.image_height, cfg.image_width, 3)}
        else:
            res = {"beauty": np.empty((cfg.image_height, cfg.image_width, 3), np.float64)}
        out.beauty = res["beauty"].ctypes.data
        for key, flag in (("albedo", cfg.use_albedo), ("normal", cfg.use_normal), ("z_depth", cfg.use_z_depth),
                          ("reflection", cfg.use_reflection), ("refraction", cfg.use_refraction)):
            if flag:
                res[key] = np.empty((cfg.image_height, cfg.image_width, 3), np.float64)
                setattr(out, key, res[key].ctypes.data)
        if want_f4:
            res["beauty_f4"] = np.empty((cfg.image_height, cfg.image_width, 4), np.float32)
            out.beauty_f4 = res["beauty_f4"].ctypes.data
        cb = _lib.PROGRESS_FN(progress) if progress else C.cast(None, _lib.PROGRESS_FN)
        cancel_p = C.byref(cancel) if cancel is not None else None
        check(self.ctx.lib.zb_render(self.ctx.h, self.h, C.byref(cam), C.byref(env), C.byref(o), C.byref(out), cancel_p,
                                     cb, None))
        assert n == res["beauty"].shape[0] * res["beauty"].shape[1]
        return res

    def refit(self, flat_with_new_transforms):
        """zb_scene_refit: take the instance table (new transforms, same geometry) of another flattened scene."""
        f = flat_with_new_transforms.contents
        check(self.ctx.lib.zb_scene_refit(self.h, f.instances, f.n_instances))

    def rebuild_tlas(self, flat_with_new_transforms):
        """zb_scene_rebuild_tlas: like refit(), and the TLAS topology is rebuilt on the device over the moved items."""
        f = flat_with_new_transforms.contents
        check(self.ctx.lib.zb_scene_rebuild_tlas(self.h, f.instances, f.n_instances))

    def render_into(self, cfg: RenderConfig, buffers: dict, seed=1, partition="auto", samples_per_launch=0):
        """zb_render into caller-owned (H, W, 3) float64 arrays keyed like render()'s result — e.g. views of one frame in
        shared memory that all ranks of a multi-process group write their bands into."""
        cam, env = camera_struct(cfg), env_struct(cfg)
        o = opts_struct(cfg, seed, None, None, False, samples_per_launch, partition)
        out = Outputs()
        n = cfg.image_width * cfg.image_height
        for key, flag in (("beauty", True), ("albedo", cfg.use_albedo), ("normal", cfg.use_normal), ("z_depth", cfg.use_z_depth),
                          ("reflection", cfg.use_reflection), ("refraction", cfg.use_refraction)):
            if flag and key in buffers:
                a = buffers[key]
                assert a.dtype == np.float64 and a.flags.c_contiguous and a.size == 3 * n, key
                setattr(out, key, a.ctypes.data)
        check(self.ctx.lib.zb_render(self.ctx.h, self.h, C.byref(cam), C.byref(env), C.byref(o), C.byref(out), None,
                                     C.cast(None, _lib.PROGRESS_FN), None))

    def render_device(self, cfg: RenderConfig, accum_ptr: int, aux_ptr: int = 0, seed=1, sample_range=None, tile=None,
                      stats=False, light_pass=False, accumulate=False, samples_per_launch=0):
        """zb_render_device: radiance SUMS of this rank's share into a device float4 buffer (async).
        light_pass: reflection sums -> accum, refraction sums -> first plane of aux (camera.hpp:492-517)."""
        cam, env = camera_struct(cfg), env_struct(cfg)
        o = opts_struct(cfg, seed, sample_range, tile, stats, samples_per_launch)
        o.light_pass = int(light_pass)
        o.accumulate = int(accumulate)
        check(self.ctx.lib.zb_render_device(self.ctx.h, self.h, C.byref(cam), C.byref(env), C.byref(o),
                                            C.c_void_p(accum_ptr), C.c_void_p(aux_ptr) if aux_ptr else None))

    def render_device_noalloc(self, cfg: RenderConfig, accum_ptr: int, seed=1, sample_range=None, tile=None, accumulate=False):
        """zb_render_device_noalloc: the beauty pass of render_device with its scratch in a static device array (no
        device allocation; frames of at most zb_noalloc_max_pixels() pixels)."""
        cam, env = camera_struct(cfg), env_struct(cfg)
        o = opts_struct(cfg, seed, sample_range, tile, False, 0)
        o.accumulate = int(accumulate)
        check(self.ctx.lib.zb_render_device_noalloc(self.ctx.h, self.h, C.byref(cam), C.byref(env), C.byref(o), C.c_void_p(accum_ptr)))

    def render_bvh_debug_device(self, cfg: RenderConfig, accum_ptr: int, level=-1, thickness=0.01, seed=1, sample_range=None):
        """zb_render_bvh_debug_device: the reference's BVH debug view (bvh.hpp:56-110), sums into a device float4 buffer."""
        cam, env = camera_struct(cfg), env_struct(cfg)
        o = opts_struct(cfg, seed, sample_range, None, False, 0)
        check(self.ctx.lib.zb_render_bvh_debug_device(self.ctx.h, self.h, C.byref(cam), C.byref(env), C.byref(o), int(level),
                                                      float(thickness), C.c_void_p(accum_ptr)))

    PROBE_DTYPE = np.dtype([("t", "f4"), ("leaf_id", "u4"), ("prim_id", "u4"), ("front_face", "u4"), ("p", "f4", 3),
                            ("normal", "f4", 3), ("u", "f4"), ("v", "f4"), ("albedo", "f4", 3), ("emitted", "f4", 3)])

    def probe(self, cfg: RenderConfig, origins, directions):
        """zb_probe_rays: per-ray hit records (structured numpy array, fields of zb_probe)."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(directions, np.float32).reshape(-1, 3)
        out = np.zeros(len(o), self.PROBE_DTYPE)
        assert out.dtype.itemsize == C.sizeof(_lib.Probe)
        env = env_struct(cfg)
        check(self.ctx.lib.zb_probe_rays(self.ctx.h, self.h, C.byref(env), len(o), o.ctypes.data, d.ctypes.data,
                                         out.ctypes.data))
        return out

    def primary_ids(self, cfg: RenderConfig):
        n = cfg.image_width * cfg.image_height
        leaf = np.empty(n, np.uint32)
        prim = np.empty(n, np.uint32)
        t = np.empty(n, np.float32)
        cam = camera_struct(cfg)
        check(self.ctx.lib.zb_primary_ids(self.ctx.h, self.h, C.byref(cam), leaf.ctypes.data, prim.ctypes.data, t.ctypes.data))
        shape = (cfg.image_height, cfg.image_width)
        return leaf.reshape(shape), prim.reshape(shape), t.reshape(shape)
