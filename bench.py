#!/usr/bin/env python
"""bench.py — Mpaths/s and Mrays/s of the path-tracing hot path on 1..8 B200 of one node (BASELINE.json `metric`; default
workload = the headline: Weekend scene, 1920x1080, 64 spp, depth 50 — SURVEY.md section 8(d)).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config headline|c1|c2|c3|c4|c5] [--partition auto|samples|tiles]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...        the reference's own CPU path on the host cores (oracle/_ref/ref_harness)

A "step" = one full frame through zb_render, the call that replaces camera::render's hot part (camera.hpp:236-250 ->
:404-579).  With N ranks (one process per GPU) the ranks form a group (zb_group_export / zb_group_join: CUDA IPC handles
exchanged through torch.distributed) and zb_render is collective: every GPU renders its sample slice (or tile stripe) into
64-bit fixed-point sums, then gathers + resolves its own band of image rows from all GPUs' sums over NVLink in one kernel
(k_gather_resolve) and copies that band home over its own PCIe link.
  value   device-timed (CUDA events on the launching stream, max over ranks): scene resident in HBM, frame left on the device
  e2e     wall clock through the same call with HOST buffers: scene upload H2D + render + frame D2H into the caller's
          3-double-per-pixel buffer (one buffer in shared memory for all ranks), barriers on both sides
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEPTH = 50
# name -> (generator, kwargs, width, height, spp, spp of one reference-arm step, what BASELINE.json calls it)
CONFIGS = {
    "headline": ("weekend", dict(seed=1), 1920, 1080, 64, 4, "Weekend final scene, 1080p 64 spp (the metric's config)"),
    "c1": ("weekend", dict(seed=1), 1200, 675, 10, 4, "configs[0]: Weekend final scene 1200x675 10 spp"),
    "c2": ("cornell", dict(seed=2), 1920, 1080, 256, 2, "configs[1]: Cornell box + rotated cubes + constant_medium fog"),
    "c3": ("mesh", dict(seed=3, grid=708, knots=12, shipped=True), 1920, 1080, 64, 1,
           "configs[2]: 999 698-triangle OBJ heightfield + 12 instances of the shipped torus_knot.obj"),
    "c4": ("exterior", dict(seed=4, small=200, hdr_size=2048, shipped=True), 1920, 1080, 128, 1,
           "configs[3]: HDRI exterior, DOF 3.0 / 9.0, the shipped JPEG bump maps and wood texture"),
    "c5": ("big", dict(seed=5, grid=2237), 3840, 2160, 1024, 1, "configs[4]: C1 spheres + 10 000 712-triangle heightfield, 4K 1024 spp"),
}
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")   # the unmodified reference (kind "reference")
PORT = os.path.join(ROOT, "oracle", "_build", "zoracle")         # CPU restatement, used only if the former is absent
ISSUE_COUNTS = os.path.join(ROOT, "profiles", "r02_issue_counts.json")
SASS_HASHES = os.path.join(ROOT, "raytracer_project_b200", "_build", "sass_hashes.json")


def workload_name(name):
    gen, kw, w, h, spp, _, _ = CONFIGS[name]
    return f"{gen}_seed{kw.get('seed', 0)}_{w}x{h}_{spp}spp_depth{DEPTH}"


def metric_name(name):
    gen, kw, w, h, spp, _, _ = CONFIGS[name]
    return "Mpaths/s (1080p 64spp Weekend, depth 50)" if name == "headline" else f"Mpaths/s ({name}: {w}x{h} {spp}spp, depth {DEPTH})"


def make_scene(name, spp=None):
    from raytracer_project_b200 import scenes
    gen, kw, w, h, spp0, _, _ = CONFIGS[name]
    r = getattr(scenes, gen)(width=w, height=h, spp=spp or spp0, max_depth=DEPTH, **kw)
    return (r[0], r[1], r[2] if len(r) > 2 else {})


def cpu_arm():
    """(binary, kind) of the CPU arm: the compiled reference if it travelled with the repo, else the restatement."""
    if os.path.exists(HARNESS):
        return HARNESS, "reference"
    if not os.path.exists(PORT):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, capture_output=True)
    return PORT, "port"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


# ------------------------------------------------------------------------------- reference arm (CPU)
def run_reference_sample(name: str, spp: int):
    """Time the reference's camera::initialize + execute_render_threads on this box's host cores."""
    from raytracer_project_b200 import scenes
    binary, kind = cpu_arm()
    txt, cfg, files = make_scene(name, spp)
    d = tempfile.mkdtemp(prefix="zb_cpu_")
    try:
        os.makedirs(os.path.join(d, "assets", "hdr_maps"), exist_ok=True)  # camera::initialize side effect (camera.hpp:123-140)
        scenes.materialize(d, txt, cfg, files)
        out = subprocess.run([binary, "render", "s.zscene", "s.cfg", "-"], cwd=d, check=True, capture_output=True,
                             text=True).stdout
        r = json.loads(out.strip().splitlines()[-1])
        r["kind"] = kind
        return r
    finally:
        shutil.rmtree(d, ignore_errors=True)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    name = args.config
    _, _, w, h, _, spp, _ = CONFIGS[name]
    for _ in range(args.warmup):
        run_reference_sample(name, spp)
    secs, threads, paths, kind = [], 0, 0.0, "reference"
    for _ in range(args.steps):
        r = run_reference_sample(name, spp)
        secs.append(r["seconds"])
        threads, paths, kind = r["threads"], r["paths"], r["kind"]
    ms = 1e3 * sum(secs) / len(secs)
    value = paths / (ms * 1e-3) * 1e-6
    sample = f"{w}x{h} x {spp} spp of the same scene and camera per step (Mpaths/s is spp-independent)"
    line = {
        "impl": "reference", "metric": metric_name(name), "value": value, "unit": "Mpaths/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(name), "sample": sample},
        "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc:
            self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 <= ts <= t1 + 0.2 and len(r) >= 9] or [r for _, r in self.rows if len(r) >= 9]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = [float(r[1]) for r in rows]
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                "samples": len(rows), "power_w_max": max(float(r[3]) for r in rows)}


# ------------------------------------------------------------------------------- roofline bookkeeping
FEATURE_NAMES = {0: "PLAIN", 96: "SPHTE", 29: "BOXES", 11: "MESH0", 43: "MESH", 117: "NOMESH", 111: "NOMEDIA", 127: "ALL"}


def kernel_name(pick: int) -> str:
    """Demangled name of the k_render_pq instantiation a render reported in zb_stats::kernel_pick."""
    return f"void k_render_pq<(int){pick >> 16}, (int){(pick >> 8) & 255}, (unsigned int){pick & 255}>"


def issue_record(name: str, pick: int):
    """Warp-instruction count of one launch of the kernel on this workload, from the ncu capture recorded in
    profiles/r02_issue_counts.json — valid only while the SHA-1 of the kernel's SASS equals the one stored with it."""
    if not (os.path.exists(ISSUE_COUNTS) and os.path.exists(SASS_HASHES)):
        return None, "no recorded capture or no sass_hashes.json (run raytracer_project_b200/build.sh)"
    rec = json.load(open(ISSUE_COUNTS)).get(workload_name(name))
    if not rec:
        return None, "no ncu capture recorded for this workload"
    want = kernel_name(pick)
    built = {k: v for k, v in json.load(open(SASS_HASHES)).items() if k.startswith(want)}
    if not built or not rec["kernel"].startswith(want):
        return None, f"capture is of {rec['kernel'][:60]}, the render ran {want}"
    sha = next(iter(built.values()))["sha1"]
    if sha != rec["sass_sha1"]:
        return None, f"kernel SASS changed since the capture ({sha[:10]} != {rec['sass_sha1'][:10]}): re-profile"
    return rec, "ok"


# ------------------------------------------------------------------------------- our arm (GPU)
def ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from raytracer_project_b200 import api, partition, scenes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # NCCL prints its version banner on STDOUT, in front of the JSON line
        dist.init_process_group("nccl", device_id=dev)
    name = args.config
    _, _, WIDTH, HEIGHT, SPP, _, described = CONFIGS[name]
    n_pix = WIDTH * HEIGHT
    paths_per_step = n_pix * SPP

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def bcast(obj):
        box = [obj]
        if world > 1:
            dist.broadcast_object_list(box, src=0)
        return box[0]

    # ---- the scene: flattened once by rank 0 (host BVH build), shared with the other ranks as a ZBFS file
    job = bcast(f"{os.getpid()}_{int(time.time()) % 100000}")
    txt, cfg, files = make_scene(name)
    # ZB_BENCH_CACHE=dir keeps the flattened scene of a config as dir/NAME.zbfs for later runs on the same box (the host
    # BVH build of the 10 M-triangle config takes ~100 s, during which every GPU of the box idles)
    cache = os.environ.get("ZB_BENCH_CACHE")
    zbfs = os.path.join(cache, f"{name}{'_devblas' if args.gpu_blas else ''}.zbfs") if cache else f"/tmp/zb_bench_{name}_{job}.zbfs"
    t_prep = time.time()
    if not files:
        flat = api.flatten_zscene(txt)
    elif cache and os.path.exists(zbfs):
        barrier()
        flat = api.load_flat(zbfs)
    else:
        flat0 = None
        if rank == 0:
            d = tempfile.mkdtemp(prefix="zb_bench_")
            flat0 = api.flatten_zscene(path=scenes.materialize(d, txt, cfg, files), gpu_blas=args.gpu_blas)
            if world > 1 or cache:
                os.makedirs(os.path.dirname(zbfs), exist_ok=True)
                api.save_flat(flat0, zbfs + ".tmp")
                os.rename(zbfs + ".tmp", zbfs)
            shutil.rmtree(d, ignore_errors=True)
            os.sync()  # the scene files (up to 0.7 GB of OBJ text) and the flat-scene file: no write-back left to run under the timed steps
        barrier()
        flat = flat0 if rank == 0 else api.load_flat(zbfs)
    prep_s = time.time() - t_prep
    del files, txt
    stream = torch.cuda.current_stream()
    ctx = api.Context(local, stream=stream.cuda_stream)
    if world > 1:
        handles = [None] * world
        dist.all_gather_object(handles, ctx.group_export(n_pix))
        ctx.group_join(rank, handles, f"/zb_bench_{job}")
    scene = ctx.upload(flat)
    scene_bytes = flat.contents.nbytes()
    n_meshes = int(flat.contents.n_meshes)
    flush = torch.empty(256 * 1024 * 1024, device=dev, dtype=torch.uint8)  # > 126 MB L2
    spl = 64 if name == "c5" else 0  # progressive launches for the 8.5 G-path frame

    # one un-timed STATS pass: step counts of the build's own BVH for the algorithmic bytes / flops per ray
    s0, s1 = partition.sample_slice(rank, world, SPP)
    st = {"rays": 0, "node_visits": 0, "prim_tests": 0, "paths": 0, "instance_visits": 0}
    if s1 > s0:
        accum = torch.empty(n_pix, 4, device=dev, dtype=torch.float32)
        scene.render_device(cfg, accum.data_ptr(), seed=1, sample_range=(s0, s1), stats=True, samples_per_launch=spl)
        torch.cuda.synchronize()
        st = ctx.stats()
        del accum
    counts = torch.tensor([st["rays"], st["node_visits"], st["prim_tests"], st["paths"], st["instance_visits"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(counts)
    rays, node_visits, prim_tests, paths_counted, inst_visits = (float(x) for x in counts.tolist())

    def step_device():
        return scene.render(cfg, seed=1, no_host=True, partition=args.partition, samples_per_launch=spl)

    for _ in range(max(args.warmup, 3) if name != "c5" else 1):
        flush.zero_()
        step_device()
    barrier()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
    kern, launches, pick = [], 0, 0
    barrier()
    t_wall0 = time.time()
    for k in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (outside the per-step timed region)
        ev[k][0].record()
        step_device()
        ev[k][1].record()
        stt = ctx.stats()
        kern.append(stt["kernel_ms"])
        launches += stt["launches"]
        pick = stt["kernel_pick"]
    barrier()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    step_ms = torch.tensor([e[0].elapsed_time(e[1]) for e in ev], device=dev, dtype=torch.float64)
    kern_ms = torch.tensor(kern, device=dev, dtype=torch.float64)
    n_launch = torch.tensor([float(launches)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(step_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(kern_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(n_launch)
    ms_per_step = float(step_ms.sum().item()) / args.steps
    kernel_ms = float(kern_ms.sum().item()) / args.steps
    value = paths_per_step / (ms_per_step * 1e-3) * 1e-6

    # ---- e2e: the public call with HOST buffers (scene H2D + zb_render + frame D2H as the reference's 3 doubles per pixel);
    # with N ranks the caller's buffer is ONE frame in shared memory that every rank's band lands in
    frame_path = f"/dev/shm/zb_bench_frame_{job}"
    if rank == 0:
        with open(frame_path, "wb") as f:
            f.truncate(n_pix * 24)
    barrier()
    host_rgb = np.memmap(frame_path, dtype=np.float64, mode="r+", shape=(n_pix, 3))
    # the caller's frame buffer is reused by every step (the reference sizes render_accumulator once): with several ranks
    # on one host, register it so that every GPU DMAs its rows straight home (zb_context_register_output); one GPU keeps
    # the staged copy + host widening, which is as fast there and leaves the buffer pageable
    direct_out = world > 1 and os.environ.get("ZB_BENCH_DIRECT", "1") != "0" and ctx.register_output(host_rgb)
    e2e_ms, e2e_parts = [], []
    e2e_warm = args.warmup if name != "c5" else 0
    n_e2e = args.steps if name != "c5" else 1

    def rank_barrier():
        """barrier + device synchronise on both sides of an e2e step: the group's own host barrier (zb_group_barrier:
        stream synchronise, then a shared-memory barrier over the ranks) — every rank's GPU is idle when it returns"""
        torch.cuda.synchronize()
        ctx.group_barrier()

    for k in range(e2e_warm + n_e2e):
        barrier()
        rank_barrier()
        t0 = time.perf_counter()
        sc2 = ctx.upload(flat)  # H2D of the flattened scene
        t_up = time.perf_counter()
        sc2.render(cfg, seed=1, beauty_out=host_rgb, partition=args.partition, samples_per_launch=spl)  # blocking, collective
        t_render = time.perf_counter()
        sc2.close()
        t_close = time.perf_counter()
        rank_barrier()
        t1 = time.perf_counter()
        if os.environ.get("ZB_BENCH_TRACE"):
            print(f"[bench e2e] step {k}: upload {(t_up - t0) * 1e3:.3f} render {(t_render - t_up) * 1e3:.3f} close {(t_close - t_render) * 1e3:.3f} "
                  f"barrier {(t1 - t_close) * 1e3:.3f} ms", file=sys.stderr, flush=True)
        if k >= e2e_warm:
            e2e_ms.append((t1 - t0) * 1e3)
            e2e_parts.append(((t_up - t0) * 1e3, (t_render - t_up) * 1e3, (t1 - t_render) * 1e3))
    if direct_out:
        ctx.unregister_output(host_rgb)
    t = torch.tensor(e2e_ms, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms_per_step = float(t.sum().item()) / len(e2e_ms)
    e2e_value = paths_per_step / (e2e_ms_per_step * 1e-3) * 1e-6

    if rank == 0:
        checksum = float(host_rgb[n_pix // 2, 0])
        mean_radiance = [float(x) for x in host_rgb.mean(axis=0)]
        hbm_peak, peak_kind, sm_max = measured_peaks()
        info = ctx.device_info()
        feat = pick & 255
        # SURVEY 8(d) algorithmic work per step; counts measured by the STATS kernel on this build's own BVH.  A node
        # visit of the 64-byte two-box node = 2 ray-box tests (32 B + 25 FLOP each); sphere 16 B + 30 FLOP, triangle
        # 36 B + 50 FLOP; 96 B + 55 FLOP per instance visited; 32 B + 60 FLOP per shaded ray; 16 B per pixel accumulated.
        prim_b, prim_f = (36.0, 50.0) if (feat & 8) else (16.0, 30.0)
        alg_bytes = 32.0 * 2 * node_visits + prim_b * prim_tests + 96.0 * inst_visits + 32.0 * rays + 16.0 * n_pix * world
        alg_flop = 25.0 * 2 * node_visits + prim_f * prim_tests + 55.0 * inst_visits + 60.0 * rays
        fp32_peak = info["sm_count"] * 128 * 2 * sm_max * 1e6 * 1e-12
        issue_peak = info["sm_count"] * 4 * sm_max * 1e6 * 1e-9  # G warp-instructions / s
        fp32 = {"bound": "fp32", "achieved": alg_flop / world / (kernel_ms * 1e-3) * 1e-12, "peak": fp32_peak, "unit": "TFLOP/s",
                "what": "algorithmic FLOP (SURVEY 8(d) per-step constants x STATS-kernel step counts) / live kernel time"}
        fp32["frac"] = fp32["achieved"] / fp32_peak
        rec, why = issue_record(name, pick) if world == 1 else (None, "recorded for the 1-GPU launch only")
        hbm = {"achieved": alg_bytes / world / (kernel_ms * 1e-3) * 1e-9, "peak": hbm_peak, "unit": "GB/s",
               "peak_kind": f"{peak_kind} burst copy bandwidth", "algorithmic_bytes_per_launch": alg_bytes / world,
               "traffic": rec.get("dram_bytes") if rec else None,
               "note": "algorithmic bytes (2 box tests x 32 B per node visit) are served by L1 / L2 when the scene fits them; "
                       "`traffic` is the DRAM traffic ncu saw"}
        hbm["frac"] = hbm["achieved"] / hbm_peak
        if rec:
            roofline = {"bound": "issue", "achieved": rec["warp_inst_per_launch"] / (kernel_ms * 1e-3) * 1e-9, "peak": issue_peak,
                        "unit": "G warp-inst/s", "traffic": rec.get("dram_bytes"),
                        "lanes_per_instruction": rec.get("lanes_per_instruction"), "kernel": rec["kernel"],
                        "source": f"smsp__inst_executed.sum of one launch on this workload ({rec['source']}), valid for SASS "
                                  f"sha1 {rec['sass_sha1'][:12]} = the kernel that ran; / live kernel time; peak = SMs x 4 schedulers x {sm_max:.0f} MHz"}
            roofline["frac"] = roofline["achieved"] / issue_peak
        else:
            roofline = dict(fp32, traffic=None, note=f"issue-slot roofline unavailable: {why}")
        line = {
            "metric": metric_name(name), "value": value, "unit": "Mpaths/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3) if name != "c5" else 1, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_per_s": value * rays / max(paths_counted, 1.0),
            "config": {"workload": workload_name(name), "described": described,
                       "partition": (args.partition if world > 1 else "one GPU") + f" x{world}",
                       "exchange": "k_gather_resolve: per-GPU row bands, integer sums of peer-mapped partials over NVLink" if world > 1 else "none",
                       "l2_flush_between_steps": True, "scene_bytes": scene_bytes, "scene_prep_s": round(prep_s, 2),
                       "rays_per_path": rays / max(paths_counted, 1.0), "node_visits_per_ray": node_visits / max(rays, 1.0),
                       "prim_tests_per_ray": prim_tests / max(rays, 1.0), "kernel": kernel_name(pick),
                       "feature_set": FEATURE_NAMES.get(feat, str(feat)),
                       "blas": "no meshes" if not n_meshes else ("device binned SAH (sah_build.cu)" if os.environ.get("ZB_GPU_BUILDER", "1") != "0" else "device LBVH (lbvh.cu)") if args.gpu_blas else "host binned SAH"},
            "kernel_ms_per_step": kernel_ms,
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "ms_per_step": e2e_ms_per_step,
                    "h2d_bytes_per_step": (scene_bytes + 512) * world, "d2h_bytes_per_step": n_pix * (24 if direct_out else 12), "checksum": checksum,
                    "steps": len(e2e_ms),
                    "parts_ms": {"scene_upload": sum(p[0] for p in e2e_parts) / len(e2e_parts),
                                 "render_gather_d2h": sum(p[1] for p in e2e_parts) / len(e2e_parts),
                                 "scene_free_barrier": sum(p[2] for p in e2e_parts) / len(e2e_parts)},
                    "output_registered": bool(direct_out),
                    "note": ("every GPU copies its own band of rows home over its own PCIe link, as 3 doubles per pixel in one DMA into "
                             "the caller's registered (page-locked, shared) frame" if direct_out else
                             "every GPU copies its own band of rows home as packed RGB float means (12 B/pixel) over its own "
                             "PCIe link; host threads of the same rank widen it to the reference's 3-double layout in the "
                             "caller's (shared) frame while later chunks are in flight")},
            "gpu_launches": int(n_launch.item()),
            "roofline": roofline,
            "roofline_fp32": fp32,
            "roofline_hbm": hbm,
            "clocks": clocks,
            "mean_radiance": mean_radiance,
            "wall_ms_per_step_incl_flush": (t_wall1 - t_wall0) * 1e3 / args.steps,
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                spp_cpu = CONFIGS[name][5] * 2 if name in ("headline", "c1") else CONFIGS[name][5]
                r = run_reference_sample(name, spp_cpu)
                line["cpu_baseline"] = {
                    "value": r["mpaths_per_s"], "unit": "Mpaths/s", "cores": r["threads"], "kind": r["kind"],
                    "sample": f"{WIDTH}x{HEIGHT} x {spp_cpu} spp of the same scene and camera, {r['seconds']:.1f} s "
                              "(reference renders with std::thread x hardware_concurrency, camera.hpp:426)"}
            except Exception as e:  # the harness is prebuilt in the build container; report, don't hide
                line["cpu_baseline"] = {"value": None, "unit": "Mpaths/s", "cores": 0, "kind": "reference",
                                        "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
    del host_rgb
    barrier()
    if rank == 0:
        for p in (frame_path,) + (() if cache else (zbfs,)):
            if os.path.exists(p):
                os.remove(p)
    scene.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="headline", choices=list(CONFIGS))
    ap.add_argument("--partition", default="auto", choices=["auto", "samples", "tiles"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gpu-blas", dest="gpu_blas", action="store_true", default=True,
                    help="mesh BLASes built on the device at upload (binned SAH, csrc/sah_build.cu; ZB_GPU_BUILDER=0: linear BVH, csrc/lbvh.cu): "
                         "the default, as in zenith::camera")
    ap.add_argument("--host-blas", dest="gpu_blas", action="store_false", help="mesh BLASes built by the host flattener (host/bvh_build.hpp)")
    ap.add_argument("--spp", type=int, default=0, help="profiling runs only: override the config's samples per pixel (the workload name says so)")
    args = ap.parse_args()
    if args.spp > 0:
        c = CONFIGS[args.config]
        CONFIGS[args.config] = c[:4] + (args.spp,) + c[5:]
    if args.impl == "reference":
        return reference_arm(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return ours(args)


if __name__ == "__main__":
    sys.exit(main())
