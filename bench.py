#!/usr/bin/env python
"""bench.py — headline benchmark: Mpaths/s (and Mrays/s) on the 1080p / 64 spp / depth-50 Weekend scene
(BASELINE.json `metric`), 1..8 B200 of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" = one full frame: every rank renders its sample slice of the whole frame into a float4 sum
buffer (k_render_pq: frame-wide path queue, fixed-point sums, + k_pq_resolve), the partial sums are combined on rank 0 with one NCCL reduce over NVLink, and rank 0
resolves sums -> means (k_resolve).  `value` is timed with inputs (the flattened scene) resident in
HBM; `e2e` goes through the public C-ABI call zb_render with host buffers (scene upload H2D and the
double-RGB frame D2H inside the timed region).  Prints ONE JSON line on rank 0.

--impl reference times the UNMODIFIED reference's CPU path (oracle/_ref/ref_harness = the reference's
own headers compiled by oracle/Makefile) on the host cores, on a bounded sample of the same workload.
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

WIDTH, HEIGHT, SPP, DEPTH = 1920, 1080, 64, 50
WORKLOAD = f"weekend_seed1_{WIDTH}x{HEIGHT}_{SPP}spp_depth{DEPTH}"
# DRAM traffic of one k_render_pq launch on the headline frame (ncu, profiles/r01_k_render_pq_plain_final_raw_key.txt): 50.3 MB read +
# 5.1 MB written while the kernel runs — the 60 KB scene is L1-resident; the traffic is the 50 MB of 64-bit fixed-point
# per-pixel sums (zeroed before the launch, RED.ADDed in L2) entering and partly leaving the 126 MB L2.
NCU_DRAM_BYTES_PER_LAUNCH = 50264576 + 5082368
# warp instructions of the same launch (smsp__inst_executed.sum, same capture); the kernel is bound by instruction issue
# (4 schedulers per SM, one warp instruction per cycle each), so this is the roofline that explains `value`
NCU_WARP_INST_PER_LAUNCH = 23784447142
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")   # the unmodified reference (kind "reference")
PORT = os.path.join(ROOT, "oracle", "_build", "zoracle")         # CPU restatement, used only if the former is absent


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
def run_reference_sample(spp: int):
    """Time the reference's camera::initialize + execute_render_threads on this box's host cores."""
    from raytracer_project_b200 import scenes
    binary, kind = cpu_arm()
    txt, cfg = scenes.weekend(seed=1, width=WIDTH, height=HEIGHT, spp=spp, max_depth=DEPTH)
    d = tempfile.mkdtemp(prefix="zb_cpu_")
    try:
        os.makedirs(os.path.join(d, "assets", "hdr_maps"))  # camera::initialize side effect (camera.hpp:123-140)
        open(os.path.join(d, "s.zscene"), "w").write(txt)
        open(os.path.join(d, "s.cfg"), "w").write(cfg.to_text())
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
    spp = 4
    for _ in range(args.warmup):
        run_reference_sample(spp)
    secs, threads, paths, kind = [], 0, 0.0, "reference"
    for _ in range(args.steps):
        r = run_reference_sample(spp)
        secs.append(r["seconds"])
        threads, paths, kind = r["threads"], r["paths"], r["kind"]
    ms = 1e3 * sum(secs) / len(secs)
    value = paths / (ms * 1e-3) * 1e-6
    sample = f"{WIDTH}x{HEIGHT} x {spp} spp of the same scene and camera per step (Mpaths/s is spp-independent)"
    line = {
        "impl": "reference", "metric": "Mpaths/s (1080p 64spp Weekend, depth 50)", "value": value, "unit": "Mpaths/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
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


# ------------------------------------------------------------------------------- our arm (GPU)
def ours(args):
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
        dist.init_process_group("nccl", device_id=dev)

    txt, cfg = scenes.weekend(seed=1, width=WIDTH, height=HEIGHT, spp=SPP, max_depth=DEPTH)
    flat = api.flatten_zscene(txt)
    stream = torch.cuda.current_stream()
    ctx = api.Context(local, stream=stream.cuda_stream)
    scene = ctx.upload(flat)
    scene_bytes = flat.contents.nbytes()
    n_pix = WIDTH * HEIGHT
    paths_per_step = n_pix * SPP
    # sample slices: rank g renders samples [g*spp/n, (g+1)*spp/n) of the whole frame
    s0, s1 = partition.sample_slice(rank, world, SPP)
    accum = torch.empty(n_pix, 4, device=dev, dtype=torch.float32)
    out_f4 = torch.empty(n_pix, 4, device=dev, dtype=torch.float32)
    flush = torch.empty(256 * 1024 * 1024, device=dev, dtype=torch.uint8)  # > 126 MB L2

    def step_device():
        scene.render_device(cfg, accum.data_ptr(), seed=1, sample_range=(s0, s1))
        partition.combine(accum, dst=0)
        if rank == 0:
            api.check(ctx.lib.zb_resolve_device(ctx.h, accum.data_ptr(), n_pix, 1.0 / SPP, out_f4.data_ptr(), None))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # one un-timed STATS pass: step counts of the build's own BVH for the algorithmic bytes/flops per ray
    scene.render_device(cfg, accum.data_ptr(), seed=1, sample_range=(s0, s1), stats=True)
    torch.cuda.synchronize()
    st = ctx.stats()
    counts = torch.tensor([st["rays"], st["node_visits"], st["prim_tests"], st["paths"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(counts)
    rays, node_visits, prim_tests, paths_counted = (float(x) for x in counts.tolist())

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step_device()
    barrier()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    barrier()
    t_wall0 = time.time()
    for k in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (not inside the per-step timed region)
        ev[k][0].record()
        scene.render_device(cfg, accum.data_ptr(), seed=1, sample_range=(s0, s1))
        ev[k][1].record()
        partition.combine(accum, dst=0)
        if rank == 0:
            api.check(ctx.lib.zb_resolve_device(ctx.h, accum.data_ptr(), n_pix, 1.0 / SPP, out_f4.data_ptr(), None))
        ev[k][2].record()
    barrier()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    step_ms = torch.tensor([e[0].elapsed_time(e[2]) for e in ev], device=dev, dtype=torch.float64)
    kern_ms = torch.tensor([e[0].elapsed_time(e[1]) for e in ev], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(step_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(kern_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(step_ms.sum().item()) / args.steps
    kernel_ms = float(kern_ms.sum().item()) / args.steps
    value = paths_per_step / (ms_per_step * 1e-3) * 1e-6
    mean_gpu = out_f4[:, :3].mean(dim=0).tolist() if rank == 0 else None

    # ---- e2e: public C-ABI call with host buffers (scene H2D + zb_render + double-RGB frame D2H)
    import numpy as np
    e2e_ms = []
    host_rgb = np.empty((n_pix, 3), np.float64) if rank == 0 else None  # the caller's std::vector<color>
    e2e_parts = []
    for k in range(args.warmup + args.steps):
        barrier()
        t0 = time.perf_counter()
        sc2 = ctx.upload(flat)  # H2D of the flattened scene
        t_up = time.perf_counter()
        if world == 1:
            res = sc2.render(cfg, seed=1, beauty_out=host_rgb)  # blocking; fills the caller's (H, W, 3) float64 buffer
            checksum = float(res["beauty"][HEIGHT // 2, WIDTH // 2, 0])
        else:
            sc2.render_device(cfg, accum.data_ptr(), seed=1, sample_range=(s0, s1))
            partition.combine(accum, dst=0)
            if rank == 0:
                api.check(ctx.lib.zb_resolve_device(ctx.h, accum.data_ptr(), n_pix, 1.0 / SPP, out_f4.data_ptr(), None))
                api.check(ctx.lib.zb_read_frame(ctx.h, out_f4.data_ptr(), n_pix, host_rgb.ctypes.data))
            torch.cuda.synchronize()
            checksum = float(host_rgb[n_pix // 2, 0]) if rank == 0 else 0.0
        t_render = time.perf_counter()
        sc2.close()
        barrier()
        if k >= args.warmup:
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
            e2e_parts.append(((t_up - t0) * 1e3, (t_render - t_up) * 1e3, (time.perf_counter() - t_render) * 1e3))
    t = torch.tensor(e2e_ms, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms_per_step = float(t.sum().item()) / args.steps
    e2e_value = paths_per_step / (e2e_ms_per_step * 1e-3) * 1e-6

    if rank == 0:
        hbm_peak, peak_kind, sm_max = measured_peaks()
        # SURVEY 8(d) algorithmic work per step: 32 B + 25 FLOP per ray-box test (2 per node visit), 16 B + 30 FLOP per
        # sphere test, 32 B + 60 FLOP per shaded ray, 16 B per pixel accumulate; counts measured by the STATS kernel.
        alg_bytes = 32.0 * 2 * node_visits + 16.0 * prim_tests + 32.0 * rays + 16.0 * n_pix * world
        alg_flop = 25.0 * 2 * node_visits + 30.0 * prim_tests + 60.0 * rays
        per_gpu_bytes = alg_bytes / world
        achieved = per_gpu_bytes / (kernel_ms * 1e-3) * 1e-9
        info = ctx.device_info()
        fp32_peak = info["sm_count"] * 128 * 2 * sm_max * 1e6 * 1e-12
        line = {
            "metric": "Mpaths/s (1080p 64spp Weekend, depth 50)", "value": value, "unit": "Mpaths/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "partition": f"sample_slices x{world}", "l2_flush_between_steps": True,
                       "scene_bytes": scene_bytes, "rays_per_path": rays / paths_counted,
                       "node_visits_per_ray": node_visits / rays, "prim_tests_per_ray": prim_tests / rays},
            "mrays_per_s": value * rays / paths_counted,
            "kernel_ms_per_step": kernel_ms,
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "ms_per_step": e2e_ms_per_step,
                    "h2d_bytes_per_step": scene_bytes + 512,
                    "d2h_bytes_per_step": n_pix * 12, "checksum": checksum,
                    "parts_ms": {"scene_upload": sum(p[0] for p in e2e_parts) / len(e2e_parts),
                                 "render_resolve_d2h": sum(p[1] for p in e2e_parts) / len(e2e_parts),
                                 "scene_free": sum(p[2] for p in e2e_parts) / len(e2e_parts)},
                    "note": "frame crosses PCIe as packed RGB float means (12 B/pixel) and is widened to the reference's "
                            "3-double layout by host threads while later chunks are in flight"},
            "gpu_launches": 3 * args.steps,  # k_render_pq + k_pq_resolve (zb_render_device) + k_resolve per step (device-timed leg)
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH if world == 1 else None,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of k_render_pq on this workload, "
                                           "ncu --set full capture summarised in profiles/r01_k_render_pq_plain_final_raw_key.txt",
                         "peak_kind": f"{peak_kind} burst copy bandwidth",
                         "kernel": "k_render_pq<8, PQ_BEAUTY, PLAIN>", "algorithmic_bytes_per_launch": per_gpu_bytes,
                         "note": "scene (60 KB) is L1/L2-resident: algorithmic bytes are served by L1, so this "
                                 "fraction can exceed 1; the binding roofline is instruction issue (roofline_issue, roofline_fp32)"},
            "roofline_fp32": {"achieved": alg_flop / world / (kernel_ms * 1e-3) * 1e-12, "peak": fp32_peak,
                              "unit": "TFLOP/s", "frac": alg_flop / world / (kernel_ms * 1e-3) * 1e-12 / fp32_peak},
            "roofline_issue": ({"achieved": NCU_WARP_INST_PER_LAUNCH / (kernel_ms * 1e-3) * 1e-9,
                                "peak": info["sm_count"] * 4 * sm_max * 1e6 * 1e-9, "unit": "G warp-inst/s",
                                "frac": NCU_WARP_INST_PER_LAUNCH / (kernel_ms * 1e-3) / (info["sm_count"] * 4 * sm_max * 1e6),
                                "lanes_per_instruction": 15.13,
                                "source": "smsp__inst_executed.sum of one k_render_pq launch on this workload "
                                          "(profiles/r01_k_render_pq_plain_final_raw_key.txt) / live kernel time"}
                               if world == 1 else None),
            "clocks": clocks,
            "mean_radiance": mean_gpu,
            "wall_ms_per_step_incl_flush": (t_wall1 - t_wall0) * 1e3 / args.steps,
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                spp_cpu = 8
                r = run_reference_sample(spp_cpu)
                line["cpu_baseline"] = {
                    "value": r["mpaths_per_s"], "unit": "Mpaths/s", "cores": r["threads"], "kind": r["kind"],
                    "sample": f"{WIDTH}x{HEIGHT} x {spp_cpu} spp of the same scene and camera, {r['seconds']:.1f} s "
                              "(reference renders with std::thread x hardware_concurrency, camera.hpp:426)"}
            except Exception as e:  # the harness is prebuilt in the build container; report, don't hide
                line["cpu_baseline"] = {"value": None, "unit": "Mpaths/s", "cores": 0, "kind": "reference",
                                        "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
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
