#!/usr/bin/env python
"""Finishing chain timing (GPU box): zb_analyze_device / zb_finish_device on a device-resident float4 frame, CUDA
events on the context stream, L2 flushed between runs.  Prints one JSON line per case with the achieved HBM GB/s
against algorithmic bytes (finish.cu header) and MEASURED_PEAKS.json."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from raytracer_project_b200 import api

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=3840)
ap.add_argument("--height", type=int, default=2160)
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--profile", action="store_true", help="one launch per case, no timing (for ncu)")
a = ap.parse_args()
peak = 6650.0
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    peak = float(json.load(open(p))["hbm_gbs"])
stream = torch.cuda.current_stream()
ctx = api.Context(0, stream=stream.cuda_stream)
n = a.width * a.height
src = torch.rand(n, 4, device="cuda") * 3.0
out_f4 = torch.empty(n, 4, device="cuda")
out_u8 = torch.empty(n, 4, device="cuda", dtype=torch.uint8)
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)


def timed(fn):
    ms = []
    for i in range(a.reps + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ms.append(e0.elapsed_time(e1))
    return min(ms), sum(ms) / len(ms)


cases = {
    # name: (callable, algorithmic bytes per pixel)
    "finish_plain_f4": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(), 0, out_f4.data_ptr(), 0), 32),
    "finish_plain_rgba8": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(), 0, 0, out_u8.data_ptr()), 20),
    "finish_sharpen_aces_hsv_rgba8": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(use_sharpening=1, use_aces=1, saturation=1.2, hue_shift=10.0), 0, 0, out_u8.data_ptr()), 20),
    "finish_bloom_sharpen_aces_rgba8 (k_finish_fused, 20 B/px)": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(use_bloom=1, use_sharpening=1, use_aces=1), 0, 0, out_u8.data_ptr()), 20),
    "finish_bloom_sharpen_aces_f4 (k_finish_fused, 32 B/px)": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(use_bloom=1, use_sharpening=1, use_aces=1), 0, out_f4.data_ptr(), 0), 32),
    "finish_bloom_sharpen_aces_rgba8 (3 kernels, 100 B/px)": (lambda: (ctx.set_tuning("fuse_finish", 0), ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(use_bloom=1, use_sharpening=1, use_aces=1), 0, 0, out_u8.data_ptr()), ctx.set_tuning("fuse_finish", 1)), 32 + 32 + 36),
    "finish_bloom_r8_rgba8 (radius 8 > fused limit: 3 kernels)": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(use_bloom=1, bloom_radius=8), 0, 0, out_u8.data_ptr()), 32 + 32 + 36),
    "aux_pass_rgba8": (lambda: ctx.finish_device(src.data_ptr(), a.width, a.height, api.post_struct(), 2, 0, out_u8.data_ptr()), 20),
}
if a.profile:
    for name, (fn, bpp) in cases.items():
        fn()
    ctx.analyze_device(src.data_ptr(), n)
    torch.cuda.synchronize()
    sys.exit(0)
for name, (fn, bpp) in cases.items():
    best, mean = timed(fn)
    gbs = n * bpp / (best * 1e-3) * 1e-9
    print(json.dumps({"case": name, "width": a.width, "height": a.height, "ms_best": round(best, 4), "ms_mean": round(mean, 4),
                      "algorithmic_bytes": n * bpp, "achieved_gbs": round(gbs, 1), "peak_gbs": peak, "frac": round(gbs / peak, 3)}), flush=True)
# analyze: includes its own small D2H + sync, timed by wall clock around the blocking call
import time
ctx.analyze_device(src.data_ptr(), n)
t = []
for _ in range(a.reps):
    flush.zero_(); torch.cuda.synchronize()
    t0 = time.perf_counter(); ctx.analyze_device(src.data_ptr(), n); t.append((time.perf_counter() - t0) * 1e3)
print(json.dumps({"case": "analyze (k_luma_stats + 1 KB D2H, blocking call)", "ms_best": round(min(t), 4), "algorithmic_bytes": n * 16,
                  "achieved_gbs": round(n * 16 / (min(t) * 1e-3) * 1e-9, 1), "peak_gbs": peak}), flush=True)
