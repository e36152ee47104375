"""Multi-GPU frames behind the C ABI (csrc/group.cu).  The exchange adds 64-bit integer sums, so a frame
rendered by n ranks must equal the one-GPU frame BIT FOR BIT, for sample slices and for tile stripes, for every output
buffer.  A device may be listed more than once in zb_context_create_multi and several processes may share a GPU, so all of
this runs on a box with a single GPU; with >= 2 GPUs the same tests also run across real peers (NVLink)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from raytracer_project_b200 import scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ("beauty", "albedo", "normal", "z_depth", "reflection", "refraction")


def _n_devices():
    import torch
    return torch.cuda.device_count()


def _device_lists():
    lists = [[0, 0], [0, 0, 0]]
    n = _n_devices()
    if n >= 2:
        lists.append([0, 1])
    if n >= 4:
        lists.append([0, 1, 2, 3])
    if n >= 8:
        lists.append(list(range(8)))
    return lists


@pytest.fixture(scope="module")
def weekend_flat():
    from raytracer_project_b200 import api
    txt, _ = scenes.weekend(seed=1)
    flat = api.flatten_zscene(txt)
    yield flat
    api.free_flat(flat)


@pytest.mark.parametrize("partition", ["samples", "tiles", "auto"])
def test_multi_context_frame_equals_one_gpu_frame(gpu_ctx, weekend_flat, partition):
    from raytracer_project_b200 import api
    cfg = scenes.weekend(seed=1, width=203, height=91, spp=24)[1]  # ragged tile edge, rows not divisible by 2 or 3
    cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = cfg.use_reflection = cfg.use_refraction = True
    one = gpu_ctx.upload(weekend_flat)
    ref = one.render(cfg, seed=7)
    one.close()
    for devs in _device_lists():
        with api.Context(devs) as ctx:
            assert ctx.device_count() == len(devs)
            sc = ctx.upload(weekend_flat)
            a = sc.render(cfg, seed=7, partition=partition)
            b = sc.render(cfg, seed=7, partition=partition, samples_per_launch=5)  # sample batches add into the same integers
            st = ctx.stats()
            sc.close()
        assert st["n_gpus"] == len(devs) and st["stack_overflows"] == 0
        assert st["paths"] == 203 * 91 * 24, st
        for k in KEYS:
            assert np.array_equal(a[k], ref[k]), (devs, partition, k, np.abs(a[k] - ref[k]).max())
            assert np.array_equal(b[k], ref[k]), (devs, partition, k, "batched")


def test_multi_context_scene_edits_reach_every_replica(gpu_ctx, tmp_path):
    """zb_scene_refit and zb_scene_rebuild_tlas on a multi-GPU context edit the scene's replica on every member; with the
    mesh BLASes built on each device at upload.  The frames equal a one-GPU render of the freshly flattened moved scene."""
    from raytracer_project_b200 import api
    flats = {}
    for m in (0.0, 1.0):
        txt, cfg, files = scenes.refit_demo(move=m)
        flats[m] = api.flatten_zscene(path=scenes.materialize(str(tmp_path / f"m{int(m)}"), txt, cfg, files), gpu_blas=True)
    fresh = gpu_ctx.upload(flats[1.0])
    ref = fresh.render(cfg, seed=5)["beauty"]
    fresh.close()
    for devs in _device_lists()[:3]:
        with api.Context(devs) as ctx:
            ctx.set_tuning("build_check", 1)
            sc = ctx.upload(flats[0.0])
            sc.refit(flats[1.0])
            a = sc.render(cfg, seed=5)["beauty"]
            sc.refit(flats[0.0])
            sc.rebuild_tlas(flats[1.0])
            b = sc.render(cfg, seed=5)["beauty"]
            st = ctx.stats()
            sc.close()
        assert st["n_gpus"] == len(devs) and st["stack_overflows"] == 0
        assert np.array_equal(a, ref), (devs, "refit", np.abs(a - ref).max())
        differ = (np.abs(b - ref).max(axis=2) > 0).mean()
        assert differ < 1e-3, (devs, "rebuild", differ)  # items re-ordered: exact ties between two items may resolve the other way
    for f in flats.values():
        api.free_flat(f)


def test_registered_output_buffers_get_the_same_frame(gpu_ctx, weekend_flat):
    """zb_context_register_output: finished rows are DMA'd into the caller's page-locked buffers as doubles instead of being
    staged as floats and widened on the host — all six output buffers equal the staged path's bit for bit, on one GPU and
    on multi-device contexts; a buffer outside the registered range falls back to the staged path."""
    from raytracer_project_b200 import api
    cfg = scenes.weekend(seed=1, width=203, height=91, spp=16)[1]
    cfg.use_albedo = cfg.use_normal = cfg.use_z_depth = cfg.use_reflection = cfg.use_refraction = True
    one = gpu_ctx.upload(weekend_flat)
    ref = one.render(cfg, seed=3)
    one.close()
    n = 203 * 91
    for devs in [[0]] + _device_lists()[:3]:
        with api.Context(devs if len(devs) > 1 else devs[0]) as ctx:
            sc = ctx.upload(weekend_flat)
            pool = np.full((6, n, 3), -1.0)  # one allocation holding the six planes: one registration
            assert ctx.register_output(pool)
            bufs = {k: pool[i].reshape(91, 203, 3) for i, k in enumerate(KEYS)}
            sc.render_into(cfg, bufs, seed=3)
            for k in KEYS:
                assert np.array_equal(bufs[k], ref[k]), (devs, k, "direct")
            # a frame whose beauty buffer is NOT registered: the staged path, same values
            other = np.full((n, 3), -1.0)
            bufs2 = dict(bufs, beauty=other.reshape(91, 203, 3))
            sc.render_into(cfg, bufs2, seed=3)
            for k in KEYS:
                assert np.array_equal(bufs2[k], ref[k]), (devs, k, "mixed -> staged")
            ctx.unregister_output(pool)
            sc.render_into(cfg, bufs, seed=3)
            assert np.array_equal(bufs["beauty"], ref["beauty"])
            sc.close()


def test_multi_context_low_spp_takes_tile_stripes(gpu_ctx, weekend_flat):
    """spp < 8 n (config C1's 10 spp on 2+ GPUs): AUTO picks tile stripes; the frame still equals the one-GPU frame."""
    from raytracer_project_b200 import api
    cfg = scenes.weekend(seed=1, width=320, height=180, spp=10)[1]
    one = gpu_ctx.upload(weekend_flat)
    ref = one.render(cfg, seed=3)["beauty"]
    one.close()
    with api.Context([0, 0]) as ctx:
        sc = ctx.upload(weekend_flat)
        a = sc.render(cfg, seed=3)["beauty"]
        sc.close()
    assert np.array_equal(a, ref)


def test_multi_context_rejects_single_gpu_only_options(weekend_flat):
    from raytracer_project_b200 import api
    cfg = scenes.weekend(seed=1, width=64, height=32, spp=8)[1]
    with api.Context([0, 0]) as ctx:
        sc = ctx.upload(weekend_flat)
        with pytest.raises(api._lib.ZbError):
            sc.render(cfg, seed=1, stats=True)
        with pytest.raises(api._lib.ZbError):
            sc.render(cfg, seed=1, tile=(0, 2))
        sc.close()


@pytest.mark.parametrize("partition", ["samples", "tiles"])
def test_process_group_frame_equals_one_gpu_frame(tmp_path, partition):
    """One PROCESS per rank (the torchrun shape): CUDA IPC buffers and events, host barrier in shared memory, one frame in
    shared memory that all ranks write their bands into."""
    n = 2
    devices = [0, 1] if _n_devices() >= 2 else [0, 0]
    d = str(tmp_path)
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "group_worker.py"), str(r), str(n), str(devices[r]), d, partition],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(n)]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    res = json.load(open(os.path.join(d, "result.json")))
    assert res["mean"] > 0.1
    for k in KEYS:
        assert res[k], (k, res)
