"""Host-side multi-GPU logic on CPU: partition arithmetic, and the exchange step (partition.combine) under
torch.distributed with the gloo backend, world_size 2 (the GPU path uses the same call over NCCL)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytracer_project_b200 import partition


def test_sample_slices_tile_the_range():
    for spp in (1, 7, 10, 64, 1024):
        for world in (1, 2, 3, 4, 8):
            cuts = [partition.sample_slice(r, world, spp) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == spp
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            assert max(e - b for b, e in cuts) - min(e - b for b, e in cuts) <= 1
    with pytest.raises(ValueError):
        partition.sample_slice(2, 2, 8)


def test_tile_stripes_cover_every_pixel_once():
    for (w, h) in ((1920, 1080), (300, 169), (17, 9)):
        for world in (1, 2, 3, 8):
            owners = np.array([[partition.owner_of_pixel(x, y, w, world) for x in range(0, w, 5)] for y in range(0, h, 3)])
            assert owners.min() >= 0 and owners.max() < world
            tx, ty = partition.tile_grid(w, h)
            assert sum(partition.owned_tiles(r, world, w, h) for r in range(world)) == tx * ty
    assert partition.choose(8, 64) == "samples" and partition.choose(8, 10) == "tiles"


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, mode, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w, h, spp = 48, 24, 16
    g = torch.Generator().manual_seed(1234)
    per_sample = torch.rand(spp, h * w, 4, generator=g)  # what every rank would compute for (pixel, sample)
    per_sample[..., 3] = 1.0
    if mode == "samples":
        b, e = partition.sample_slice(rank, world, spp)
        acc = per_sample[b:e].sum(0)
    else:
        ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
        own = torch.tensor([[partition.owner_of_pixel(int(x), int(y), w, world) for x in range(w)] for y in range(h)])
        acc = per_sample.sum(0) * (own.reshape(-1, 1) == rank)
    partition.combine(acc, dst=0)
    if rank == 0:
        full = per_sample.sum(0)
        out.put((mode, bool(torch.allclose(acc, full, rtol=1e-6, atol=1e-6)), bool(torch.equal(acc, full))))
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["samples", "tiles"])
def test_exchange_step_gloo_world2(mode):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got_mode, close, exact = out.get(timeout=10)
    assert got_mode == mode and close
    if mode == "tiles":
        assert exact  # disjoint pixels: the reduce is a gather, bit-identical to one rank
