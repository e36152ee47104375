"""Multi-GPU partition of one frame and the exchange step that follows it.

Every (pixel, sample) path is independent and the scene is replicated, so the only communication is the
combination of the per-rank float4 radiance SUMS at the end of the frame:

  sample slices  rank g renders samples [g*spp/N, (g+1)*spp/N) of the whole frame; the partial sums are
                 added with one reduce (NCCL over NVLink on GPUs; gloo in the CPU tests).
  tile stripes   tile t (16x8 pixels, row-major) belongs to rank t mod N; every rank's buffer is zero outside
                 its tiles, so the same sum-reduce is a gather and the union is bit-identical to one GPU.

The reference has one strategy only: static contiguous row blocks over std::thread (camera.hpp:557-566).
"""
from __future__ import annotations

TILE_W, TILE_H = 16, 8  # ZB_TILE_W / ZB_TILE_H of csrc/scene_dev.cuh


def sample_slice(rank: int, world: int, spp: int) -> tuple[int, int]:
    """[begin, end) of the global sample indices rendered by `rank`; slices tile [0, spp) exactly."""
    if not (0 <= rank < world) or spp < 0:
        raise ValueError("bad rank / world / spp")
    return rank * spp // world, (rank + 1) * spp // world


def choose(world: int, spp: int) -> str:
    """Sample slices when every rank gets at least 8 samples, else tile stripes."""
    return "samples" if spp >= 8 * world else "tiles"


def tile_grid(width: int, height: int) -> tuple[int, int]:
    return (width + TILE_W - 1) // TILE_W, (height + TILE_H - 1) // TILE_H


def owned_tiles(rank: int, world: int, width: int, height: int) -> int:
    tx, ty = tile_grid(width, height)
    total = tx * ty
    return (total - rank + world - 1) // world


def owner_of_pixel(x: int, y: int, width: int, world: int) -> int:
    tx = (width + TILE_W - 1) // TILE_W
    return ((y // TILE_H) * tx + x // TILE_W) % world


def combine(accum, dst: int = 0):
    """Exchange step: sum the per-rank float4 sum buffers onto rank `dst` (in place).  `accum` is a torch
    tensor on the device of the process group's backend."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum
