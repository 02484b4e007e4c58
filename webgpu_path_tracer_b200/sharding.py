"""Multi-GPU partitioning of the hot path: one process per GPU, scene replicated,
work split by image tiles and/or frame strata, accumulation buffers combined by ONE sum-reduce
(NCCL over NVLink on GPUs; any torch.distributed backend works, gloo in the CPU tests).

The arithmetic here mirrors csrc/scene.cuh shard_pixel() and fast.cu's frame selection exactly, so
the host can reason about which rank owns what without touching the device.
"""
import numpy as np


def tile_rows_of_rank(height, rank, world, tile_rows):
    """Image rows rendered by `rank`: row-blocks of `tile_rows` rows, block b -> rank b % world."""
    rows = np.arange(height)
    return rows[(rows // tile_rows) % world == rank]


def tile_mask(width, height, rank, world, tile_rows):
    """Boolean (H*W,) mask of the pixels `rank` renders."""
    m = np.zeros((height, width), bool)
    m[tile_rows_of_rank(height, rank, world, tile_rows)] = True
    return m.reshape(-1)


def shard_pixel_count(width, height, rank, world, tile_rows):
    """shard.n_pixels as wgpt_create computes it (the last block may run past H: masked on the device)."""
    if world == 1:
        return width * height
    nblocks = (height + tile_rows - 1) // tile_rows
    mine = (nblocks - 1 - rank) // world + 1 if nblocks > rank else 0
    return mine * tile_rows * width


def strata_frames(first_frame, n_frames, rank, world):
    """Frame numbers `rank` renders out of [first_frame, first_frame + n_frames)."""
    return [f for f in range(first_frame, first_frame + n_frames) if f % world == rank]


def shard_of_rank(rank, world, tile_groups):
    """(tile_rank, tile_world, strata_rank, strata_world) of `rank` in a tile x strata grid -- the assignment
    wgpt_group_create uses: rank r renders tile shard r % tile_groups of strata shard r // tile_groups."""
    if world % tile_groups:
        raise ValueError("world must be a multiple of tile_groups")
    return rank % tile_groups, tile_groups, rank // tile_groups, world // tile_groups


def default_tile_groups(world):
    """Tile groups bench.py uses at `world` GPUs: 1, 1, 2, 4 for 1, 2, 4, 8 (north_star: C4 = 4 tile groups x 2 strata)."""
    return 4 if world >= 8 and world % 4 == 0 else (2 if world >= 4 and world % 2 == 0 else 1)


def init_comm(pt, dist, rank, world):
    """One process per GPU: rank 0 makes the NCCL id, `dist` (any initialised torch.distributed backend: the plumbing)
    broadcasts its 128 bytes, every rank joins with its context.  The reduce itself is the library's (wgpt_reduce_framebuffer)."""
    from .api import nccl_unique_id
    box = [nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    pt.comm_init(box[0], rank, world)


def reduce_framebuffer(fb_tensor, dist, dst=0):
    """CPU stand-in of wgpt_reduce_framebuffer for the gloo tests (no NCCL without GPUs): the same sum-reduce of binding 3
    through torch.distributed.  Tile shards have disjoint support (pixels a rank does not own stay 0, and x + 0 == x),
    so the result is bit-identical to one GPU; frame strata change the per-pixel summation order (f32 rounding)."""
    dist.reduce(fb_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return fb_tensor
