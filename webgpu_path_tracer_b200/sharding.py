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


def reduce_framebuffer(fb_tensor, dist, dst=0):
    """Sum-reduce binding 3 of every rank onto `dst`.  Tile shards have disjoint support (pixels a
    rank does not own stay 0, and x + 0 == x), so the result is bit-identical to one GPU; frame
    strata change the per-pixel summation order, so they agree to f32 rounding only."""
    dist.reduce(fb_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return fb_tensor
