"""World-size-2 test of the N>1 host path on CPU (gloo): shard arithmetic + one sum-reduce.
The renderer behind each rank is the CPU oracle here (no GPU in this container); the GPU
version of the same check is tests/test_gpu_multi.py."""
import os
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from webgpu_path_tracer_b200 import sharding
from tests.util import bits

W, H, FRAMES = 40, 28, 4


def _render(frames, mask=None):
    from webgpu_path_tracer_b200 import Scene, Camera
    from oracle.binding import Oracle
    s = Scene()
    s.init_mesh_data()
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
    o = Oracle(W, H, threads=2)
    o.write_scene(s.buffers())
    for f in frames:
        o.write_uniforms(f, 0, cam.viewMatrix)
        o.dispatch()
    fb = o.read_framebuffer()
    if mask is not None:
        fb[~mask] = 0.0
    return fb


def _worker(rank, world, port, split, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    if split == "tile":
        fb = _render(range(1, FRAMES + 1), sharding.tile_mask(W, H, rank, world, 8))
    else:
        fb = _render(sharding.strata_frames(1, FRAMES, rank, world))
    t = torch.from_numpy(fb.reshape(-1).copy())
    sharding.reduce_framebuffer(t, dist, dst=0)
    if rank == 0:
        np.save(out_path, t.numpy().reshape(-1, 4))
    dist.destroy_process_group()


@pytest.mark.parametrize("split", ["tile", "strata"])
def test_two_ranks_reduce_to_single_render(tmp_path, split):
    out = str(tmp_path / "fb.npy")
    port = 29500 + (os.getpid() % 2000) + (0 if split == "tile" else 1)
    mp.spawn(_worker, args=(2, port, split, out), nprocs=2, join=True)
    got = np.load(out)
    want = _render(range(1, FRAMES + 1))
    if split == "tile":
        got[:, 3] = np.where(got[:, 3] > 0, 1.0, 0.0)
        assert np.array_equal(bits(got), bits(want))                 # disjoint support: bit-exact
    else:
        ok = np.isfinite(want[:, :3]).all(axis=1)
        assert np.allclose(got[ok, :3], want[ok, :3], rtol=1e-5, atol=1e-6)   # summation order differs: tolerance 1e-5


def test_shard_arithmetic():
    for (w, h, world, rows) in ((40, 28, 2, 8), (1920, 1080, 8, 8), (33, 17, 4, 3), (16, 16, 3, 16)):
        total = np.zeros(w * h, int)
        for r in range(world):
            m = sharding.tile_mask(w, h, r, world, rows)
            total += m
            assert sharding.shard_pixel_count(w, h, r, world, rows) >= m.sum()
            assert sharding.shard_pixel_count(w, h, r, world, rows) - m.sum() < rows * w
        assert (total == 1).all()
    assert sharding.strata_frames(1, 8, 1, 4) == [1, 5] and sharding.strata_frames(1, 8, 0, 4) == [4, 8]
    assert sum(len(sharding.strata_frames(3, 17, r, 8)) for r in range(8)) == 17
