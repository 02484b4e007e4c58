"""Worker of tests/test_gpu_multi.py::test_one_process_per_gpu_reduce (launched under torchrun, one rank per GPU)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch.distributed as dist
from webgpu_path_tracer_b200 import PathTracer, Scene, Camera, MODE_FAST, sharding

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
dist.init_process_group("gloo")
W, H, FRAMES = 256, 160, 10
s = Scene()
s.init_mesh_data(); s.create_meshes(); s.create_bvh()
cam = Camera()
cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
res = {}
for name, tiles in (("tiles", world), ("strata", 1)):
    tr, tw, sr, sw = sharding.shard_of_rank(rank, world, tiles)
    pt = PathTracer(W, H, mode=MODE_FAST, device=local, tile_rank=tr, tile_world=tw, strata_rank=sr, strata_world=sw)
    sharding.init_comm(pt, dist, rank, world)
    pt.write_scene(s.buffers())
    pt.write_uniforms(1, 0, cam.viewMatrix)
    pt.render(1, FRAMES)
    pt.reduce_framebuffer(0)
    res[name] = pt.read_framebuffer()
    pt.close()
if rank == 0:
    one = PathTracer(W, H, mode=MODE_FAST, device=local)
    one.write_scene(s.buffers())
    one.write_uniforms(1, 0, cam.viewMatrix)
    one.render(1, FRAMES)
    np.savez(sys.argv[1], single=one.read_framebuffer(), **res)
dist.barrier()
dist.destroy_process_group()
