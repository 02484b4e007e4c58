"""Multi-GPU inside the product (csrc/mgpu.cu): tile x strata sharding, ncclReduce issued by the library on the contexts'
own streams.  The reduced N-GPU framebuffer against the 1-GPU one: bit-exact for tile shards (disjoint support),
<= 1e-5 relative for frame strata (summation order).  Needs >= 2 GPUs for the NCCL paths (skipped on a 1-GPU box: the
1-GPU group and the error paths still run); world-size-2 host logic on CPU is tests/test_multirank_gloo.py."""
import os
import subprocess
import sys

import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, Group, WgptError, MODE_FAST
from tests import scenes
from tests.conftest import _cuda_device_count
from tests.util import bits

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = _cuda_device_count()
W, H, FRAMES = 320, 200, 12


def scene_and_single(golden_meshes, **kw):
    data = {"vertices": golden_meshes["monkey_968_vertices"], "normals": golden_meshes["monkey_968_normals"]}
    scene, cam = scenes.mesh_in_room(data, scenes.DRAGON_MAT, [("scale", 0.6, 0.6, 0.6), ("translate", 0.1, -0.3, 0)])
    one = PathTracer(W, H, mode=MODE_FAST, **kw)
    one.write_scene(scene.buffers())
    one.write_uniforms(1, 0, cam.viewMatrix)
    one.render(1, FRAMES)
    return scene, cam, one.read_framebuffer()


def run_group(scene, cam, n, tile_groups, **kw):
    g = Group(W, H, n, tile_groups=tile_groups, **kw)
    g.write_scene(scene.buffers())
    g.write_uniforms(1, 0, cam.viewMatrix)
    g.render(1, FRAMES)
    g.reduce(0)
    fb = g.member(0).read_framebuffer()
    g.close()
    return fb


def test_group_of_one_is_the_plain_context(golden_meshes):
    scene, cam, want = scene_and_single(golden_meshes)
    got = run_group(scene, cam, 1, 1)
    assert np.array_equal(bits(want), bits(got))
    pt = PathTracer(64, 64, mode=MODE_FAST)
    with pytest.raises(WgptError):
        pt.reduce_framebuffer(0)                 # no communicator on this context


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_tile_shards_reduce_bit_exact(golden_meshes):
    scene, cam, want = scene_and_single(golden_meshes)
    n = 2 if NGPU < 4 else 4
    got = run_group(scene, cam, n, n)
    assert np.array_equal(bits(want), bits(got))


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_strata_shards_reduce_to_rounding(golden_meshes):
    scene, cam, want = scene_and_single(golden_meshes)
    got = run_group(scene, cam, 2, 1)
    assert np.allclose(want, got, rtol=1e-5, atol=1e-6, equal_nan=True)
    assert (got[:, 3] == 1.0).all()
    if NGPU >= 4:
        got = run_group(scene, cam, 4, 2)        # 2 tile groups x 2 strata
        assert np.allclose(want, got, rtol=1e-5, atol=1e-6, equal_nan=True)


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_one_process_per_gpu_reduce(tmp_path):
    """torchrun, one rank per GPU: the NCCL id travels over torch.distributed (gloo), the reduce is the library's."""
    out = tmp_path / "mp.npz"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29613", os.path.join(ROOT, "tests", "mgpu_worker.py"), str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    d = np.load(out)
    assert np.array_equal(bits(d["single"]), bits(d["tiles"]))
    assert np.allclose(d["single"], d["strata"], rtol=1e-5, atol=1e-6, equal_nan=True)
