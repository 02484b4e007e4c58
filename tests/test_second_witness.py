"""The oracle against a SECOND WITNESS (tests/witness.py: numpy float32, written from the WGSL, hitRec-as-global shape):
primary hit and per-invocation radiance bit for bit on 256 pixels x 2 frames of the default scene and on a glass mesh
scene with mixture importance sampling, plus a stratified multi-sample case."""
import numpy as np
import pytest

from oracle.binding import Oracle
from tests import scenes
from tests.util import bits
from tests.witness import Witness


def compare(scene, cam, W, H, pixels, frames, **kw):
    o = Oracle(W, H, **kw)
    o.write_scene(scene.buffers())
    w = Witness(W, H, scene.buffers(), cam.viewMatrix, **kw)
    n_hit = 0
    for frame in frames:
        o.write_uniforms(frame, 0, cam.viewMatrix)
        rad, ptype, pindex, pt = o.probe()
        for p in pixels:
            col, prim = w.invocation(int(p), frame)
            assert (int(ptype[p]), int(pindex[p])) == (prim[0], prim[1]), (frame, p, prim, ptype[p], pindex[p])
            assert np.array_equal(bits(np.float32(prim[2])), bits(pt[p])), (frame, p)
            assert np.array_equal(bits(np.array(col, np.float32)), bits(rad[p])), (frame, p, col, rad[p])
            n_hit += prim[0] >= 0
    return n_hit


def test_default_scene_256_pixels_2_frames(default_scene):
    scene, cam = default_scene
    W, H = 64, 48
    pixels = np.random.default_rng(1).choice(W * H, 256, replace=False)
    assert compare(scene, cam, W, H, pixels, (1, 2)) > 400


def test_glass_mesh_with_importance_sampling(golden_meshes):
    data = {"vertices": golden_meshes["monkey_968_vertices"], "normals": golden_meshes["monkey_968_normals"]}
    scene, cam = scenes.mesh_in_room(data, scenes.GLASS_MAT, [("scale", 0.6, 0.6, 0.6), ("rotate", 0.5, [0, 1, 0]), ("translate", 0.1, -0.3, 0)])
    W, H = 48, 32
    pixels = np.random.default_rng(2).choice(W * H, 256, replace=False)
    assert compare(scene, cam, W, H, pixels, (1, 2), importance_sampling=True) > 300


def test_stratified_multisample(default_scene):
    scene, cam = default_scene
    W, H = 32, 24
    pixels = np.random.default_rng(3).choice(W * H, 48, replace=False)
    compare(scene, cam, W, H, pixels, (3,), num_samples=4, stratify=True, importance_sampling=True)
