"""Parity of both CUDA paths with the ORACLE at the real size of every BASELINE.json configuration, and of the
paths no small test reaches: the resetBuffer = 1 sequence of renderAnimation (renderer.js:172-184, main.wgsl:22-25,
fragment.js:30-33), the work counters bench.py derives B_ray from, and direction components that are -0.0.
Everything is required BIT-EXACT (tests/util.py `bits` only canonicalises NaN payloads)."""
import hashlib
import json
import os

import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, Renderer, Scene, Camera, MODE_EXACT, MODE_FAST
from oracle.binding import Oracle
from tests import scenes
from tests.util import bits

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "primary_sha.json")))


def setup(x, scene, cam, frame=1):
    x.write_scene(scene.buffers())
    x.write_uniforms(frame, 0, cam.viewMatrix)
    return x


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("cfg", ["C1", "C2", "C3", "C4", "C5"])
def test_full_size_frame_fast_and_exact_equal_the_oracle(cfg):
    """One frame of the configuration at its benchmark size (C1 512x512; C2-C4 1920x1080; C5 3840x2160 with 4
    stratified samples and two meshes): primary-hit type / index / t of EVERY pixel and the frame's radiance,
    fast mode and exact mode, against the CPU oracle; and the oracle against its committed digests."""
    import bench_scenes
    scene, cam, W, H, kw = bench_scenes.build(cfg)
    gold = GOLDEN[cfg]
    frame = gold["frame"]
    o = setup(Oracle(W, H, **kw), scene, cam, frame=frame)
    ro, to, io, do = o.probe()
    assert [W, H] == gold["size"]
    assert sha(to.astype(np.int32)) == gold["type"] and sha(io.astype(np.int32)) == gold["index"]
    assert sha(bits(do)) == gold["t"] and sha(bits(ro)) == gold["radiance"]
    for mode in (MODE_FAST, MODE_EXACT):
        g = setup(PathTracer(W, H, mode=mode, **kw), scene, cam, frame=frame)
        rg, tg, ig, dg = g.probe()
        g.close()
        assert int((to != tg).sum()) == 0 and int((io != ig).sum()) == 0, "primary ids, mode %d" % mode
        assert np.array_equal(bits(do), bits(dg)), "primary t, mode %d" % mode
        assert np.array_equal(bits(ro), bits(rg)), "radiance, mode %d" % mode
    if cfg != "C1":
        assert (to == 2).sum() > 10000          # the mesh is really in view


def test_c1_as_defined_512x512_16spp_accumulated():
    """BASELINE.json configs[0] as written: default scene, 512x512, frames 1..16 summed into binding 3."""
    import bench_scenes
    scene, cam, W, H, kw = bench_scenes.build("C1")
    assert (W, H) == (512, 512)
    o = setup(Oracle(W, H, **kw), scene, cam)
    o.render(1, 16)
    want = bits(o.read_framebuffer())
    for mode in (MODE_FAST, MODE_EXACT):
        g = setup(PathTracer(W, H, mode=mode, **kw), scene, cam)
        g.render(1, 16)
        got = bits(g.read_framebuffer())
        g.close()
        assert np.array_equal(want, got), "mode %d: %d pixels differ" % (mode, int((want != got).any(axis=1).sum()))


@pytest.mark.parametrize("mode", [MODE_FAST, MODE_EXACT])
def test_reset_sequence_of_render_animation(default_scene, mode):
    """renderAnimation with a camera move in the middle: frames 1..3 accumulate; the move makes the next dispatch run
    with resetBuffer = 1 and the STALE frameNum 4 (the frame OVERWRITES binding 3, main.wgsl:22-25), `fs` divides by
    4 and CLEARS binding 3 (fragment.js:30-33), frameNum restarts (renderer.js:174-180) and frame 2 adds to zeros."""
    scene, cam = default_scene
    W, H = 96, 64
    cam2 = Camera()
    cam2.set_camera([0.4, 0.1, 2.4], [0.5, 0, 0], [0, 1, 0])
    o = setup(Oracle(W, H), scene, cam)
    r = Renderer.create(mode=mode)
    r.initBuffers(scene, cam, W, H)
    r.setRenderParameters({}, cam, W, H)

    def oracle_frame(frame, reset, c):
        o.write_uniforms(frame, reset, c.viewMatrix)
        o.dispatch()
        fb = o.read_framebuffer()
        return fb, o.present()

    for f in (1, 2, 3):
        fb_o, img_o = oracle_frame(f, 0, cam)
        img_g = r.renderAnimation()
        assert np.array_equal(bits(img_o), bits(img_g.reshape(-1, 4)))
        assert np.array_equal(bits(o.read_framebuffer()), bits(r.webGPU.read_framebuffer()))
    # the camera moves: this frame is dispatched with frameNum = 4, resetBuffer = 1
    r.camera = cam2
    cam2.MOVING = 1
    fb_o, img_o = oracle_frame(4, 1, cam2)
    img_g = r.renderAnimation()
    assert r.frameNum == 1
    assert np.array_equal(bits(img_o), bits(img_g.reshape(-1, 4)))
    assert not r.webGPU.read_framebuffer().any() and not o.read_framebuffer().any()      # fs cleared binding 3
    cam2.MOVING = 0
    for f in (2, 3):
        fb_o, img_o = oracle_frame(f, 0, cam2)
        img_g = r.renderAnimation()
        assert r.frameNum == f
        assert np.array_equal(bits(img_o), bits(img_g.reshape(-1, 4)))
        assert np.array_equal(bits(o.read_framebuffer()), bits(r.webGPU.read_framebuffer()))
    # reset WITHOUT a present in between: the overwritten frame stays in binding 3
    o.write_uniforms(7, 1, cam.viewMatrix)
    o.dispatch()
    r.webGPU.write_uniforms(7, 1, cam.viewMatrix)
    r.webGPU.dispatch()
    assert np.array_equal(bits(o.read_framebuffer()), bits(r.webGPU.read_framebuffer()))


@pytest.mark.parametrize("which", ["C1", "mesh"])
def test_exact_mode_counters_equal_the_oracle_counters(default_scene, golden_meshes, which):
    """bench.py builds B_ray / B_sample / Mrays/s from the exact-mode kernel's counter block (SURVEY.md §8(d) defines
    them on the reference's traversal): that block must equal the oracle's, counter by counter."""
    if which == "C1":
        scene, cam = default_scene
        kw = {}
    else:
        data = {"vertices": golden_meshes["monkey_smooth_15744_vertices"], "normals": golden_meshes["monkey_smooth_15744_normals"]}
        scene, cam = scenes.mesh_in_room(data, scenes.GLASS_MAT, [("scale", 0.6, 0.6, 0.6), ("translate", 0.1, -0.3, 0)])
        kw = dict(importance_sampling=True)
    W, H = 160, 96
    o = setup(Oracle(W, H, **kw), scene, cam)
    e = setup(PathTracer(W, H, mode=MODE_EXACT, **kw), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    o.counters(reset=True)
    e.set_counters(True)
    g.set_counters(True)
    o.render(1, 2)
    e.render(1, 2)
    g.render(1, 2)
    co, ce, cg = o.counters(), e.counters(), g.counters()
    assert co == ce
    # the fast path walks another tree (node visits / triangle tests differ by design) but must do the same
    # closest-hit queries, analytic tests, RNG draws and samples
    for k in ("rays", "sphere_tests", "quad_tests", "rng_draws", "samples"):
        assert co[k] == cg[k], k


def test_negative_zero_direction_components(golden_meshes):
    """a -0.0 direction component clamps to -eps; near / far planes must follow the sign of the clamped
    value.  Rays with +0.0 and -0.0 in every component (origins inside the mesh's slab), fast against exact, and
    a render whose centre row / column produce -0.0 (shootRay.wgsl:39 `-1 * (2y - 1)` = -0.0) with an axis-aligned camera."""
    data = {"vertices": golden_meshes["monkey_smooth_15744_vertices"], "normals": golden_meshes["monkey_smooth_15744_normals"]}
    scene, cam = scenes.mesh_in_room(data, scenes.MIRROR_MAT, [("scale", 0.7, 0.7, 0.7), ("translate", 0.0, -0.2, 0)])
    e = PathTracer(8, 8, mode=MODE_EXACT)
    g = PathTracer(8, 8, mode=MODE_FAST)
    e.write_scene(scene.buffers())
    g.write_scene(scene.buffers())
    rng = np.random.default_rng(5)
    n = 60000
    o = rng.uniform(-0.5, 0.5, (n, 3)).astype(np.float32)
    o[:, 1] -= 0.2
    d = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    for k in range(6):                       # component k % 3 is +0.0 (k < 3) or -0.0
        sl = slice(k * n // 8, (k + 1) * n // 8)
        d[sl, k % 3] = 0.0 if k < 3 else -0.0
    sl = slice(6 * n // 8, 7 * n // 8)      # two zero components: axis-parallel rays of both signs
    d[sl, 0] = -0.0
    d[sl, 1] = 0.0
    sl = slice(7 * n // 8, n)
    d[sl, 1] = -0.0
    d[sl, 2] = -0.0
    rays = np.concatenate([o, d], axis=1)
    assert np.signbit(rays[:, 3:]).any() and (rays[:, 3:] == 0).any()
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    te, ie, de = e.trace_rays(rays, seeds)
    tg, ig, dg = g.trace_rays(rays, seeds)
    assert (te == 2).sum() > 1000
    assert int((te != tg).sum()) == 0 and int((ie != ig).sum()) == 0
    assert np.array_equal(bits(de), bits(dg))
    # identity-rotation camera below / left of the origin, odd image size: the centre row and column have s or t = -0.0
    cam = Camera()
    cam.set_camera([-0.2, -0.3, 2.5], [-0.2, -0.3, 0], [0, 1, 0])
    W, H = 129, 97
    o_ = setup(Oracle(W, H), scene, cam)
    gf = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    ro, to, io, do = o_.probe()
    rg, tg, ig, dg = gf.probe()
    assert int((to != tg).sum() + (io != ig).sum()) == 0
    assert np.array_equal(bits(do), bits(dg)) and np.array_equal(bits(ro), bits(rg))
