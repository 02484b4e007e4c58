"""Edge cases of the hot path on the GPU, through the C ABI, bit-exact against the CPU oracle: scenes with no
triangles / nothing at all, image sizes that are no multiple of a workgroup or a tile, several meshes with
different transforms, a traversal stack small enough for the reference's guard (hitRay.wgsl:100-104) to fire,
the lowest bounce limits, and tile shards that must add up to the unsharded image."""
import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, Scene, Camera, MODE_EXACT, MODE_FAST
from oracle.binding import Oracle
from tests import scenes
from tests.util import bits

pytestmark = pytest.mark.gpu


def setup(x, scene, cam, frame=1):
    x.write_scene(scene.buffers())
    x.write_uniforms(frame, 0, cam.viewMatrix)
    return x


def check_all_modes(scene, cam, W, H, frames=3, **kw):
    o = setup(Oracle(W, H, **kw), scene, cam)
    ro, to, io, do = o.probe()
    o.render(1, frames)
    fo = o.read_framebuffer()
    for mode in (MODE_EXACT, MODE_FAST):
        g = setup(PathTracer(W, H, mode=mode, **kw), scene, cam)
        rg, tg, ig, dg = g.probe()
        assert int((to != tg).sum() + (io != ig).sum()) == 0, mode
        assert np.array_equal(bits(do), bits(dg)), mode
        assert np.array_equal(bits(ro), bits(rg)), mode
        g.render(1, frames)
        assert np.array_equal(bits(fo), bits(g.read_framebuffer())), mode
        g.close()
    return fo


def room_only():
    s = Scene(default=False)
    s.add_sphere([0.3, -0.6, 0.1], 0.4, s.add_material("ball", *scenes.MIRROR_MAT))
    s.add_sphere([-0.4, -0.7, -0.3], 0.3, s.add_material("glass", *scenes.GLASS_MAT))
    s.objs.append(list(s.spheres))
    scenes.cornell_room(s)
    cam = Camera()
    cam.set_camera([0, 0, 2.5], [0, 0, 0], [0, 1, 0])
    return s, cam


@pytest.mark.parametrize("W,H", [(37, 23), (64, 1), (1, 64), (130, 67)])
def test_scene_without_triangles_and_ragged_image_sizes(W, H):
    """No mesh, no BVH (bindings 5, 6, 7, 9 empty): the traversal stage has nothing to do; sizes that are
    no multiple of the 64-thread workgroup (renderer.js:187-188 dispatches W*H/64 + 1 groups) or of a tile."""
    scene, cam = room_only()
    check_all_modes(scene, cam, W, H)


def test_empty_scene_is_all_background():
    s = Scene(default=False)
    cam = Camera()
    cam.set_camera([0, 0, 2.5], [0, 0, 0], [0, 1, 0])
    fb = check_all_modes(s, cam, 48, 32, frames=2)
    fb = np.asarray(fb).reshape(-1, 4)
    assert np.array_equal(fb[:, :3], np.tile(np.float32([0, 2, 2]), (48 * 32, 1)))          # traceRay.wgsl:12-16, two frames


def test_two_meshes_with_different_transforms(golden_meshes):
    """Two meshes, two materials, two inverse model matrices: the object-space ray is per (ray, mesh)
    (common.wgsl:193-197) and a ray may test triangles of both."""
    s = Scene(default=False)
    scenes.cornell_room(s)
    for name, mat, tr in (("monkey_968", scenes.DRAGON_MAT, [("scale", 0.45, 0.45, 0.45), ("rotate", 0.6, [0, 1, 0]), ("translate", -0.45, -0.4, 0.1)]),
                          ("icosphere", scenes.GLASS_MAT, [("scale", 0.4, 0.5, 0.4), ("translate", 0.45, -0.3, -0.2)])):
        data = {"vertices": golden_meshes[name + "_vertices"], "normals": golden_meshes[name + "_normals"]}
        m = s.add_mesh(data, s.add_material(name, *mat))
        m.transform.update(*[getattr(m.transform, t[0])(*t[1:]) for t in tr])
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera([0, 0, 2.5], [0, 0, 0], [0, 1, 0])
    check_all_modes(s, cam, 128, 80, importance_sampling=True)


def test_stack_guard_of_the_reference_fires_identically(golden_meshes):
    """STACK_SIZE below the tree depth: the reference's walk aborts when its stack would overflow
    (hitRay.wgsl:100-104) and returns whatever it had; the fast path must fall back to that very walk."""
    data = {"vertices": golden_meshes["monkey_968_vertices"], "normals": golden_meshes["monkey_968_normals"]}
    scene, cam = scenes.mesh_in_room(data, scenes.DRAGON_MAT, [("scale", 0.7, 0.7, 0.7), ("translate", 0, -0.3, 0)])
    for stack_size in (4, 7):
        check_all_modes(scene, cam, 96, 64, frames=2, stack_size=stack_size)


@pytest.mark.parametrize("max_bounces", [1, 2])
def test_lowest_bounce_limits(default_scene, max_bounces):
    scene, cam = default_scene
    check_all_modes(scene, cam, 80, 48, frames=2, max_bounces=max_bounces)


def test_tile_shards_add_up_to_the_unsharded_image(default_scene):
    """interleaved row-block shards render disjoint pixels in frame order, so their
    framebuffers add up (zeros elsewhere) to the single-GPU image bit for bit; 3 shards, 5-row tiles, a
    height that is no multiple of either."""
    scene, cam = default_scene
    W, H, frames = 96, 53, 3
    whole = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    whole.render(1, frames)
    ref = np.asarray(whole.read_framebuffer()).reshape(-1, 4)
    total = np.zeros_like(ref)
    for r in range(3):
        g = setup(PathTracer(W, H, mode=MODE_FAST, tile_rank=r, tile_world=3, tile_rows=5), scene, cam)
        g.render(1, frames)
        part = np.asarray(g.read_framebuffer()).reshape(-1, 4)
        assert not np.any((part[:, 3] != 0) & (total[:, 3] != 0))                      # shards are disjoint
        total += part
        g.close()
    assert np.array_equal(bits(ref), bits(total))


def synthetic_mesh(kind, seed=5):
    """Triangle soups that stress the hierarchy build and the quantised boxes."""
    rng = np.random.default_rng(seed)
    if kind == "duplicates_and_degenerates":
        # every triangle twice (an exact tie on every hit), plus zero-area and zero-extent triangles
        base = rng.uniform(-0.6, 0.6, (60, 1, 3)).astype(np.float32) + rng.uniform(-0.25, 0.25, (60, 3, 3)).astype(np.float32)
        line = np.stack([base[:8, 0], base[:8, 1], (base[:8, 0] + base[:8, 1]) * np.float32(0.5)], axis=1)     # collinear
        point = np.repeat(base[8:12, :1], 3, axis=1)                                                                # a single point
        v = np.concatenate([base, base, line, point], axis=0)
    elif kind == "flat":
        # all triangles in the plane z = 0.125: every box has zero extent in z (the reference pads it, AABB.js:35-51)
        xy = rng.uniform(-0.8, 0.8, (200, 1, 2)).astype(np.float32) + rng.uniform(-0.15, 0.15, (200, 3, 2)).astype(np.float32)
        v = np.concatenate([xy, np.full((200, 3, 1), 0.125, np.float32)], axis=2)
    elif kind == "scale_disparity":
        # two huge triangles spanning the room and a cloud of tiny ones: one quantisation frame holds both scales
        big = np.float32([[[-0.95, -0.9, -0.5], [0.95, -0.9, -0.5], [0.0, 0.9, -0.45]],
                          [[-0.95, -0.9, 0.2], [0.0, 0.9, 0.25], [0.95, -0.9, 0.2]]])
        c = rng.uniform(-0.5, 0.5, (400, 1, 3)).astype(np.float32)
        tiny = c + rng.uniform(-2e-3, 2e-3, (400, 3, 3)).astype(np.float32)
        v = np.concatenate([big, tiny], axis=0)
    else:
        raise KeyError(kind)
    v = np.ascontiguousarray(v, np.float32)
    e1, e2 = v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]
    n = np.cross(e1, e2)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    n = np.where(ln > 0, n / np.maximum(ln, 1e-30), np.float32([0, 0, 1])).astype(np.float32)
    return {"vertices": v.reshape(-1), "normals": np.repeat(n[:, None, :], 3, axis=1).reshape(-1)}


@pytest.mark.parametrize("kind", ["duplicates_and_degenerates", "flat", "scale_disparity"])
def test_hard_triangle_soups(kind):
    scene, cam = scenes.mesh_in_room(synthetic_mesh(kind), scenes.DRAGON_MAT, [("rotate", 0.3, [0, 1, 0])])
    check_all_modes(scene, cam, 128, 96, frames=2)
    rays, seeds = scenes.random_rays(200_000, 11)
    e = setup(PathTracer(8, 8, mode=MODE_EXACT), scene, cam)
    g = setup(PathTracer(8, 8, mode=MODE_FAST), scene, cam)
    te, ie, de = e.trace_rays(rays, seeds)
    tg, ig, dg = g.trace_rays(rays, seeds)
    assert int((te != tg).sum()) == 0 and int((ie != ig).sum()) == 0
    assert np.array_equal(bits(de), bits(dg))


def test_scene_far_from_the_origin(golden_meshes):
    """Everything translated by (1000, -500, 250): f32 coordinates lose 10 bits, the quantisation frames sit on
    large origins, the reference's +-5e-5 box pad falls below one ulp of the coordinates."""
    off = np.float32([1000.0, -500.0, 250.0])
    s = Scene(default=False)
    am = s.add_material
    s.add_sphere(list(np.float32([0.4, -0.6, 0.2]) + off), 0.35, am("ball", *scenes.MIRROR_MAT))
    s.objs.append(list(s.spheres))
    e = 20.0
    s.add_quad(list(np.float32([-0.35, 0.9999, -0.25]) + off), [0.7, 0, 0], [0, 0, 0.5], am("light", 0, [0, 0, 0], [0, 0, 0], [e, e, e], 0, 0, 0))
    s.add_quad(list(np.float32([-1, -1, -1]) + off), [2, 0, 0], [0, 2, 0], am("back", 0, [0.7, 0.7, 0.7], [0.7, 0.7, 0.7], [0, 0, 0], 0.05, 0.95, 0))
    s.add_quad(list(np.float32([-1, -1, -1]) + off), [2, 0, 0], [0, 0, 2], am("floor", 0, [0.6, 0.6, 0.3], [0.6, 0.6, 0.3], [0, 0, 0], 0.05, 0.95, 0))
    s.lights.append(s.quads[0])
    s.objs.append(list(s.quads))
    data = {"vertices": golden_meshes["monkey_968_vertices"], "normals": golden_meshes["monkey_968_normals"]}
    m = s.add_mesh(data, am("meshMat", *scenes.DRAGON_MAT))
    m.transform.update(m.transform.scale(0.6, 0.6, 0.6), m.transform.translate(float(off[0]), float(off[1]) - 0.3, float(off[2])))
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera(list(np.float32([0, 0, 2.5]) + off), list(off), [0, 1, 0])
    check_all_modes(s, cam, 128, 96, frames=2)


@pytest.mark.parametrize("pool,ring", [(1024, 2), (3072, 5), (65536, 64)])
def test_the_image_does_not_depend_on_the_schedule(golden_meshes, monkeypatch, pool, ring):
    """Which slot runs which (pixel, frame) item, how many frames are in flight and when the drain phase switches the
    shade stage to its live-slot list are scheduling matters: a pool much smaller than a frame with a 2-frame window
    (slots idle at the window, many start/finish generations, a long drain) renders the same bits as the oracle.
    Default scene: fog, glass and mirror paths of very different lengths; mesh scene: glass mesh with the mixture pdf
    (ties and the marked-slots-first order of the shade grid included)."""
    monkeypatch.setenv("WGPT_POOL", str(pool))
    monkeypatch.setenv("WGPT_RING", str(ring))
    s = Scene()
    s.init_mesh_data(); s.create_meshes(); s.create_bvh()
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
    cases = [(s, cam, {}, 9)]
    data = {"vertices": golden_meshes["monkey_968_vertices"], "normals": golden_meshes["monkey_968_normals"]}
    ms, mcam = scenes.mesh_in_room(data, scenes.GLASS_MAT, [("scale", 0.6, 0.6, 0.6), ("rotate", 0.5, [0, 1, 0]), ("translate", 0.1, -0.3, 0)])
    cases.append((ms, mcam, {"importance_sampling": True}, 7))
    W, H = 80, 48
    for scene, c, kw, frames in cases:
        o = setup(Oracle(W, H, **kw), scene, c)
        o.render(1, frames)
        g = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, c)
        g.render(1, frames)
        st = g.fast_statistics()
        assert st["pool_slots"] == pool and st["ring_frames"] == ring
        assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))
        # a second render on the same context starts from a clean control block (drain / live-list state reset)
        g.write_buffer(3, np.zeros(W * H * 4, np.float32))
        g.render(1, frames)
        assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))
        g.close()
