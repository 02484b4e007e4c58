"""GPU parity of the fast path (K2 wide BVH + K3 wavefront + K4 accumulate) against the CPU oracle
and against the exact-mode kernel, through the C ABI.  Everything is required BIT-EXACT: primary
hit (type, index, t), per-frame radiance and the accumulated framebuffer."""
import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, MODE_EXACT, MODE_FAST
from oracle.binding import Oracle
from tests import scenes
from tests.util import bits

pytestmark = pytest.mark.gpu




def setup(x, scene, cam, frame=1):
    x.write_scene(scene.buffers())
    x.write_uniforms(frame, 0, cam.viewMatrix)
    return x


def test_default_scene_fast_probe_matches_oracle(default_scene):
    scene, cam = default_scene
    W = H = 128
    o = setup(Oracle(W, H), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    for frame in (1, 5):
        o.write_uniforms(frame, 0, cam.viewMatrix)
        g.write_uniforms(frame, 0, cam.viewMatrix)
        ro, to, io, do = o.probe()
        rg, tg, ig, dg = g.probe()
        assert np.array_equal(to, tg) and np.array_equal(io, ig)
        assert np.array_equal(bits(do), bits(dg))
        assert np.array_equal(bits(ro), bits(rg))


def test_default_scene_fast_accumulation_matches_oracle(default_scene):
    scene, cam = default_scene
    W, H = 96, 64
    o = setup(Oracle(W, H), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    o.render(1, 6)
    g.render(1, 6)
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))
    # a second call continues the progressive sum (frames 7..9), as renderAnimation would
    o.render(7, 3)
    g.render(7, 3)
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))


@pytest.mark.parametrize("mesh,material,is_on", [
    ("monkey_968", scenes.DRAGON_MAT, False),
    ("monkey_968", scenes.GLASS_MAT, True),
    ("icosphere", scenes.MIRROR_MAT, True),
    ("monkey_smooth_15744", scenes.DRAGON_MAT, True),
])
def test_mesh_scene_fast_matches_oracle(golden_meshes, mesh, material, is_on):
    data = {"vertices": golden_meshes[mesh + "_vertices"], "normals": golden_meshes[mesh + "_normals"]}
    scene, cam = scenes.mesh_in_room(data, material, [("scale", 0.6, 0.6, 0.6), ("rotate", 0.5, [0, 1, 0]), ("translate", 0.1, -0.3, 0)])
    W, H = 160, 96
    kw = dict(importance_sampling=is_on)
    o = setup(Oracle(W, H, **kw), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    ro, to, io, do = o.probe()
    rg, tg, ig, dg = g.probe()
    assert int((to != tg).sum() + (io != ig).sum()) == 0
    assert np.array_equal(bits(do), bits(dg))
    assert np.array_equal(bits(ro), bits(rg))
    o.render(1, 3)
    g.render(1, 3)
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))


def test_stratified_multisample_fast_matches_oracle(default_scene):
    scene, cam = default_scene
    W, H = 64, 48
    kw = dict(num_samples=4, stratify=True, importance_sampling=True)
    o = setup(Oracle(W, H, **kw), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    e = setup(PathTracer(W, H, mode=MODE_EXACT, **kw), scene, cam)
    for x in (o, g, e):
        x.render(1, 3)
    want = bits(o.read_framebuffer())
    assert np.array_equal(want, bits(e.read_framebuffer()))
    assert np.array_equal(want, bits(g.read_framebuffer()))
    kw = dict(num_samples=3)
    o = setup(Oracle(W, H, **kw), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    o.render(1, 2)
    g.render(1, 2)
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))


def test_wide_traversal_equals_reference_walk_on_random_rays(golden_meshes):
    """K2 validation: same closest hit as the reference-order walk."""
    data = {"vertices": golden_meshes["monkey_smooth_15744_vertices"], "normals": golden_meshes["monkey_smooth_15744_normals"]}
    scene, cam = scenes.mesh_in_room(data, scenes.DRAGON_MAT, [("scale", 0.7, 0.7, 0.7), ("translate", 0.0, -0.2, 0)])
    e = PathTracer(8, 8, mode=MODE_EXACT)
    g = PathTracer(8, 8, mode=MODE_FAST)
    e.write_scene(scene.buffers())
    g.write_scene(scene.buffers())
    rays, seeds = scenes.random_rays(2_000_000, 7)
    te, ie, de = e.trace_rays(rays, seeds)
    tg, ig, dg = g.trace_rays(rays, seeds)
    assert int((te != tg).sum()) == 0 and int((ie != ig).sum()) == 0
    assert np.array_equal(bits(de), bits(dg))
    st = g.fast_statistics()
    assert st["wide_tris"] == 15744 and st["wide_nodes"] > 0
    # and against the oracle's brute force on a subset
    o = Oracle(8, 8)
    o.write_scene(scene.buffers())
    tb, ib, db = o.trace_rays(rays[:20000], seeds[:20000], brute_force=False)
    assert np.array_equal(tb, tg[:20000]) and np.array_equal(ib, ig[:20000]) and np.array_equal(bits(db), bits(dg[:20000]))


@pytest.mark.parametrize("tree", ["device", "reference"])
def test_collapse_plan_and_greedy_collapse_give_the_same_hits(golden_meshes, monkeypatch, tree):
    """K2's SAH-optimal collapse plan (default) and the largest-area-first expansion (WGPT_COLLAPSE=greedy) are two
    shapes of the same superset search: identical closest hits on random rays, every triangle in exactly one leaf slot."""
    data = {"vertices": golden_meshes["monkey_smooth_15744_vertices"], "normals": golden_meshes["monkey_smooth_15744_normals"]}
    scene, cam = scenes.mesh_in_room(data, scenes.DRAGON_MAT, [("scale", 0.7, 0.7, 0.7), ("translate", 0.0, -0.2, 0)])
    rays, seeds = scenes.random_rays(500_000, 11)
    monkeypatch.setenv("WGPT_BVH", tree)
    out = {}
    for how in ("greedy", "plan"):
        monkeypatch.setenv("WGPT_COLLAPSE", how)
        g = PathTracer(8, 8, mode=MODE_FAST)
        g.write_scene(scene.buffers())
        out[how] = g.trace_rays(rays, seeds) + (g.fast_statistics(),)
        g.close()
    e = PathTracer(8, 8, mode=MODE_EXACT)
    e.write_scene(scene.buffers())
    te, ie, de = e.trace_rays(rays, seeds)
    for how in ("greedy", "plan"):
        t, i, d, st = out[how]
        assert np.array_equal(t, te) and np.array_equal(i, ie) and np.array_equal(bits(d), bits(de))
        assert st["wide_tris"] == 15744
        assert 15744 // 8 < st["wide_nodes"] < 15744


def test_coplanar_ties_resolve_like_the_reference_walk():
    """The atrium stand-in is full of coplanar, overlapping faces (column caps on the floor and under the
    lintels): rays hit two triangles at exactly the same distance and the reference keeps the one it visits
    last.  The fast path must reproduce that through its tie resolver (restricted reference-order walk)."""
    from webgpu_path_tracer_b200 import meshgen, Scene, Camera
    m = meshgen.atrium(target_tris=9000)
    s = Scene(default=False)
    s.objs.append([])
    am = s.add_material
    s.add_quad([-0.5, 0.995, -0.4], [1.0, 0, 0], [0, 0, 0.8], am("light", 0, [0, 0, 0], [0, 0, 0], [20, 20, 20], 0, 0, 0))
    s.lights.append(s.quads[0])
    s.objs.append(list(s.quads))
    s.add_mesh(m, am("stone", 0, [0.76, 0.70, 0.51], [0.76, 0.70, 0.51], [0, 0, 0], 0.05, 0.95, 0))
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera([-1.8, -0.3, 0.0], [1.0, -0.1, 0.1], [0, 1, 0])
    W, H = 192, 108
    o = setup(Oracle(W, H), s, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST), s, cam)
    ro, to, io, do = o.probe()
    rg, tg, ig, dg = g.probe()
    assert int((to != tg).sum() + (io != ig).sum()) == 0
    assert np.array_equal(bits(do), bits(dg))
    o.render(1, 4)
    g.render(1, 4)
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))
    assert g.fast_statistics()["fallback_rays"] > 0          # the resolver really ran
    # closest-hit hook on random rays inside the atrium, against the exact-mode kernel
    e = PathTracer(8, 8, mode=MODE_EXACT)
    e.write_scene(s.buffers())
    rays, seeds = scenes.random_rays(500_000, 11)
    te, ie, de = e.trace_rays(rays, seeds)
    tg, ig, dg = g.trace_rays(rays, seeds)
    assert int((te != tg).sum()) == 0 and int((ie != ig).sum()) == 0
    assert np.array_equal(bits(de), bits(dg))


def test_renderer_mirror_saves_the_presented_image(default_scene, tmp_path):
    """Renderer.renderFrames + save_image (N4): the file holds exactly wgpt_present's RGBA8 output, and the
    reset semantics of renderAnimation (camera moving -> resetBuffer = 1, frameNum back to 1) hold in fast mode."""
    from webgpu_path_tracer_b200 import Renderer, Scene, Camera, imageio
    scene = Scene()
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
    r = Renderer.create(mode=MODE_FAST)
    r.initBuffers(scene, cam, 96, 64)
    r.setRenderParameters({}, cam, 96, 64)
    r.renderFrames(4)
    path = tmp_path / "frame.png"
    rgba8 = r.save_image(path)
    assert np.array_equal(imageio.decode_png(path.read_bytes()), rgba8[:, :, :3])
    o = Oracle(96, 64)
    o.write_scene(scene.buffers())
    o.write_uniforms(1, 0, cam.viewMatrix)
    o.render(1, 4)
    assert np.array_equal(bits(o.read_framebuffer()), bits(r.webGPU.read_framebuffer()))


def test_bytes_only_present_equals_the_pinned_pow_for_every_float(default_scene):
    """K4 present to the 8-bit canvas skips the binary64 pow when an SFU estimate already decides the byte.
    Exhaustive proof: round(255 pow(r, 1/2.2)) agrees with the pinned pow for EVERY float r in [0, 1] (the
    tone map clamps to that range), for -0, and for NaN / out-of-range patterns (which take the pinned path);
    and on a rendered image the bytes-only present equals the bytes of the float present."""
    scene, cam = default_scene
    g = setup(PathTracer(96, 64, mode=MODE_FAST), scene, cam)
    bad, slow = g.present_quant_check(0, 0x3f800000 + 1)
    assert bad == 0
    assert slow < 0.01 * 0x3f800000          # the pinned path is the exception
    for first, count in ((0x3f800001, 1 << 16), (0x7f800000, 1 << 16), (0x80000000, 1 << 16), (0xffc00000, 16)):
        assert g.present_quant_check(first, count)[0] == 0
    g.render(1, 5)
    g.write_uniforms(5, 0, cam.viewMatrix)
    both = g.present(rgba8=True)[1]
    only = g.present(rgba8="only")
    assert np.array_equal(both, only)
    o = Oracle(96, 64)
    o.write_scene(scene.buffers())
    o.write_uniforms(1, 0, cam.viewMatrix)
    o.render(1, 5)
    o.write_uniforms(5, 0, cam.viewMatrix)
    img = np.asarray(o.present()).reshape(-1, 4)
    assert np.array_equal(only[:, :3], np.rint(img[:, :3] * np.float32(255.0)).astype(np.uint8))
