"""The reference's own call pattern (renderer.js:163-191): one wgpt_dispatch per frame, enqueued without waiting,
one `fs` pass when the pixels are needed.  In fast mode the dispatched frames join ONE living wavefront (fast.cu
sessions); the result must equal the batched wgpt_render and the oracle bit for bit."""
import time

import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, Camera, MODE_EXACT, MODE_FAST
from oracle.binding import Oracle
from tests import scenes
from tests.util import bits

pytestmark = pytest.mark.gpu


def setup(x, scene, cam, frame=1):
    x.write_scene(scene.buffers())
    x.write_uniforms(frame, 0, cam.viewMatrix)
    return x


@pytest.mark.parametrize("kw", [dict(), dict(num_samples=4, stratify=True, importance_sampling=True)])
def test_n_dispatches_and_one_present_equal_render(default_scene, kw):
    scene, cam = default_scene
    W, H, N = 160, 96, 12
    o = setup(Oracle(W, H, **kw), scene, cam)
    o.render(1, N)
    o.write_uniforms(N, 0, cam.viewMatrix)
    want_fb, want_img = bits(o.read_framebuffer()), bits(o.present())
    a = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    for f in range(1, N + 1):
        a.write_uniforms(f, 0, cam.viewMatrix)
        a.dispatch()
    img = a.present()
    assert np.array_equal(bits(img), want_img)
    assert np.array_equal(bits(a.read_framebuffer()), want_fb)
    b = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    b.render(1, N)
    assert np.array_equal(bits(b.read_framebuffer()), want_fb)
    # mixed: render, then dispatches that continue it, then a render again
    c = setup(PathTracer(W, H, mode=MODE_FAST, **kw), scene, cam)
    c.render(1, 4)
    for f in (5, 6, 7):
        c.write_uniforms(f, 0, cam.viewMatrix)
        c.dispatch()
    c.render(8, N - 7)
    assert np.array_equal(bits(c.read_framebuffer()), want_fb)


def test_dispatch_sessions_break_where_they_must(default_scene, golden_meshes):
    """A new view, a frame-number gap, a reset or a buffer write between dispatches must not leak into frames already
    submitted (they keep the uniforms they were dispatched with)."""
    scene, cam = default_scene
    cam2 = Camera()
    cam2.set_camera([0.3, 0.2, 2.2], [0.5, 0, 0], [0, 1, 0])
    W, H = 128, 80
    o = setup(Oracle(W, H), scene, cam)
    g = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    seq = [(1, 0, cam), (2, 0, cam), (3, 0, cam2), (4, 0, cam2), (9, 0, cam2), (10, 0, cam), (11, 1, cam), (2, 0, cam), (3, 0, cam)]
    for f, reset, c in seq:
        o.write_uniforms(f, reset, c.viewMatrix)
        o.dispatch()
        g.write_uniforms(f, reset, c.viewMatrix)
        g.dispatch()
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))
    # a scene write between dispatches: frames 4, 5 see the old materials, 6, 7 the new ones
    bufs = scene.buffers()
    mats = np.array(bufs[8], np.float32).copy()
    mats[0:3] = [0.9, 0.1, 0.9]
    for x in (o, g):
        x.write_buffer(3, np.zeros(W * H * 4, np.float32))
        for f in (4, 5):
            x.write_uniforms(f, 0, cam.viewMatrix)
            x.dispatch()
        x.write_buffer(8, mats)
        for f in (6, 7):
            x.write_uniforms(f, 0, cam.viewMatrix)
            x.dispatch()
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))


def test_dispatch_returns_before_the_frames_are_done(default_scene):
    """wgpt_dispatch enqueues and returns (include/wgpt.h): 32 dispatches of a 1024x768 frame take far less host time
    than the render they start; the flush happens in read_buffer."""
    scene, cam = default_scene
    W, H, N = 1024, 768, 32
    g = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    g.render(1, 2)
    g.write_buffer(3, np.zeros(W * H * 4, np.float32))
    g.synchronize()
    t0 = time.perf_counter()
    for f in range(1, N + 1):
        g.write_uniforms(f, 0, cam.viewMatrix)
        g.dispatch()
    t_enqueue = time.perf_counter() - t0
    got = g.read_framebuffer()
    t_total = time.perf_counter() - t0
    assert t_enqueue < 0.5 * t_total, (t_enqueue, t_total)
    e = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    e.render(1, N)
    assert np.array_equal(bits(e.read_framebuffer()), bits(got))


def test_progressive_refresh_callback_and_pfm(default_scene, tmp_path):
    """renderProgressive: dispatches between refreshes, `on_refresh` sees exactly what `fs` shows at that frame count
    (the oracle presenting after the same number of frames); the .pfm file holds the float image."""
    from webgpu_path_tracer_b200 import Renderer, Scene, Camera, imageio
    scene = Scene()
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
    W, H = 96, 64
    r = Renderer.create(mode=MODE_FAST)
    r.initBuffers(scene, cam, W, H)
    r.setRenderParameters({}, cam, W, H)
    seen = []
    r.renderProgressive(10, 4, lambda f, img: seen.append((f, img.copy())))
    assert [f for f, _ in seen] == [4, 8, 10]
    o = Oracle(W, H)
    o.write_scene(scene.buffers())
    done = 0
    for f, img in seen:
        o.write_uniforms(done + 1, 0, cam.viewMatrix)
        o.render(done + 1, int(f) - done)
        done = int(f)
        o.write_uniforms(f, 0, cam.viewMatrix)
        assert np.array_equal(bits(o.present()), bits(img.reshape(-1, 4)))
    path = tmp_path / "hdr.pfm"
    img = r.save_image(path)
    assert np.array_equal(bits(imageio.read_pfm(path)), bits(img[:, :, :3]))
