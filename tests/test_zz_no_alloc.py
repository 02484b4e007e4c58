"""wgpt_present_rgba8_static (include/wgpt.h): the allocation-free present, through the C ABI, against the existing
entry point wgpt_present(ctx, NULL, rgba8) -- same kernel, so the bytes must be IDENTICAL (the tolerance of every
byte / integer test here: 0) -- on two calls in a row (the static staging array keeps the previous call's contents),
on a second image size after a larger one, and with the size limit enforced.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, WgptError, MODE_EXACT, MODE_FAST

pytestmark = pytest.mark.gpu


def setup(x, scene, cam, frame=1):
    x.write_scene(scene.buffers())
    x.write_uniforms(frame, 0, cam.viewMatrix)
    return x


def test_static_present_equals_present_on_two_calls_in_a_row(default_scene):
    scene, cam = default_scene
    W, H = 200, 120
    g = setup(PathTracer(W, H, mode=MODE_FAST), scene, cam)
    g.render(1, 3)
    g.write_uniforms(3, 0, cam.viewMatrix)
    want1 = g.present(rgba8="only")
    got1 = g.present_rgba8_static()                    # call 1
    assert np.array_equal(want1, got1) and got1[:, :3].any() and (got1[:, 3] == 255).all()
    g.render(4, 2)                                      # another image: call 2 must not show call 1's bytes
    g.write_uniforms(5, 0, cam.viewMatrix)
    got2 = g.present_rgba8_static()                    # call 2, right after call 1
    want2 = g.present(rgba8="only")
    assert np.array_equal(want2, got2)
    assert not np.array_equal(got1, got2)
    got3 = g.present_rgba8_static()                    # and again, same input: same output
    assert np.array_equal(got2, got3)
    # frames dispatched but not yet flushed are presented too (the entry point flushes like wgpt_present)
    for f in (6, 7):
        g.write_uniforms(f, 0, cam.viewMatrix)
        g.dispatch()
    got4 = g.present_rgba8_static()
    assert np.array_equal(got4, g.present(rgba8="only"))


def test_static_present_smaller_image_after_a_larger_one_and_exact_mode(default_scene):
    scene, cam = default_scene
    big = setup(PathTracer(256, 160, mode=MODE_EXACT), scene, cam)
    big.render(1, 2)
    big.write_uniforms(2, 0, cam.viewMatrix)
    assert np.array_equal(big.present_rgba8_static(), big.present(rgba8="only"))
    small = setup(PathTracer(64, 48, mode=MODE_EXACT), scene, cam)
    small.render(1, 2)
    small.write_uniforms(2, 0, cam.viewMatrix)
    assert np.array_equal(small.present_rgba8_static(), small.present(rgba8="only"))
    # resetBuffer = 1: the same clear of binding 3 as fs (fragment.js:30-33)
    small.write_uniforms(2, 1, cam.viewMatrix)
    small.present_rgba8_static()
    assert not small.read_framebuffer().any()


def test_static_present_rejects_images_beyond_its_maximum():
    g = PathTracer(2048, 1025, mode=MODE_EXACT)        # 2,099,200 pixels > 2 Mi
    g.write_uniforms(1, 0, np.eye(4, dtype=np.float32).reshape(-1))
    with pytest.raises(WgptError, match="at most 2097152"):
        g.present_rgba8_static()
    ok = PathTracer(1920, 1080, mode=MODE_EXACT)       # the benchmark size fits
    ok.write_uniforms(1, 0, np.eye(4, dtype=np.float32).reshape(-1))
    out = ok.present_rgba8_static()
    assert out.shape == (1920 * 1080, 4) and np.array_equal(out, ok.present(rgba8="only"))
