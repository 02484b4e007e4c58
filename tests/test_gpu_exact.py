"""GPU parity of the exact-mode CUDA path (K1) against the CPU oracle, through the C ABI.
Integer/index results (primary hit type/index) and, because both sides evaluate the same pinned
IEEE arithmetic, the f32 results too are required to be BIT-EXACT."""
import numpy as np
import pytest

from webgpu_path_tracer_b200 import PathTracer, MODE_EXACT
from oracle.binding import Oracle
from tests import scenes
from tests.util import bits

pytestmark = pytest.mark.gpu




def run_pair(scene, cam, W, H, frames, **cfg):
    o = Oracle(W, H, **cfg)
    g = PathTracer(W, H, mode=MODE_EXACT, **cfg)
    for x in (o, g):
        x.write_scene(scene.buffers())
        x.write_uniforms(1, 0, cam.viewMatrix)
    return o, g


def test_pinned_math_on_gpu_matches_cpu():
    from oracle.binding import pinned_math
    g = PathTracer(8, 8)
    rng = np.random.default_rng(1)
    x = rng.uniform(0, 6.2831855, 200000).astype(np.float32)
    for fn in ("sin", "cos"):
        assert np.array_equal(bits(g.pinned_math(fn, x)), bits(pinned_math(fn, x)))
    a = rng.uniform(-1, 1, 200000).astype(np.float32)
    assert np.array_equal(bits(g.pinned_math("acos", a)), bits(pinned_math("acos", a)))
    u = rng.uniform(0, 1, 200000).astype(np.float32)
    u[:3] = [0.0, 1.0, 1e-38]
    assert np.array_equal(bits(g.pinned_math("log", u)), bits(pinned_math("log", u)))
    y = rng.choice(np.array([5.0, 2.0, 0.45454547], np.float32), 200000)
    assert np.array_equal(bits(g.pinned_math("pow", u * 1.5, y)), bits(pinned_math("pow", u * 1.5, y)))


def test_default_scene_primary_hits_and_radiance_bit_exact(default_scene):
    scene, cam = default_scene
    W = H = 128
    o, g = run_pair(scene, cam, W, H, 1)
    for frame in (1, 2, 7):
        for x in (o, g):
            x.write_uniforms(frame, 0, cam.viewMatrix)
        ro, to, io, do = o.probe()
        rg, tg, ig, dg = g.probe()
        assert np.array_equal(to, tg) and np.array_equal(io, ig)
        assert np.array_equal(bits(do), bits(dg))
        assert np.array_equal(bits(ro), bits(rg))


def test_default_scene_accumulation_and_present_bit_exact(default_scene):
    scene, cam = default_scene
    W, H = 96, 64
    o, g = run_pair(scene, cam, W, H, 4)
    o.render(1, 4)
    g.render(1, 4)
    assert np.array_equal(bits(o.read_framebuffer()), bits(g.read_framebuffer()))
    for x in (o, g):
        x.write_uniforms(4, 0, cam.viewMatrix)
    assert np.array_equal(bits(o.present()), bits(g.present()))


def test_device_only_math_forms_equal_the_plain_ones():
    """The shading code divides a vector by a scalar with ONE reciprocal refinement shared by the three
    components, and evaluates sin and cos of an angle with one range reduction; both must produce the bits
    of IEEE division / of the separate pinned functions."""
    from oracle.binding import pinned_math
    g = PathTracer(8, 8, mode=MODE_EXACT)
    rng = np.random.default_rng(11)
    n = 4_000_000
    # numerators / divisors over many binades, plus the edge cases (zeros, denormals, huge, tiny, inf, nan)
    x = (rng.standard_normal(n) * np.exp2(rng.integers(-70, 70, n))).astype(np.float32)
    y = (np.abs(rng.standard_normal(n)) * np.exp2(rng.integers(-70, 70, n))).astype(np.float32)
    edge = np.array([0.0, -0.0, 1e-45, -1e-45, 1e-38, 3e38, -3e38, np.inf, -np.inf, np.nan, 1.0, -1.0, 1e-19, 1e19], np.float32)
    ex, ey = np.meshgrid(edge, edge)
    x[:ex.size] = ex.ravel()
    y[:ey.size] = ey.ravel()
    # unit-length style operands, the common case in normalize()
    x[1000:2_000_000] = rng.uniform(-1, 1, 2_000_000 - 1000).astype(np.float32)
    y[1000:2_000_000] = rng.uniform(0.05, 2, 2_000_000 - 1000).astype(np.float32)
    with np.errstate(all="ignore"):
        want = (x / y).astype(np.float32)
    got = g.pinned_math("vdiv", x, y)
    assert np.array_equal(bits(got), bits(want))
    a = (rng.uniform(-1, 1, 1_000_000) * 7.0).astype(np.float32)
    assert np.array_equal(bits(g.pinned_math("sincos_sin", a)), bits(pinned_math("sin", a)))
    assert np.array_equal(bits(g.pinned_math("sincos_cos", a)), bits(pinned_math("cos", a)))


def test_k4_stream_timing_hook_reports_time_and_bytes():
    """wgpt_k4_timing (the roofline hook of the accumulate / present passes) leaves binding 3 alone."""
    g = PathTracer(256, 128, mode=MODE_EXACT)
    before = np.asarray(g.read_framebuffer()).copy()
    for kernel, per_px in (("accumulate", 48), ("present_rgba8", 20), ("present_f32", 32)):
        ms, nbytes = g.k4_timing(kernel, reps=5)
        assert ms > 0.0 and nbytes == 256 * 128 * per_px
    assert np.array_equal(before, np.asarray(g.read_framebuffer()))
