"""A SECOND WITNESS for the oracle -- TEST INFRASTRUCTURE.

An independent float32 transliteration of the reference's shaders, written from the WGSL files themselves
(/root/reference/shaders/*.wgsl, cited per function) and NOT from oracle/*.cpp or the CUDA sources: numpy float32
scalars, one operation per WGSL operator, in WGSL's own shape (hitRec is a mutable global that every accepted hit
overwrites, scatterRec / doSpecular / the ONB vectors are globals, the BVH loop is the one of hitRay.wgsl).  What WGSL
leaves loose is taken from SURVEY.md §8(c) "pinned semantics" (left-to-right f32 sums, no contraction, mix = a(1-t)+bt,
normalize = v / length ...); sin / cos / acos / log / pow go through the shared pinned-math header's host hook
(include/wgpt_pinned_math.h via oracle.binding.pinned_math), the one piece both sides must share by construction.
tests/test_second_witness.py compares it with the oracle bit for bit: a misreading of the WGSL shared by the oracle and
the kernels (both written from one reading) would show up here.
"""
import numpy as np

from oracle.binding import pinned_math

f32 = np.float32
u32 = np.uint32
np.seterr(all="ignore")

PI = f32(3.1415926535897932385)
TWO_PI = f32(2 * 3.1415926535897932385)            # `2 * PI` folds in abstract float, then becomes f32
MIN_FLOAT = f32(0.0001)
MAX_FLOAT = f32(999999999.999)
MAX_FLOAT_P1 = f32(999999999.999 + 1)
LAMBERTIAN, MIRROR, GLASS, ISOTROPIC = 0, 1, 2, 3


def pm(fn, x, y=None):
    return f32(pinned_math(fn, np.array([x], np.float32), None if y is None else np.array([y], np.float32))[0])


# ---- vec3 as tuples of float32, WGSL operators spelled out ----
def V(x, y, z):
    return (f32(x), f32(y), f32(z))


def add(a, b): return (a[0] + b[0], a[1] + b[1], a[2] + b[2])
def sub(a, b): return (a[0] - b[0], a[1] - b[1], a[2] - b[2])
def mul(a, b): return (a[0] * b[0], a[1] * b[1], a[2] * b[2])
def scale(a, s): return (a[0] * s, a[1] * s, a[2] * s)
def divs(a, s): return (a[0] / s, a[1] / s, a[2] / s)
def neg(a): return (-a[0], -a[1], -a[2])
def dot(a, b): return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]
def cross(a, b): return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])
def length(a): return np.sqrt(dot(a, a))
def normalize(a): return divs(a, length(a))
def mix(a, b, t): return add(scale(a, f32(1) - t), scale(b, t))
def reflect(e1, e2): return sub(e1, scale(e2, f32(2) * dot(e2, e1)))


def refract(e1, e2, e3):
    d = dot(e2, e1)
    k = f32(1) - e3 * e3 * (f32(1) - d * d)
    if k < 0:
        return V(0, 0, 0)
    return sub(scale(e1, e3), scale(e2, e3 * d + np.sqrt(k)))


def fmin(a, b):
    if np.isnan(a):
        return b
    if np.isnan(b):
        return a
    if a == b:                                    # min(-0, +0) = -0
        return f32(np.array([np.array([a], np.float32).view(np.uint32)[0] | np.array([b], np.float32).view(np.uint32)[0]], np.uint32).view(np.float32)[0])
    return a if a < b else b


def fmax(a, b):
    if np.isnan(a):
        return b
    if np.isnan(b):
        return a
    if a == b:                                    # max(-0, +0) = +0
        return f32(np.array([np.array([a], np.float32).view(np.uint32)[0] & np.array([b], np.float32).view(np.uint32)[0]], np.uint32).view(np.float32)[0])
    return a if a > b else b


def mat_vec4(m, v):
    """mat4x4f * vec4f, m column-major 16 floats: ((c0 x + c1 y) + c2 z) + c3 w per row."""
    return tuple(((m[0 + r] * v[0] + m[4 + r] * v[1]) + m[8 + r] * v[2]) + m[12 + r] * v[3] for r in range(4))


def matT_vec4(m, v):
    """transpose(m) * v"""
    return tuple(((m[4 * r] * v[0] + m[4 * r + 1] * v[1]) + m[4 * r + 2] * v[2]) + m[4 * r + 3] * v[3] for r in range(4))


class HitRecord:
    def __init__(self):
        self.p = V(0, 0, 0)
        self.t = f32(0)
        self.normal = V(0, 0, 0)
        self.front_face = False
        self.material = None           # index into materials (header.wgsl:119-125 copies the struct; an index is the same data)
        self.prim = (-1, -1)           # (type, index) of the accepted primitive: the parity probe, not part of the WGSL
