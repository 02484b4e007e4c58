"""Restatement of the gl-matrix 3.x functions the reference's host code calls.

The reference imports gl-matrix from an UNPINNED CDN URL (/root/reference/lib/scene.js:1,
lib/camera.js:1, lib/transform.js:1, lib/primitives/quad.js:1, lib/primitives/triangle.js:3);
it is not vendored, so these are the library's published formulas restated from its public
documentation -- parity with the real library is UNPINNED.  gl-matrix
computes in JS doubles (= Python floats) and stores into Float32Array (`glMatrix.ARRAY_TYPE`);
`f32()` marks every such rounding point.  This code only manufactures the INPUT buffers of the
hot path; the hot path itself consumes the resulting f32 arrays as given data.
"""
import math
import numpy as np


def f32(x):
    """Round a Python float to binary32 and widen it back (a Float32Array store)."""
    return float(np.float32(x))


def f32v(v):
    return [f32(x) for x in v]


def mat4_create():
    return [1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0]


def mat4_from_translation(v):
    out = mat4_create()
    out[12], out[13], out[14] = f32(v[0]), f32(v[1]), f32(v[2])
    return out


def mat4_from_scaling(v):
    out = [0.0] * 16
    out[0], out[5], out[10], out[15] = f32(v[0]), f32(v[1]), f32(v[2]), 1.0
    return out


def mat4_from_rotation(rad, axis):
    x, y, z = axis
    ln = math.hypot(x, y, z)
    if ln < 0.000001:
        return None
    ln = 1.0 / ln
    x *= ln
    y *= ln
    z *= ln
    s, c = math.sin(rad), math.cos(rad)
    t = 1.0 - c
    out = [x * x * t + c, y * x * t + z * s, z * x * t - y * s, 0.0,
           x * y * t - z * s, y * y * t + c, z * y * t + x * s, 0.0,
           x * z * t + y * s, y * z * t - x * s, z * z * t + c, 0.0,
           0.0, 0.0, 0.0, 1.0]
    return f32v(out)


def mat4_mul(a, b):
    """out = a * b (column-major), rounded to f32 on store."""
    a00, a01, a02, a03, a10, a11, a12, a13, a20, a21, a22, a23, a30, a31, a32, a33 = a
    out = [0.0] * 16
    for col in range(4):
        b0, b1, b2, b3 = b[4 * col:4 * col + 4]
        out[4 * col + 0] = b0 * a00 + b1 * a10 + b2 * a20 + b3 * a30
        out[4 * col + 1] = b0 * a01 + b1 * a11 + b2 * a21 + b3 * a31
        out[4 * col + 2] = b0 * a02 + b1 * a12 + b2 * a22 + b3 * a32
        out[4 * col + 3] = b0 * a03 + b1 * a13 + b2 * a23 + b3 * a33
    return f32v(out)


def mat4_invert(a):
    a00, a01, a02, a03, a10, a11, a12, a13, a20, a21, a22, a23, a30, a31, a32, a33 = a
    b00 = a00 * a11 - a01 * a10
    b01 = a00 * a12 - a02 * a10
    b02 = a00 * a13 - a03 * a10
    b03 = a01 * a12 - a02 * a11
    b04 = a01 * a13 - a03 * a11
    b05 = a02 * a13 - a03 * a12
    b06 = a20 * a31 - a21 * a30
    b07 = a20 * a32 - a22 * a30
    b08 = a20 * a33 - a23 * a30
    b09 = a21 * a32 - a22 * a31
    b10 = a21 * a33 - a23 * a31
    b11 = a22 * a33 - a23 * a32
    det = b00 * b11 - b01 * b10 + b02 * b09 + b03 * b08 - b04 * b07 + b05 * b06
    if not det:
        return None
    det = 1.0 / det
    out = [(a11 * b11 - a12 * b10 + a13 * b09) * det,
           (a02 * b10 - a01 * b11 - a03 * b09) * det,
           (a31 * b05 - a32 * b04 + a33 * b03) * det,
           (a22 * b04 - a21 * b05 - a23 * b03) * det,
           (a12 * b08 - a10 * b11 - a13 * b07) * det,
           (a00 * b11 - a02 * b08 + a03 * b07) * det,
           (a32 * b02 - a30 * b05 - a33 * b01) * det,
           (a20 * b05 - a22 * b02 + a23 * b01) * det,
           (a10 * b10 - a11 * b08 + a13 * b06) * det,
           (a01 * b08 - a00 * b10 - a03 * b06) * det,
           (a30 * b04 - a31 * b02 + a33 * b00) * det,
           (a21 * b02 - a20 * b04 - a23 * b00) * det,
           (a11 * b07 - a10 * b09 - a12 * b06) * det,
           (a00 * b09 - a01 * b07 + a02 * b06) * det,
           (a31 * b01 - a30 * b03 - a32 * b00) * det,
           (a20 * b03 - a21 * b01 + a22 * b00) * det]
    return f32v(out)


def mat4_target_to(eye, target, up):
    """Camera-to-world matrix (lib/camera.js:32)."""
    eyex, eyey, eyez = eye
    upx, upy, upz = up
    z0, z1, z2 = eyex - target[0], eyey - target[1], eyez - target[2]
    ln = z0 * z0 + z1 * z1 + z2 * z2
    if ln > 0:
        ln = 1.0 / math.sqrt(ln)
        z0 *= ln
        z1 *= ln
        z2 *= ln
    x0, x1, x2 = upy * z2 - upz * z1, upz * z0 - upx * z2, upx * z1 - upy * z0
    ln = x0 * x0 + x1 * x1 + x2 * x2
    if ln > 0:
        ln = 1.0 / math.sqrt(ln)
        x0 *= ln
        x1 *= ln
        x2 *= ln
    out = [x0, x1, x2, 0.0,
           z1 * x2 - z2 * x1, z2 * x0 - z0 * x2, z0 * x1 - z1 * x0, 0.0,
           z0, z1, z2, 0.0,
           eyex, eyey, eyez, 1.0]
    return f32v(out)


def vec3_cross(a, b):
    return f32v([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def vec3_normalize(a):
    ln = a[0] * a[0] + a[1] * a[1] + a[2] * a[2]
    if ln > 0:
        ln = 1.0 / math.sqrt(ln)
    return f32v([a[0] * ln, a[1] * ln, a[2] * ln])


def vec3_dot(a, b):
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def transform_mat4_many(points, m):
    """vec3.transformMat4 over an (N,3) float64 array; result rounded to f32 (N,3) float64."""
    p = np.asarray(points, dtype=np.float64)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    w = m[3] * x + m[7] * y + m[11] * z + m[15]
    w = np.where(w == 0.0, 1.0, w)
    out = np.stack([(m[0] * x + m[4] * y + m[8] * z + m[12]) / w,
                    (m[1] * x + m[5] * y + m[9] * z + m[13]) / w,
                    (m[2] * x + m[6] * y + m[10] * z + m[14]) / w], axis=1)
    return out.astype(np.float32).astype(np.float64)
