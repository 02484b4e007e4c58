/*
 * oracle/ovec.h -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product path.
 *
 * f32 vector/matrix helpers with the operation order PINNED as SURVEY.md §8(c) "pinned
 * semantics" states (real WGSL leaves it loose): IEEE binary32 round-to-nearest, no FMA
 * contraction (compile with -ffp-contract=off), left-to-right sums.
 * WGSL built-ins restated here are used by /root/reference/shaders/ (all .wgsl files).
 */
#ifndef WGPT_ORACLE_OVEC_H
#define WGPT_ORACLE_OVEC_H
#include <math.h>
#include <string.h>

namespace orc {

struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct mat4 { float c[4][4]; };   /* c[col][row], column-major like WGSL mat4x4f */

static inline vec3 V3(float x, float y, float z) { vec3 r = {x, y, z}; return r; }
static inline vec3 V3s(float s) { return V3(s, s, s); }
static inline vec4 V4(vec3 a, float w) { vec4 r = {a.x, a.y, a.z, w}; return r; }
static inline vec3 xyz(vec4 a) { return V3(a.x, a.y, a.z); }

static inline vec3 operator+(vec3 a, vec3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline vec3 operator-(vec3 a, vec3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline vec3 operator*(vec3 a, vec3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline vec3 operator/(vec3 a, vec3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline vec3 operator*(vec3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline vec3 operator*(float s, vec3 a) { return V3(s * a.x, s * a.y, s * a.z); }
static inline vec3 operator/(vec3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
static inline vec3 operator/(float s, vec3 a) { return V3(s / a.x, s / a.y, s / a.z); }
static inline vec3 operator-(vec3 a) { return V3(-a.x, -a.y, -a.z); }

/* min/max with the result BITS pinned (WGSL leaves NaN and signed-zero handling loose):
 * NaN -> the other operand; min(-0,+0) = -0, max(-0,+0) = +0.  The CUDA side (csrc/pvec.cuh
 * pmin/pmax) implements the same rule. */
static inline float fmin_(float a, float b) {
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) { unsigned x, y; memcpy(&x, &a, 4); memcpy(&y, &b, 4); x |= y; float r; memcpy(&r, &x, 4); return r; }
    return a < b ? a : b;
}
static inline float fmax_(float a, float b) {
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) { unsigned x, y; memcpy(&x, &a, 4); memcpy(&y, &b, 4); x &= y; float r; memcpy(&r, &x, 4); return r; }
    return a > b ? a : b;
}
static inline vec3 vmin(vec3 a, vec3 b) { return V3(fmin_(a.x, b.x), fmin_(a.y, b.y), fmin_(a.z, b.z)); }
static inline vec3 vmax(vec3 a, vec3 b) { return V3(fmax_(a.x, b.x), fmax_(a.y, b.y), fmax_(a.z, b.z)); }

static inline float dot(vec3 a, vec3 b) { return ((a.x * b.x) + (a.y * b.y)) + (a.z * b.z); }
static inline float dot4(vec4 a, vec4 b) { return (((a.x * b.x) + (a.y * b.y)) + (a.z * b.z)) + (a.w * b.w); }
static inline vec3 cross(vec3 a, vec3 b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(vec3 a) { return sqrtf(dot(a, a)); }
static inline vec3 normalize(vec3 a) { return a / length(a); }
static inline vec4 normalize4(vec4 a) {
    const float l = sqrtf(dot4(a, a));
    vec4 r = {a.x / l, a.y / l, a.z / l, a.w / l};
    return r;
}
/* mix(a,b,t) = a*(1-t) + b*t */
static inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
/* reflect(e1,e2) = e1 - 2*dot(e2,e1)*e2 */
static inline vec3 reflect(vec3 e1, vec3 e2) { return e1 - (2.0f * dot(e2, e1)) * e2; }
/* refract(e1,e2,e3): k = 1 - e3^2 (1 - dot(e2,e1)^2); k<0 -> 0 else e3*e1 - (e3*dot(e2,e1)+sqrt(k))*e2 */
static inline vec3 refract(vec3 e1, vec3 e2, float e3) {
    const float d = dot(e2, e1);
    const float k = 1.0f - (e3 * e3) * (1.0f - d * d);
    if (k < 0.0f) return V3s(0.0f);
    return e3 * e1 - (e3 * d + sqrtf(k)) * e2;
}
/* M * v = ((c0*x + c1*y) + c2*z) + c3*w */
static inline vec4 mul(const mat4& m, vec4 v) {
    vec4 r;
    r.x = ((m.c[0][0] * v.x + m.c[1][0] * v.y) + m.c[2][0] * v.z) + m.c[3][0] * v.w;
    r.y = ((m.c[0][1] * v.x + m.c[1][1] * v.y) + m.c[2][1] * v.z) + m.c[3][1] * v.w;
    r.z = ((m.c[0][2] * v.x + m.c[1][2] * v.y) + m.c[2][2] * v.z) + m.c[3][2] * v.w;
    r.w = ((m.c[0][3] * v.x + m.c[1][3] * v.y) + m.c[2][3] * v.z) + m.c[3][3] * v.w;
    return r;
}
/* transpose(M) * v : component i = dot4-style sum over column i of M */
static inline vec4 mulT(const mat4& m, vec4 v) {
    vec4 r;
    r.x = ((m.c[0][0] * v.x + m.c[0][1] * v.y) + m.c[0][2] * v.z) + m.c[0][3] * v.w;
    r.y = ((m.c[1][0] * v.x + m.c[1][1] * v.y) + m.c[1][2] * v.z) + m.c[1][3] * v.w;
    r.z = ((m.c[2][0] * v.x + m.c[2][1] * v.y) + m.c[2][2] * v.z) + m.c[2][3] * v.w;
    r.w = ((m.c[3][0] * v.x + m.c[3][1] * v.y) + m.c[3][2] * v.z) + m.c[3][3] * v.w;
    return r;
}
static inline float comp(vec3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

}  /* namespace orc */
#endif
