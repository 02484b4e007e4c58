// shade.cuh -- hit finalisation, sampling helpers, BSDF sampling and the per-bounce update.
// The function bodies live in shade_body.inc and are compiled twice: against the pinned arithmetic (namespace wg: what
// every bit-exact path uses) and against fast arithmetic (namespace wg::tol: WGPT_MODE_FAST_TOL, opt-in).  Restates /root/reference/shaders/importanceSampling.wgsl:1-16,35-45,56-125,
// scatterRay.wgsl:1-95 and the loop body of traceRay.wgsl:10-80.
#pragma once
#include "intersect.cuh"

namespace wg {
#include "shade_body.inc"
}  // namespace wg

// ---------------------------------------------------------------------------------------------------------------
// wg::tol -- the SAME shading code on fast arithmetic (WGPT_MODE_FAST_TOL, include/wgpt.h).  north_star asks for
// bit-exact PRIMARY hits and 1e-4 relative error on per-sample radiance, not for bit-exact everything: in this mode the
// camera ray, every intersection (spheres, quads, BVH, triangles) and the RNG stream stay pinned, and only the shading
// of a hit (normal, BSDF sample, pdfs, throughput) runs on: plain * + - (ptxas contracts them into FMAs), MUFU
// reciprocal / reciprocal square root with one Newton step, and CUDA's f32 libm (sincosf, acosf, logf: <= 2 ulp) instead
// of the binary64-evaluated pinned header.  Names inside this namespace shadow the pinned ones of namespace wg, so
// shade_body.inc compiles unchanged.
namespace wg {
namespace tol {

__device__ __forceinline__ float fm(float a, float b) { return a * b; }
__device__ __forceinline__ float fa(float a, float b) { return a + b; }
__device__ __forceinline__ float fs(float a, float b) { return a - b; }
__device__ __forceinline__ float frcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r * (2.0f - x * r); }
__device__ __forceinline__ float fd(float a, float b) { return a * frcp(b); }
__device__ __forceinline__ float fsq(float a) { return sqrtf(a); }

struct v3 { float x, y, z; };
__device__ __forceinline__ v3 V3(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ v3 V3s(float s) { return V3(s, s, s); }
__device__ __forceinline__ v3 operator+(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ v3 operator-(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ v3 operator*(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ v3 operator*(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ v3 operator*(float s, v3 a) { return V3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ v3 operator/(v3 a, float s) { const float r = frcp(s); return V3(a.x * r, a.y * r, a.z * r); }
__device__ __forceinline__ v3 operator-(v3 a) { return V3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float pmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ float pmax(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ v3 cross(v3 a, v3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float length(v3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ v3 normalize(v3 a) {
    const float q = dot(a, a);
    float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(q));
    r = r * (1.5f - 0.5f * q * r * r);                       // one Newton step: ~1 ulp
    return V3(a.x * r, a.y * r, a.z * r);
}
__device__ __forceinline__ v3 mix(v3 a, v3 b, float t) { return a * (1.0f - t) + b * t; }
__device__ __forceinline__ v3 reflect(v3 e1, v3 e2) { return e1 - (2.0f * dot(e2, e1)) * e2; }
__device__ __forceinline__ v3 refract(v3 e1, v3 e2, float e3) {
    const float d = dot(e2, e1);
    const float k = 1.0f - e3 * e3 * (1.0f - d * d);
    if (k < 0.0f) return V3s(0.0f);
    return e3 * e1 - (e3 * d + sqrtf(k)) * e2;
}
__device__ __forceinline__ v3 mulT4_xyz0(const float* __restrict__ m, v3 v) {
    return V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
struct Ray { v3 o, d; };
__device__ __forceinline__ v3 ld3(const float* p) { return V3(p[0], p[1], p[2]); }
__device__ __forceinline__ v3 at(const Ray& r, float t) { return r.o + t * r.d; }

namespace wgpt_pm {       // shadows ::wgpt_pm (the pinned, binary64-evaluated header) for the code of this namespace
__device__ __forceinline__ void sincosf_(float x, float* s, float* c) { ::sincosf(x, s, c); }
__device__ __forceinline__ float acosf_(float x) { return ::acosf(x); }
__device__ __forceinline__ float logf_(float x) { return ::logf(x); }
__device__ __forceinline__ float powf_(float x, float y) {
    if (y == 5.0f) { const float x2 = x * x; return x2 * x2 * x; }        // reflectance(): (1 - cos)^5
    if (y == 2.0f) return x * x;                                          // Henyey-Greenstein
    return ::powf(x, y);
}
}  // namespace wgpt_pm

#include "shade_body.inc"

// pinned <-> fast value types: same bits
__device__ __forceinline__ v3 from_pinned(wg::v3 a) { return V3(a.x, a.y, a.z); }
__device__ __forceinline__ wg::v3 to_pinned(v3 a) { return wg::V3(a.x, a.y, a.z); }

}  // namespace tol

// One bounce AFTER a closest hit, either flavour: finalize the hit, sample the BSDF, update radiance / throughput
// (traceRay.wgsl:18-79).  Returns false when the path ends; `ray` becomes the scattered ray.
template <bool SPH, int IS, bool TOL>
__device__ __forceinline__ bool shade_bounce(const DeviceScene& S, const HitLite& h, const TriNormals& tn, int bounce, Ray& ray,
                                             uint32_t& rng, unsigned& n_draws, v3& acc, v3& thr) {
    // (the material is loaded AFTER finalize_hit in both flavours: loading it first costs registers across the normal
    //  computation -- measured, shade +1 %)
    bool early;
    if (TOL) {
        tol::Ray r; r.o = tol::from_pinned(ray.o); r.d = tol::from_pinned(ray.d);
        tol::ShadeState st;
        st.p = st.n = st.unit_w = st.u = st.v = tol::V3s(0.0f);
        st.front_face = false; st.skip_pdf = false; st.doSpecular = 0.0f;
        st.skip_ray.o = st.skip_ray.d = tol::V3s(0.0f);
        tol::TriNormals t2; t2.a = tn.a; t2.b = tn.b; t2.c = tn.c;
        tol::finalize_hit<SPH>(S, r, h, t2, st);
        const Material m = S.materials[h.mat_id];
        tol::v3 a = tol::from_pinned(acc), t = tol::from_pinned(thr);
        const bool cont = tol::bounce_update<IS>(S, m, bounce, r, st, rng, n_draws, a, t, early);
        acc = tol::to_pinned(a); thr = tol::to_pinned(t);
        ray.o = tol::to_pinned(r.o); ray.d = tol::to_pinned(r.d);
        return cont;
    }
    ShadeState st;
    st.p = st.n = st.unit_w = st.u = st.v = V3s(0.0f);
    st.front_face = false; st.skip_pdf = false; st.doSpecular = 0.0f;
    st.skip_ray.o = st.skip_ray.d = V3s(0.0f);
    finalize_hit<SPH>(S, ray, h, tn, st);
    const Material m = S.materials[h.mat_id];
    return bounce_update<IS>(S, m, bounce, ray, st, rng, n_draws, acc, thr, early);
}

}  // namespace wg
