// shade.cuh -- hit finalisation, sampling helpers, BSDF sampling and the per-bounce update,
// pinned arithmetic.  Restates /root/reference/shaders/importanceSampling.wgsl:1-16,35-45,56-125,
// scatterRay.wgsl:1-95 and the loop body of traceRay.wgsl:10-80.
#pragma once
#include "intersect.cuh"

namespace wg {

// The reference's private (per-invocation) shading state: HitRecord (header.wgsl:119-125),
// ScatterRecord (:127-131), doSpecular (scatterRay.wgsl:1), ONB (importanceSampling.wgsl:56-58).
struct ShadeState {
    v3 p, n;             // hitRec.p, hitRec.normal
    bool front_face;
    bool skip_pdf;
    Ray skip_ray;
    float doSpecular;
    v3 unit_w, u, v;
};

// The three 16-byte records of a Triangle that hold its vertex normals (+ local_id, mesh_id in the
// w lanes: header.wgsl:88-98, offsets 48 / 64 / 80), fetched ahead of the shading code.
struct TriNormals { float4 a, b, c; };
__device__ __forceinline__ TriNormals load_tri_normals(const DeviceScene& S, int index) {
    const float4* tp = reinterpret_cast<const float4*>(&S.triangles[index]);
    TriNormals t; t.a = __ldg(&tp[3]); t.b = __ldg(&tp[4]); t.c = __ldg(&tp[5]);
    return t;
}

// What hit_sphere / hit_volume / hit_quad / hit_triangle store into hitRec on accept
// (common.wgsl:54-71, 140-143, 176-185, 224-239), computed once for the closest hit.  SPH = false: no spheres exist.
template <bool SPH = true>
__device__ __forceinline__ void finalize_hit(const DeviceScene& S, const Ray& ray, const HitLite& h,
                                             const TriNormals& tn, ShadeState& st) {
    st.p = at(ray, h.t);
    if (SPH && h.kind == HIT_SPHERE) {
        const Sphere s = S.spheres[h.index];
        st.n = normalize((st.p - V3(s.cx, s.cy, s.cz)) / s.r);
        st.front_face = dot(ray.d, st.n) < 0.0f;
        if (!st.front_face) st.n = -st.n;
    } else if (SPH && h.kind == HIT_VOLUME) {
        const Sphere s = S.spheres[h.index];
        st.n = normalize(st.p - V3(s.cx, s.cy, s.cz));
        st.front_face = true;
    } else if (h.kind == HIT_QUAD) {
        st.n = normalize(ld3(S.quads[h.index].normal));
        st.front_face = dot(ray.d, st.n) < 0.0f;
        if (!st.front_face) st.n = -st.n;
    } else {
        const float ww = fs(fs(1.0f, h.uu), h.vv);
        const v3 nn = (V3(tn.a.x, tn.a.y, tn.a.z) * ww + V3(tn.b.x, tn.b.y, tn.b.z) * h.uu) + V3(tn.c.x, tn.c.y, tn.c.z) * h.vv;
        const Mesh m = S.meshes[(int)tn.c.w];
        st.n = normalize(mulT4_xyz0(S.transforms[m.global_id].invModelMatrix, nn));
        st.front_face = dot(ray.d, st.n) < 0.0f;
        if (!st.front_face) st.n = -st.n;
    }
}

// importanceSampling.wgsl:1-5
__device__ __forceinline__ float reflectance(float cosine, float ref_idx) {
    float r0 = fd(fs(1.0f, ref_idx), fa(1.0f, ref_idx));
    r0 = fm(r0, r0);
    return fa(r0, fm(fs(1.0f, r0), wgpt_pm::powf_(fs(1.0f, cosine), 5.0f)));
}

// importanceSampling.wgsl:7-16 (`rand2D() * 2.0 * PI` = (r*2)*f32(PI))
__device__ __forceinline__ v3 uniform_random_in_unit_sphere(uint32_t& rng) {
    const float phi = fm(fm(rand2D(rng), 2.0f), WGPT_PI_F);
    const float theta = wgpt_pm::acosf_(fs(fm(2.0f, rand2D(rng)), 1.0f));
    float st, ct, sp, cp;
    wgpt_pm::sincosf_(theta, &st, &ct);
    wgpt_pm::sincosf_(phi, &sp, &cp);
    const float x = fm(st, cp);
    const float y = fm(st, sp);
    const float z = ct;
    return normalize(V3(x, y, z));
}

// importanceSampling.wgsl:35-45 (`2 * PI * r1` = f32(2*PI)*r1)
__device__ __forceinline__ v3 cosine_sampling_wrt_Z(uint32_t& rng) {
    const float r1 = rand2D(rng);
    const float r2 = rand2D(rng);
    const float phi = fm(WGPT_TWO_PI_F, r1);
    const float sr2 = fsq(r2);
    float sp, cp;
    wgpt_pm::sincosf_(phi, &sp, &cp);
    return V3(fm(cp, sr2), fm(sp, sr2), fsq(fs(1.0f, r2)));
}

// importanceSampling.wgsl:60-71
__device__ __forceinline__ void onb_build_from_w(ShadeState& st, v3 w) {
    st.unit_w = normalize(w);
    const v3 a = (fabsf(st.unit_w.x) > 0.9f) ? V3(0, 1, 0) : V3(1, 0, 0);
    st.v = normalize(cross(st.unit_w, a));
    st.u = cross(st.unit_w, st.v);
}
__device__ __forceinline__ v3 onb_get_local(const ShadeState& st, v3 a) {
    return (st.u * a.x + st.v * a.y) + st.unit_w * a.z;
}

// importanceSampling.wgsl:88-125; the area term reads the global `lights` (:124) == S.lights
__device__ __forceinline__ float light_pdf(const DeviceScene& S, const Ray& ray) {
    const Quad& q = S.lights;
    const v3 qn = ld3(q.normal);
    if (dot(ray.d, qn) > 0.0f) return WGPT_MIN_FLOAT;
    const float denom = dot(qn, ray.d);
    if (fabsf(denom) < 1e-8f) return WGPT_MIN_FLOAT;
    const float t = fd(fs(q.D, dot(qn, ray.o)), denom);
    if (t <= 0.001f || t >= WGPT_MAX_FLOAT) return WGPT_MIN_FLOAT;
    const v3 planar = at(ray, t) - ld3(q.Q);
    const float alpha = dot(ld3(q.w), cross(planar, ld3(q.v)));
    const float beta = dot(ld3(q.w), cross(ld3(q.u), planar));
    if (alpha < 0.0f || 1.0f < alpha || beta < 0.0f || 1.0f < beta) return WGPT_MIN_FLOAT;
    v3 hitNormal = qn;
    if (!(dot(ray.d, qn) < 0.0f)) hitNormal = -hitNormal;
    const float len = length(ray.d);
    const float distance_squared = fm(fm(fm(t, t), len), len);
    const float cosine = fabsf(fd(dot(ray.d, hitNormal), len));
    return fd(distance_squared, fm(cosine, length(cross(ld3(q.u), ld3(q.v)))));
}

// scatterRay.wgsl:2-95.  Returns the scattered ray; updates doSpecular / skip_pdf / ONB state.
__device__ __forceinline__ Ray material_scatter(const Material& m, const Ray& ray_in, ShadeState& st,
                                                uint32_t& rng, unsigned& draws) {
    Ray scattered; scattered.o = V3s(0.0f); scattered.d = V3s(0.0f);
    st.doSpecular = 0.0f;
    if (m.material_type == 0.0f) {                                   // LAMBERTIAN (+ glossy lobe)
        onb_build_from_w(st, st.n);
        v3 diffuse_dir = cosine_sampling_wrt_Z(rng);
        diffuse_dir = normalize(onb_get_local(st, diffuse_dir));
        st.doSpecular = (rand2D(rng) < m.specularStrength) ? 1.0f : 0.0f;
        draws += 3;
        v3 specular_dir = reflect(ray_in.d, st.n);
        specular_dir = normalize(mix(specular_dir, diffuse_dir, m.roughness));
        scattered.o = st.p;
        scattered.d = normalize(mix(diffuse_dir, specular_dir, st.doSpecular));
        st.skip_pdf = false;
        if (st.doSpecular == 1.0f) { st.skip_pdf = true; st.skip_ray = scattered; }
    } else if (m.material_type == 1.0f) {                            // MIRROR
        const v3 reflected = reflect(ray_in.d, st.n);
        scattered.o = st.p;
        scattered.d = normalize(reflected + m.roughness * uniform_random_in_unit_sphere(rng));
        draws += 2;
        st.skip_pdf = true; st.skip_ray = scattered;
    } else if (m.material_type == 2.0f) {                            // GLASS
        float ir = m.eta;
        if (st.front_face) ir = fd(1.0f, ir);
        const v3 unit_direction = normalize(ray_in.d);
        const float cos_theta = pmin(dot(-unit_direction, st.n), 1.0f);
        const float sin_theta = fsq(fs(1.0f, fm(cos_theta, cos_theta)));
        bool refl = fm(ir, sin_theta) > 1.0f;
        if (!refl) { draws++; refl = reflectance(cos_theta, ir) > rand2D(rng); }   // short-circuit ||
        const v3 direction = refl ? reflect(unit_direction, st.n) : refract(unit_direction, st.n, ir);
        scattered.o = st.p;
        scattered.d = normalize(direction);
        st.skip_pdf = true; st.skip_ray = scattered;
    } else if (m.material_type == 3.0f) {                            // ISOTROPIC (Henyey-Greenstein)
        const float g = m.specularStrength;
        const float ratio = fd(fs(1.0f, fm(g, g)), fa(fs(1.0f, g), fm(fm(2.0f, g), rand2D(rng))));
        const float cos_hg = fd(fs(fa(1.0f, fm(g, g)), wgpt_pm::powf_(ratio, 2.0f)), fm(2.0f, g));
        const float sin_hg = fsq(fs(1.0f, fm(cos_hg, cos_hg)));
        const float phi = fm(WGPT_TWO_PI_F, rand2D(rng));
        draws += 2;
        float sp, cp;
        wgpt_pm::sincosf_(phi, &sp, &cp);
        const v3 hg_dir = V3(fm(sin_hg, cp), fm(sin_hg, sp), cos_hg);
        onb_build_from_w(st, ray_in.d);
        scattered.o = st.p;
        scattered.d = normalize(onb_get_local(st, hg_dir));
        st.skip_pdf = true; st.skip_ray = scattered;
    }
    return scattered;
}

// One iteration of the bounce loop AFTER a hit (traceRay.wgsl:18-79): emission, scatter,
// mixture-pdf weighting, Russian roulette.  Returns false when the path ends here.
// `early_return` is set when traceRay.wgsl:51-53 fires (the path's radiance is REPLACED by
// emission * throughput).  IS: -1 = header.wgsl:12 IMPORTANCE_SAMPLING read at run time, 0 / 1 = known to the caller
// (the other branch is not compiled in).
template <int IS = -1>
__device__ __forceinline__ bool bounce_update(const DeviceScene& S, const Material& m, int i, Ray& ray,
                                              ShadeState& st, uint32_t& rng, unsigned& draws,
                                              v3& acc_radiance, v3& throughput, bool& early_return) {
    early_return = false;
    v3 emission = ld3(m.emissionColor);
    if (!st.front_face) emission = V3s(0.0f);

    const bool importance_sampling = IS < 0 ? S.K.IMPORTANCE_SAMPLING != 0 : IS != 0;
    if (importance_sampling) {
        const Ray scattered_surface = material_scatter(m, ray, st, rng, draws);
        if (st.skip_pdf) {
            acc_radiance = acc_radiance + emission * throughput;
            throughput = throughput * mix(ld3(m.color), ld3(m.specularColor), st.doSpecular);
            ray = st.skip_ray;
            return true;                                     // `continue`: no Russian roulette
        }
        // get_random_on_quad, importanceSampling.wgsl:78-81 (u coefficient drawn first)
        const float ru = rand2D(rng);
        const v3 pu = ld3(S.lights.Q) + ru * ld3(S.lights.u);
        const float rv = rand2D(rng);
        const v3 pl = pu + rv * ld3(S.lights.v);
        Ray scattered; scattered.o = st.p; scattered.d = normalize(pl - st.p);
        const float rnd = rand2D(rng);
        draws += 3;
        if (rnd > 0.2f) scattered = scattered_surface;

        const float lambertian_pdf = pmax(0.0f, fd(dot(normalize(scattered.d), st.unit_w), WGPT_PI_F));
        const float lpdf = light_pdf(S, scattered);
        const float pdf = fa(fm(0.2f, lpdf), fm(0.8f, lambertian_pdf));
        if (pdf <= 0.00001f) { acc_radiance = emission * throughput; early_return = true; return false; }

        acc_radiance = acc_radiance + emission * throughput;
        throughput = throughput * ((lambertian_pdf * mix(ld3(m.color), ld3(m.specularColor), st.doSpecular)) / pdf);
        ray = scattered;
    } else {
        const Ray scattered = material_scatter(m, ray, st, rng, draws);
        acc_radiance = acc_radiance + emission * throughput;
        throughput = throughput * mix(ld3(m.color), ld3(m.specularColor), st.doSpecular);
        ray = scattered;
    }

    if (i > 2) {                                             // Russian roulette, traceRay.wgsl:71-79
        const float p = pmax(throughput.x, pmax(throughput.y, throughput.z));
        draws++;
        if (rand2D(rng) > p) return false;
        throughput = throughput * fd(1.0f, p);
    }
    return true;
}

}  // namespace wg
