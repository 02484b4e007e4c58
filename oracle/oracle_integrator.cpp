/*
 * oracle/oracle_integrator.cpp -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product path.
 * Restatement of importanceSampling.wgsl, scatterRay.wgsl, traceRay.wgsl, shootRay.wgsl and
 * main.wgsl of the reference; see oracle_shader.h.  PARITY UNPINNED (no reference tests).
 */
#include "oracle_shader.h"

namespace orc {

using wgpt_pm::sinf_; using wgpt_pm::cosf_; using wgpt_pm::acosf_; using wgpt_pm::powf_;

/* importanceSampling.wgsl:1-5 */
float Invocation::reflectance(float cosine, float ref_idx) {
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    return r0 + (1.0f - r0) * powf_((1.0f - cosine), 5.0f);
}

/* importanceSampling.wgsl:7-16 -- `rand2D() * 2.0 * PI` is (r*2)*f32(PI) */
vec3 Invocation::uniform_random_in_unit_sphere() {
    const float phi = (rand2D() * 2.0f) * PI_F;
    const float theta = acosf_(2.0f * rand2D() - 1.0f);
    const float x = sinf_(theta) * cosf_(phi);
    const float y = sinf_(theta) * sinf_(phi);
    const float z = cosf_(theta);
    return normalize(V3(x, y, z));
}

/* importanceSampling.wgsl:35-45 -- `2 * PI * r1` is f32(2*PI)*r1 */
vec3 Invocation::cosine_sampling_wrt_Z() {
    const float r1 = rand2D();
    const float r2 = rand2D();
    const float phi = TWO_PI_F * r1;
    const float x = cosf_(phi) * sqrtf(r2);
    const float y = sinf_(phi) * sqrtf(r2);
    const float z = sqrtf(1.0f - r2);
    return V3(x, y, z);
}

/* importanceSampling.wgsl:60-67 */
void Invocation::onb_build_from_w(vec3 w) {
    unit_w = normalize(w);
    const vec3 a = (fabsf(unit_w.x) > 0.9f) ? V3(0, 1, 0) : V3(1, 0, 0);
    v = normalize(cross(unit_w, a));
    u = cross(unit_w, v);
}

/* importanceSampling.wgsl:73-76 */
float Invocation::onb_lambertian_scattering_pdf(const Ray& scattered) {
    const float cosine_theta = dot(normalize(scattered.dir), unit_w);
    return fmax_(0.0f, cosine_theta / PI_F);
}

/* importanceSampling.wgsl:78-81 -- left-to-right evaluation: the u coefficient is drawn first */
Ray Invocation::get_random_on_quad(const Quad& q, vec3 origin) {
    const float ru = rand2D();
    const vec3 pu = v3(q.Q) + (ru * v3(q.u));
    const float rv = rand2D();
    const vec3 p = pu + (rv * v3(q.v));
    Ray r = {origin, normalize(p - origin)};
    return r;
}

/* importanceSampling.wgsl:88-125 -- the area term uses the global `lights` (:124) */
float Invocation::light_pdf(const Ray& ray, const Quad& quad) {
    const vec3 qn = v3(quad.normal);
    if (dot(ray.dir, qn) > 0) return MIN_FLOAT;
    const float denom = dot(qn, ray.dir);
    if (fabsf(denom) < 1e-8f) return MIN_FLOAT;
    const float t = (quad.D - dot(qn, ray.origin)) / denom;
    if (t <= 0.001f || t >= MAX_FLOAT) return MIN_FLOAT;

    const vec3 intersection = at(ray, t);
    const vec3 planar_hitpt_vector = intersection - v3(quad.Q);
    const float alpha = dot(v3(quad.w), cross(planar_hitpt_vector, v3(quad.v)));
    const float beta = dot(v3(quad.w), cross(v3(quad.u), planar_hitpt_vector));
    if (alpha < 0 || 1 < alpha || beta < 0 || 1 < beta) return MIN_FLOAT;

    vec3 hitNormal = qn;
    const bool front_face = dot(ray.dir, qn) < 0;
    if (front_face == false) hitNormal = -hitNormal;

    const float distance_squared = ((t * t) * length(ray.dir)) * length(ray.dir);
    const float cosine = fabsf(dot(ray.dir, hitNormal) / length(ray.dir));
    return distance_squared / (cosine * length(cross(v3(lights.u), v3(lights.v))));
}

/* scatterRay.wgsl:2-95 */
Ray Invocation::material_scatter(const Ray& ray_in) {
    Ray scattered = {V3s(0), V3s(0)};
    doSpecular = 0;
    const Material m = mat();

    if (m.material_type == LAMBERTIAN) {
        onb_build_from_w(hit_normal);
        vec3 diffuse_dir = cosine_sampling_wrt_Z();
        diffuse_dir = normalize(onb_get_local(diffuse_dir));
        doSpecular = (rand2D() < m.specularStrength) ? 1.0f : 0.0f;

        vec3 specular_dir = reflect(ray_in.dir, hit_normal);
        specular_dir = normalize(mix(specular_dir, diffuse_dir, m.roughness));
        scattered.origin = hit_p;
        scattered.dir = normalize(mix(diffuse_dir, specular_dir, doSpecular));

        skip_pdf = false;
        if (doSpecular == 1.0f) { skip_pdf = true; skip_pdf_ray = scattered; }
    }
    else if (m.material_type == MIRROR) {
        const vec3 reflected = reflect(ray_in.dir, hit_normal);
        scattered.origin = hit_p;
        scattered.dir = normalize(reflected + m.roughness * uniform_random_in_unit_sphere());
        skip_pdf = true; skip_pdf_ray = scattered;
    }
    else if (m.material_type == GLASS) {
        float ir = m.eta;
        if (hit_front_face == true) ir = (1.0f / ir);

        const vec3 unit_direction = normalize(ray_in.dir);
        const float cos_theta = fmin_(dot(-unit_direction, hit_normal), 1.0f);
        const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);

        vec3 direction;
        /* short-circuit ||: no draw on total internal reflection (scatterRay.wgsl:55) */
        if (ir * sin_theta > 1.0f || reflectance(cos_theta, ir) > rand2D())
            direction = reflect(unit_direction, hit_normal);
        else
            direction = refract(unit_direction, hit_normal, ir);
        /* near_zero() is always false (common.wgsl:25-27) */
        scattered.origin = hit_p;
        scattered.dir = normalize(direction);
        skip_pdf = true; skip_pdf_ray = scattered;
    }
    else if (m.material_type == ISOTROPIC) {
        const float g = m.specularStrength;
        const float ratio = (1.0f - g * g) / ((1.0f - g) + (2.0f * g) * rand2D());
        const float cos_hg = ((1.0f + g * g) - powf_(ratio, 2.0f)) / (2.0f * g);
        const float sin_hg = sqrtf(1.0f - cos_hg * cos_hg);
        const float phi = TWO_PI_F * rand2D();
        const vec3 hg_dir = V3(sin_hg * cosf_(phi), sin_hg * sinf_(phi), cos_hg);
        onb_build_from_w(ray_in.dir);
        scattered.origin = hit_p;
        scattered.dir = normalize(onb_get_local(hg_dir));
        skip_pdf = true; skip_pdf_ray = scattered;
    }
    return scattered;
}

/* traceRay.wgsl:3-83 */
vec3 Invocation::ray_color(const Ray& incidentRay) {
    Ray currRay = incidentRay;
    vec3 acc_radiance = V3s(0);
    vec3 throughput = V3s(1);
    const vec3 background_color = V3(0, 1, 1);

    for (int i = 0; i < K.MAX_BOUNCES; i++) {
        const bool hit = hitScene(currRay);
        if (!primary_done) {
            primary_done = true;
            primary_type = hit ? hit_type : -1;
            primary_index = hit ? hit_index : -1;
            primary_t = hit ? hit_t : 0.0f;
        }
        if (hit == false) {
            acc_radiance = acc_radiance + (background_color * throughput);
            break;
        }
        const Material m = mat();
        vec3 emissionColor = v3(m.emissionColor);
        if (!hit_front_face) emissionColor = V3s(0);

        if (K.IMPORTANCE_SAMPLING) {
            const Ray scatterred_surface = material_scatter(currRay);
            if (skip_pdf) {
                acc_radiance = acc_radiance + emissionColor * throughput;
                throughput = throughput * mix(v3(m.color), v3(m.specularColor), doSpecular);
                currRay = skip_pdf_ray;
                continue;                      /* skips Russian roulette (traceRay.wgsl:35) */
            }
            const Ray scattered_light = get_random_on_quad(lights, hit_p);
            Ray scattered = scattered_light;
            const float rnd = rand2D();
            if (rnd > 0.2f) scattered = scatterred_surface;

            const float lambertian_pdf = onb_lambertian_scattering_pdf(scattered);
            const float lpdf = light_pdf(scattered, lights);
            const float pdf = 0.2f * lpdf + 0.8f * lambertian_pdf;
            if (pdf <= 0.00001f) return emissionColor * throughput;

            acc_radiance = acc_radiance + emissionColor * throughput;
            throughput = throughput * ((lambertian_pdf * mix(v3(m.color), v3(m.specularColor), doSpecular)) / pdf);
            currRay = scattered;
        } else {
            const Ray scattered = material_scatter(currRay);
            acc_radiance = acc_radiance + emissionColor * throughput;
            throughput = throughput * mix(v3(m.color), v3(m.specularColor), doSpecular);
            currRay = scattered;
        }

        if (i > 2) {
            const float p = fmax_(throughput.x, fmax_(throughput.y, throughput.z));
            if (rand2D() > p) break;
            throughput = throughput * (1.0f / p);
        }
    }
    return acc_radiance;
}

/* shootRay.wgsl:54-60 -- the vec4 (w = 0) is normalised, then .xyz */
Ray Invocation::getCameraRay(float s, float t) {
    const vec4 d = normalize4(mul(S.uniforms.viewMatrix, V4(V3(s, t, -fovFactor), 0.0f)));
    Ray r = {cam_origin, xyz(d)};
    return r;
}

/* shootRay.wgsl:5-49 */
vec3 Invocation::pathTrace() {
    vec3 pixColor = V3s(0);
    const float W = S.uniforms.screenW, H = S.uniforms.screenH;

    if (K.STRATIFY) {
        const float sqrt_spp = (float)sqrt((double)K.NUM_SAMPLES);
        const float recip_sqrt_spp = 1.0f / (float)((int)sqrt_spp);
        float numSamples = 0.0f;
        for (float i = 0.0f; i < sqrt_spp; i += 1.0f) {
            for (float j = 0.0f; j < sqrt_spp; j += 1.0f) {
                const float rx = rand2D();
                const float sx = (W / H) * (2.0f * (((pixelCoords.x - 0.5f) + (recip_sqrt_spp * (i + rx))) / W) - 1.0f);
                const float ry = rand2D();
                const float sy = -1.0f * (2.0f * (((pixelCoords.y - 0.5f) + (recip_sqrt_spp * (j + ry))) / H) - 1.0f);
                const Ray ray = getCameraRay(sx, sy);
                pixColor = pixColor + ray_color(ray);
                cnt.samples++;
                numSamples += 1.0f;
            }
        }
        pixColor = pixColor / numSamples;
    } else {
        for (int i = 0; i < K.NUM_SAMPLES; i++) {
            const float rx = rand2D();
            const float sx = (W / H) * (2.0f * (((pixelCoords.x - 0.5f) + rx) / W) - 1.0f);
            const float ry = rand2D();
            const float sy = -1.0f * (2.0f * (((pixelCoords.y - 0.5f) + ry) / H) - 1.0f);
            const Ray ray = getCameraRay(sx, sy);
            pixColor = pixColor + ray_color(ray);
            cnt.samples++;
        }
        pixColor = pixColor / (float)K.NUM_SAMPLES;
    }
    return pixColor;
}

/* main.wgsl:1-28, minus the framebuffer read-modify-write (done by the caller).
 * Q1: pixelCoords = (f32(idx) % W, f32(idx) / W) -- float remainder, unfloored quotient. */
vec3 Invocation::computeFrameBuffer(uint32_t pixelIndex) {
    const float W = S.uniforms.screenW;
    const float fi = (float)pixelIndex;
    pixelCoords = V3(fi - W * truncf(fi / W), fi / W, 1.0f);

    fovFactor = 1.7320508075688774f;            /* 1 / tan(60 * (PI / 180) / 2), const-folded */
    cam_origin = xyz(mul(S.uniforms.viewMatrix, V4(V3(0, 0, 0), 1.0f)));
    randState = pixelIndex + (uint32_t)S.uniforms.frameNum * 719393u;
    get_lights();
    return pathTrace();
}

}  /* namespace orc */
