// camera.cuh -- pixel coordinates, sub-pixel jitter and the pinhole camera ray, pinned arithmetic.
// Restates /root/reference/shaders/main.wgsl:5-8,16 and shootRay.wgsl:11-40,54-60.
#pragma once
#include "intersect.cuh"

namespace wg {

struct PixelSetup {
    float px, py;        // pixelCoords.xy (main.wgsl:5: float remainder, UNFLOORED quotient)
    v3 cam_origin;       // viewMatrix * (0,0,0,1)
};

__device__ __forceinline__ PixelSetup pixel_setup(const DeviceScene& S, uint32_t pixelIndex) {
    PixelSetup p;
    const float W = S.uniforms.screenW;
    const float fi = (float)pixelIndex;
    p.px = fs(fi, fm(W, truncf(fd(fi, W))));
    p.py = fd(fi, W);
    p.cam_origin = mul4_xyz(S.uniforms.viewMatrix, V3s(0.0f), 1.0f);
    return p;
}

// main.wgsl:16
__device__ __forceinline__ uint32_t pixel_seed(uint32_t pixelIndex, float frameNum) {
    return pixelIndex + (uint32_t)frameNum * 719393u;
}

// shootRay.wgsl:54-60: normalize(viewMatrix * vec4(s, t, -fovFactor, 0)).xyz -- the vec4 is normalised;
// fovFactor = 1 / tan(60 * (PI / 180) / 2) const-folds to 1.7320508
__device__ __forceinline__ Ray camera_ray(const DeviceScene& S, v3 cam_origin, float s, float t) {
    const v4 d = mul4(S.uniforms.viewMatrix, V3(s, t, -1.7320508075688774f), 0.0f);
    const float l = fsq(fa(fa(fa(fm(d.x, d.x), fm(d.y, d.y)), fm(d.z, d.z)), fm(d.w, d.w)));
    Ray r; r.o = cam_origin; r.d = V3(d.x, d.y, d.z) / l;      // three IEEE quotients by one divisor (pvec.cuh operator/)
    return r;
}

// The camera ray of one sample.  Plain loop (shootRay.wgsl:34-43): two draws, x then y.
// Stratified (shootRay.wgsl:11-30): cell (si, sj) of a sqrt_spp x sqrt_spp grid.
__device__ __forceinline__ Ray sample_ray(const DeviceScene& S, const PixelSetup& p, int si, int sj, uint32_t& rng) {
    const float W = S.uniforms.screenW, H = S.uniforms.screenH;
    const float aspect = S.aspect;                              // fd(W, H), invocation-invariant: divided once on the host
    float ox, oy;
    if (S.K.STRATIFY) {
        const float sqrt_spp = (float)sqrt((double)S.K.NUM_SAMPLES);
        const float recip = fd(1.0f, (float)((int)sqrt_spp));
        const float rx = rand2D(rng);
        ox = fm(recip, fa((float)si, rx));
        const float ry = rand2D(rng);
        oy = fm(recip, fa((float)sj, ry));
    } else {
        ox = rand2D(rng);
        oy = rand2D(rng);
    }
    const float sx = fm(aspect, fs(fm(2.0f, fd(fa(fs(p.px, 0.5f), ox), W)), 1.0f));
    const float sy = fm(-1.0f, fs(fm(2.0f, fd(fa(fs(p.py, 0.5f), oy), H)), 1.0f));
    return camera_ray(S, p.cam_origin, sx, sy);
}

// number of cells per side of the stratified grid: the float loop `i < sqrt_spp` of shootRay.wgsl:15
__device__ __forceinline__ int strat_side(const DeviceScene& S) {
    const float sqrt_spp = (float)sqrt((double)S.K.NUM_SAMPLES);
    int n = 0;
    for (float i = 0.0f; i < sqrt_spp; i += 1.0f) n++;
    return n;
}

}  // namespace wg
