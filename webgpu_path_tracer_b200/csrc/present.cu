// present.cu -- K4: vectorised float4 accumulate and present (tonemap) passes.
// present restates `fs` of /root/reference/shaders/fragment.js:22-36 with aces_approx of
// shaders/common.wgsl:273-282: c = fb.xyz / frameNum; c = clamp(aces(0.6 c)); c = pow(c, 1/2.2);
// and clears binding 3 when uniforms.resetBuffer == 1.  accumulate is main.wgsl:20-27 applied to
// a whole frame of per-pixel radiance (used by the wavefront path, which stages frames so that
// the per-pixel float sum is taken in frame order).
// Both are pure streams: 16 B in + 16 B out per pixel (present: +4 B when RGBA8 is requested),
// one float4 per thread, grid-stride, fully coalesced.
#include "kernels.cuh"
#include "pvec.cuh"
#include "../../include/wgpt_pinned_math.h"

namespace wg {

__device__ __forceinline__ float tonemap1(float sum, float frameNum) {
    const float col = fd(sum, frameNum);
    const float v1 = fm(col, 0.6f);
    float r = fd(fm(v1, fa(fm(2.51f, v1), 0.03f)), fa(fm(v1, fa(fm(2.43f, v1), 0.59f)), 0.14f));
    r = pmin(pmax(r, 0.0f), 1.0f);
    return wgpt_pm::powf_(r, 0.45454545454545453f);
}

__global__ void __launch_bounds__(256) k_present(const float4* __restrict__ fb_in, float4* __restrict__ fb_clear,
                                                 float4* __restrict__ rgba_out, uchar4* __restrict__ rgba8_out,
                                                 uint32_t npix, float frameNum, int clear) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float4 s = fb_in[i];
        float4 c;
        c.x = tonemap1(s.x, frameNum); c.y = tonemap1(s.y, frameNum); c.z = tonemap1(s.z, frameNum); c.w = 1.0f;
        if (rgba_out) rgba_out[i] = c;
        if (rgba8_out) {
            // canvas quantisation: round(c * 255); NaN (0/0 before the first frame) maps to 0
            rgba8_out[i] = make_uchar4((unsigned char)__float2uint_rn(c.x * 255.0f), (unsigned char)__float2uint_rn(c.y * 255.0f),
                                       (unsigned char)__float2uint_rn(c.z * 255.0f), 255);
        }
        if (clear) fb_clear[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

__global__ void __launch_bounds__(256) k_accumulate(float4* __restrict__ fb, const float4* __restrict__ frame,
                                                    uint32_t npix, int reset) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float4 c = frame[i];
        if (c.w == 0.0f) continue;            // pixel not rendered by this shard: leave binding 3 untouched
        float4 o;
        if (!reset) { const float4 s = fb[i]; o = make_float4(fa(s.x, c.x), fa(s.y, c.y), fa(s.z, c.z), 1.0f); }
        else o = make_float4(c.x, c.y, c.z, 1.0f);
        fb[i] = o;
    }
}

__global__ void __launch_bounds__(256) k_fb_to_rgb(const float4* __restrict__ fb, float* __restrict__ rgb, uint32_t npix) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float4 s = fb[i];
        rgb[3 * i] = s.x; rgb[3 * i + 1] = s.y; rgb[3 * i + 2] = s.z;
    }
}

static unsigned stream_grid(uint32_t n, unsigned threads) {
    unsigned blocks = (n + threads - 1) / threads;
    const unsigned cap = 148u * 8u;          // 148 SMs x 8 resident 256-thread CTAs
    return blocks < cap ? (blocks ? blocks : 1u) : cap;
}

void launch_present(const float4* fb_in, float4* fb_clear, float4* rgba_out, uchar4* rgba8_out, uint32_t npix,
                    float frameNum, int clear, cudaStream_t stream) {
    k_present<<<stream_grid(npix, 256), 256, 0, stream>>>(fb_in, fb_clear, rgba_out, rgba8_out, npix, frameNum, clear);
}

void launch_fb_to_rgb(const float4* fb, float* rgb, uint32_t npix, cudaStream_t stream) {
    k_fb_to_rgb<<<stream_grid(npix, 256), 256, 0, stream>>>(fb, rgb, npix);
}

void launch_accumulate(float4* fb, const float4* frame, uint32_t npix, int reset, cudaStream_t stream) {
    k_accumulate<<<stream_grid(npix, 256), 256, 0, stream>>>(fb, frame, npix, reset);
}

}  // namespace wg
