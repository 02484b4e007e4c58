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

// aces + clamp part of tonemap1 (exact float; NaN stays NaN through pmin/pmax like the oracle's)
template <bool BYTES_ONLY>
__device__ __forceinline__ float tonemap_pre(float sum, float frameNum, float rcp_frame, bool frame_ok) {
    const float col = fdiv_shared(sum, frameNum, rcp_frame, frame_ok);      // == fd(sum, frameNum), one reciprocal per kernel
    const float v1 = fm(col, 0.6f);
    const float r = fd(fm(v1, fa(fm(2.51f, v1), 0.03f)), fa(fm(v1, fa(fm(2.43f, v1), 0.59f)), 0.14f));
    // the sign of a zero cannot reach a byte (pow(+-0) = 0), so the bytes-only pass uses the bare FMNMX
    if (BYTES_ONLY) return fminf(fmaxf(r, 0.0f), 1.0f);
    return pmin(pmax(r, 0.0f), 1.0f);
}
__device__ __forceinline__ unsigned char quant8_exact(float r) {
    return (unsigned char)__float2uint_rn(wgpt_pm::powf_(r, 0.45454545454545453f) * 255.0f);
}
// round(255 * pow(r, 1/2.2)) for the 8-bit canvas WITHOUT the binary64 pow whenever the answer is
// already decided: p~ = ex2(lg2(r) / 2.2) on the SFU has relative error < (2.5 + |t| / 2) 2^-22
// (lg2.approx: 2^-22 absolute on the mantissa part; ex2.approx: 2 ulp), so if 255 p~ is further than
// that from a rounding boundary k + 1/2 the rounded byte equals the pinned one.  Otherwise (about 3
// pixels in 10^4) the pinned pow decides.  The test suite checks the byte against quant8_exact for
// EVERY float in [0, 1] (wgpt_present_quant_check), so the bound is verified, not trusted.
__device__ __forceinline__ unsigned char quant8(float r, unsigned* slow) {
    float l, p;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(r));
    const float t = l * 0.45454545454545453f;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p) : "f"(t));
    const float v = p * 255.0f;
    const float m = v + 12582912.0f;                       // 1.5 * 2^23: the sum's low mantissa bits are rint(v), no F2I
    const float k = m - 12582912.0f;
    const float err = fmaf(v * (2.5f + 0.5f * fabsf(t)), 2.3841858e-07f, 1.0e-7f);
    // decided <=> v is at least err away from both k - 1/2 and k + 1/2 (false for NaN, which takes the pinned path)
    if (0.5f - fabsf(v - k) > err) return (unsigned char)(__float_as_uint(m) & 0xffu);
    if (slow) (*slow)++;
    return quant8_exact(r);
}

__global__ void __launch_bounds__(256) k_present(const float4* fb_in, float4* fb_clear,      // may alias (wgpt_present clears binding 3 in place): no __restrict__
                                                 float4* __restrict__ rgba_out, uchar4* __restrict__ rgba8_out,
                                                 uint32_t npix, float frameNum, int clear) {
    const bool frame_ok = frameNum > 8.67e-19f && frameNum < 1.15e18f;
    float r0;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(frameNum));
    const float rcp_frame = __fmaf_rn(r0, __fmaf_rn(-frameNum, r0, 1.0f), r0);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float4 s = fb_in[i];
        if (rgba_out) {
            float4 c;
            c.x = tonemap1(s.x, frameNum); c.y = tonemap1(s.y, frameNum); c.z = tonemap1(s.z, frameNum); c.w = 1.0f;
            rgba_out[i] = c;
            if (rgba8_out) {
                // canvas quantisation: round(c * 255); NaN (0/0 before the first frame) maps to 0
                rgba8_out[i] = make_uchar4((unsigned char)__float2uint_rn(c.x * 255.0f), (unsigned char)__float2uint_rn(c.y * 255.0f),
                                           (unsigned char)__float2uint_rn(c.z * 255.0f), 255);
            }
        } else if (rgba8_out) {
            // bytes only (what the reference's 8-bit canvas shows): 16 B in, 4 B out, no binary64 in the common case
            rgba8_out[i] = make_uchar4(quant8(tonemap_pre<true>(s.x, frameNum, rcp_frame, frame_ok), nullptr),
                                       quant8(tonemap_pre<true>(s.y, frameNum, rcp_frame, frame_ok), nullptr),
                                       quant8(tonemap_pre<true>(s.z, frameNum, rcp_frame, frame_ok), nullptr), 255);
        }
        if (clear) fb_clear[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

// every float r with bit pattern in [first, first + n): does quant8(r) equal quant8_exact(r)?
__global__ void __launch_bounds__(256) k_quant_check(uint32_t first, uint32_t n, unsigned long long* __restrict__ out) {
    unsigned bad = 0, slow = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float r = __uint_as_float(first + i);
        if (quant8(r, &slow) != quant8_exact(r)) bad++;
    }
    for (int o = 16; o > 0; o >>= 1) { bad += __shfl_down_sync(0xffffffffu, bad, o); slow += __shfl_down_sync(0xffffffffu, slow, o); }
    if ((threadIdx.x & 31) == 0) { if (bad) atomicAdd(&out[0], (unsigned long long)bad); if (slow) atomicAdd(&out[1], (unsigned long long)slow); }
}
void launch_quant_check(uint32_t first, uint32_t n, unsigned long long* out, cudaStream_t stream) {
    if (n) k_quant_check<<<148 * 8, 256, 0, stream>>>(first, n, out);
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

// after a sum-reduce over shards the w lane counts the shards that rendered the pixel: back to the reference's 1.0
__global__ void __launch_bounds__(256) k_fix_alpha(float4* __restrict__ fb, uint32_t npix) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        float* w = reinterpret_cast<float*>(&fb[i]) + 3;
        if (*w != 0.0f && *w != 1.0f) *w = 1.0f;
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

void launch_fix_alpha(float4* fb, uint32_t npix, cudaStream_t stream) {
    k_fix_alpha<<<stream_grid(npix, 256), 256, 0, stream>>>(fb, npix);
}

void launch_fb_to_rgb(const float4* fb, float* rgb, uint32_t npix, cudaStream_t stream) {
    k_fb_to_rgb<<<stream_grid(npix, 256), 256, 0, stream>>>(fb, rgb, npix);
}

void launch_accumulate(float4* fb, const float4* frame, uint32_t npix, int reset, cudaStream_t stream) {
    k_accumulate<<<stream_grid(npix, 256), 256, 0, stream>>>(fb, frame, npix, reset);
}

}  // namespace wg

// ---------------------------------------------------------------------------------------------------------------
// wgpt_present_rgba8_static -- the bytes-only present WITHOUT any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  wgpt_present cudaMallocs its device staging on a context's first call; this variant never allocates: it
// stages the RGBA8 image in a file-scope device array of WGPT_STATIC_PRESENT_MAX_PIXELS x 4 bytes (8 MiB), which every
// GPU that loads the library holds until exit.
//   * In HOST code the array's name is a host-side stand-in, not a device address: its device address is looked up once
//     per device, with that device current, by cudaGetSymbolAddress, and THAT pointer goes to the kernel.
//   * Every call on a device shares the array and it keeps the last call's contents; the kernel overwrites all 4 x npix
//     bytes it then copies out, so nothing has to start at zero.  Calls that overlap on the GPU (two contexts of one
//     device on their own streams, two host threads) RACE on it: serialise them.
//   * Same kernel, same bytes as wgpt_present(ctx, NULL, rgba8_out), including the clear of binding 3 when
//     uniforms.resetBuffer == 1 (fragment.js:30-33).
#include <string>
#include "context.h"

// WGPT_STATIC_PRESENT_MAX_PIXELS = 2 Mi pixels (include/wgpt.h): 1920 x 1080 fits (2,073,600)
static __device__ __align__(256) unsigned char g_static_present_rgba8[(size_t)WGPT_STATIC_PRESENT_MAX_PIXELS * 4];

#define WGPT_STATIC_MAX_DEVICES 64
static uchar4* g_static_present_ptr[WGPT_STATIC_MAX_DEVICES];      // device address of the array, per device ordinal

extern "C" int wgpt_present_rgba8_static(wgpt_ctx* c, uint8_t* rgba8_out) {
    if (!c || !rgba8_out) return WGPT_E_INVALID;
    if (!c->have_uniforms) return wg::set_err(c, WGPT_E_STATE, "uniforms (binding 0) not written");
    const uint32_t npix = (uint32_t)c->cfg.width * (uint32_t)c->cfg.height;
    if (npix > WGPT_STATIC_PRESENT_MAX_PIXELS)
        return wg::set_err(c, WGPT_E_INVALID, "wgpt_present_rgba8_static: image has " + std::to_string(npix) +
                           " pixels, the static staging array holds at most " + std::to_string(WGPT_STATIC_PRESENT_MAX_PIXELS) +
                           " (use wgpt_present)");
    const int dev = c->cfg.device;
    if (dev < 0 || dev >= WGPT_STATIC_MAX_DEVICES) return wg::set_err(c, WGPT_E_INVALID, "wgpt_present_rgba8_static: device ordinal >= 64");
    cudaError_t e = cudaSetDevice(dev);
    if (e != cudaSuccess) return wg::cuda_err(c, e, "present_rgba8_static");
    if (!g_static_present_ptr[dev]) {                      // set-up, once per device, with that device current
        void* p = nullptr;
        if ((e = cudaGetSymbolAddress(&p, g_static_present_rgba8)) != cudaSuccess) return wg::cuda_err(c, e, "cudaGetSymbolAddress(static present staging)");
        g_static_present_ptr[dev] = (uchar4*)p;
    }
    uchar4* d_rgba8 = g_static_present_ptr[dev];
    {   // `fs` runs after the compute passes submitted before it (launches only: the wavefront pool exists already)
        const int rc = wg::fast_flush(c);
        if (rc) return rc;
    }
    const int clear = c->uniforms.resetBuffer == 1.0f;
    wg::launch_present((const float4*)c->dbuf[3], (float4*)c->dbuf[3], nullptr, d_rgba8, npix, c->uniforms.frameNum, clear, c->stream);
    if ((e = cudaGetLastError()) != cudaSuccess) return wg::cuda_err(c, e, "present_rgba8_static");
    if ((e = cudaMemcpyAsync(rgba8_out, d_rgba8, (size_t)npix * 4, cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess)
        return wg::cuda_err(c, e, "present_rgba8_static");
    return wg::cuda_err(c, cudaStreamSynchronize(c->stream), "present_rgba8_static");
}
