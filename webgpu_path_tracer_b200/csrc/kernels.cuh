// kernels.cuh -- host-callable launchers of the CUDA kernels (one per .cu file).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "scene.cuh"

namespace wg {

// exact.cu (K1)
void launch_exact_frame(const DeviceScene& S, float4* fb, float* radiance_out, int* prim_type, int* prim_index,
                        float* prim_t, cudaStream_t stream);
void launch_trace_rays_binary(const DeviceScene& S, int n, const float* rays, const uint32_t* seeds, int* out_type,
                              int* out_index, float* out_t, cudaStream_t stream);
void launch_pinned_math(int fn, int n, const float* x, const float* y, float* out, cudaStream_t stream);

// present.cu (K4)
void launch_present(const float4* fb_in, float4* fb_clear, float4* rgba_out, uchar4* rgba8_out, uint32_t npix,
                    float frameNum, int clear, cudaStream_t stream);
void launch_fix_alpha(float4* fb, uint32_t npix, cudaStream_t stream);
void launch_fb_to_rgb(const float4* fb, float* rgb, uint32_t npix, cudaStream_t stream);
void launch_quant_check(uint32_t first, uint32_t n, unsigned long long* out, cudaStream_t stream);
void launch_accumulate(float4* fb, const float4* frame, uint32_t npix, int reset, cudaStream_t stream);

}  // namespace wg
