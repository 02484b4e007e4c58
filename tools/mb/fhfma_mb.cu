// Issue-rate microbenchmark: FFMA vs FHFMA (fma.rn.f32.f16) vs HADD2.F32 + FFMA, 8 independent chains per thread.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fhfma_mb tools/mb/fhfma_mb.cu ; run on the B200 box.
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
__device__ __forceinline__ float fhfma(uint32_t a, uint32_t b, float c) {
    float d; asm volatile("fma.rn.f32.f16 %0, %1, %2, %3;" : "=f"(d) : "h"((unsigned short)a), "h"((unsigned short)b), "f"(c)); return d;
}
template <int MODE>
__global__ void k(float* out, uint32_t a, uint32_t b, float fa, float fb, int iters) {
    float acc[8];
    #pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = (float)(threadIdx.x + i);
    for (int it = 0; it < iters; it++) {
        #pragma unroll
        for (int r = 0; r < 8; r++) {
            #pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) acc[i] = fmaf(fa, fb, acc[i]);
                else if (MODE == 1) acc[i] = fhfma(a, b, acc[i]);
                else { float f; asm volatile("cvt.f32.f16 %0, %1;" : "=f"(f) : "h"((unsigned short)(a + i))); acc[i] = fmaf(f, fb, acc[i]); }
            }
        }
    }
    float s = 0; for (int i = 0; i < 8; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096;
    for (int mode = 0; mode < 3; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(out, 0x6412, 0x3c00, 1.0001f, 0.5f, iters);
            if (mode == 1) k<1><<<148 * 8, 256>>>(out, 0x6412, 0x3c00, 1.0001f, 0.5f, iters);
            if (mode == 2) k<2><<<148 * 8, 256>>>(out, 0x6412, 0x3c00, 1.0001f, 0.5f, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double inst = 148.0 * 8 * 256 / 32 * (double)iters * 64 * (mode == 2 ? 2 : 1);
            if (rep) printf("%s: %.3f ms, %.1f G warp-inst/s (%.2f per clk per SM at 1.9 GHz)\n",
                            mode == 0 ? "FFMA" : mode == 1 ? "FHFMA" : "HADD2.F32+FFMA", ms, inst / ms / 1e6, inst / ms / 1e6 / 148 / 1.9);
        }
    }
    return 0;
}
