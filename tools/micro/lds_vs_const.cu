// lds_vs_const.cu — cost of warp-uniform weights: broadcast LDS.128 from shared memory vs constant-bank FFMA operands.
// One CTA per SM, 256 threads (2 warps / SMSP, like update_tc_kernel): per "unit" 8 FFMA + the weight fetch.
#include <cstdio>
#include <cuda_runtime.h>
__constant__ float cw[64 * 8];
template <int MODE>
__global__ void k(float* out, long long* cyc, int iters, const float* gw) {
    __shared__ __align__(16) float sw[64 * 8];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) sw[i] = gw[i];
    __syncthreads();
    float xe[8];
    for (int i = 0; i < 8; i++) xe[i] = threadIdx.x * 0.001f + i;
    float tot = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int j = 0; j < 64; j++) {
            float acc;
            if (MODE == 0) {
                float4 w0, w1;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(w0.x), "=f"(w0.y), "=f"(w0.z), "=f"(w0.w)
                             : "r"((unsigned)__cvta_generic_to_shared(sw + j * 8)));
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(w1.x), "=f"(w1.y), "=f"(w1.z), "=f"(w1.w)
                             : "r"((unsigned)__cvta_generic_to_shared(sw + j * 8 + 4)));
                acc = xe[0] * w0.x;
                acc = fmaf(xe[1], w0.y, acc); acc = fmaf(xe[2], w0.z, acc); acc = fmaf(xe[3], w0.w, acc);
                acc = fmaf(xe[4], w1.x, acc); acc = fmaf(xe[5], w1.y, acc); acc = fmaf(xe[6], w1.z, acc);
                acc = fmaf(xe[7], w1.w, acc);
            } else {
                acc = xe[0] * cw[j * 8];
#pragma unroll
                for (int i = 1; i < 8; i++) acc = fmaf(xe[i], cw[j * 8 + i], acc);
            }
            tot += acc;
        }
        xe[0] += tot * 1e-20f;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = tot;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
    float *out, *gw; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8); cudaMalloc(&gw, 2048);
    float h[512]; for (int i = 0; i < 512; i++) h[i] = 0.001f * i;
    cudaMemcpy(gw, h, 2048, cudaMemcpyHostToDevice); cudaMemcpyToSymbol(cw, h, 2048);
    const int iters = 256;
    for (int threads : {128, 256, 512}) {
        long long c;
        k<0><<<148, threads>>>(out, cyc, iters, gw); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("threads=%d  smem broadcast LDS.128: %.1f cycles per 64-unit layer\n", threads, (double)c / iters);
        k<1><<<148, threads>>>(out, cyc, iters, gw); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("threads=%d  constant-bank operands: %.1f cycles per 64-unit layer\n", threads, (double)c / iters);
    }
    return 0;
}
