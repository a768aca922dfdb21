// fma_rate.cu — issue rate of FFMA vs FFMA2 (fma.rn.f32x2) on sm_100a: cycles per warp-instruction per SMSP.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_rate fma_rate.cu && ./fma_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, long long* cyc, int iters, float a0, float b0) {
    float a = a0 + threadIdx.x, b = b0;
    float acc[16];
    for (int i = 0; i < 16; i++) acc[i] = i;
    unsigned long long A, B, C[8];
    asm("mov.b64 %0, {%1, %2};" : "=l"(A) : "f"(a), "f"(a + 1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(B) : "f"(b), "f"(b + 1));
    for (int i = 0; i < 8; i++) asm("mov.b64 %0, {%1, %2};" : "=l"(C[i]) : "f"((float)i), "f"((float)i + 8));
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(a, b, acc[i]);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(C[i]) : "l"(A), "l"(B));
        } else {   // distinct b operands per instruction (no operand reuse)
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(acc[(i + 1) & 15], b, acc[i]);
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 16; i++) s += acc[i];
    for (int i = 0; i < 8; i++) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(C[i])); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 4096;
    for (int threads : {128, 256, 512, 1024}) {
        for (int mode = 0; mode < 3; mode++) {
            if (mode == 0) k<0><<<148, threads>>>(out, cyc, iters, 1.0f, 1e-9f);
            if (mode == 1) k<1><<<148, threads>>>(out, cyc, iters, 1.0f, 1e-9f);
            if (mode == 2) k<2><<<148, threads>>>(out, cyc, iters, 1.0f, 1e-9f);
            long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            const int per_iter = mode == 1 ? 8 : 16, warps_per_smsp = threads / 128;
            printf("threads=%4d mode=%d (%s): %.2f cycles per warp-instr per SMSP, %.1f fp32 FMA/clk/SM\n", threads, mode,
                   mode == 0 ? "FFMA same a,b" : mode == 1 ? "FFMA2" : "FFMA no reuse",
                   (double)c / iters / per_iter / warps_per_smsp,
                   (mode == 1 ? 64.0 : 32.0) * per_iter * warps_per_smsp * 4 * iters / (double)c);
        }
    }
    return 0;
}
