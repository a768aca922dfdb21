// mlp.cuh — CTA-level building blocks of the actor / critic MLP (Linear-Tanh-Linear-Tanh-Linear,
// PPO/Agent.cpp:25-59) on a tile of TM rows held in shared memory, fp32 FMA path.
//
// A CTA has MLP_THREADS = 256 threads.  Activations live in shared memory as [TM][LDA] with
// LDA = H + 4 (rows stay 16 B aligned, consecutive rows shift by 4 banks).  The H x H weight is
// streamed through a 64 x 64 shared-memory chunk Ws[64][68] (it is L2-resident: 16 KB at H=64,
// 256 KB at H=256).  Small operands (W1, W3, biases) are read through the read-only path.
//
// The three GEMM shapes of one hidden layer:
//   gemm_nt : C[m][n] = sum_k A[m][k] * W[n][k]        forward   (y = x W^T)
//   gemm_nn : C[m][k] = sum_n D[m][n] * W[n][k]        backward  (dX = dY W)
//   gemm_tn : G[n][k] = sum_m D[m][n] * A[m][k]        backward  (dW = dY^T X)
// The tcgen05 (TF32, TMEM accumulators) versions of the same three live in update_tc.cu.
#pragma once
#include "common.cuh"

namespace ppo {

constexpr int MLP_THREADS = 256;
constexpr int WS_LD = 68;  // Ws[64][68]

__host__ __device__ constexpr int lda_of(int H) { return H + 4; }

__device__ inline float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
// 4 consecutive floats from global memory through the read-only path.  The flat parameter slab keeps the
// reference's order (PPO/Agent.cpp:41-66), in which the critic's 1-element output bias leaves every actor
// tensor off 16 B alignment, so the vector load is only taken when the address allows it.
__device__ inline float4 ldg4_any(const float* p) {
    if ((reinterpret_cast<uintptr_t>(p) & 15u) == 0) return __ldg(reinterpret_cast<const float4*>(p));
    return make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3));
}

// Stage W[n0:n0+64, k0:k0+64] (row-major, leading dim ldw) into Ws[64][WS_LD]; coalesced along k.
__device__ inline void stage_w(float* Ws, const float* __restrict__ W, int ldw, int n0, int k0) {
    for (int e = threadIdx.x; e < 64 * 16; e += MLP_THREADS) {
        const int r = e >> 4, c4 = (e & 15) * 4;
        const float4 v = ldg4_any(W + (size_t)(n0 + r) * ldw + k0 + c4);
        *reinterpret_cast<float4*>(Ws + r * WS_LD + c4) = v;
    }
}

// ---- layer 1: Hout[m][j] = tanh(b1[j] + sum_i X[m][i] W1[j][i]),  X smem [TM][O] ----------
template <int TM, int H>
__device__ inline void layer1_fwd(const float* Xs, int O, const float* __restrict__ W1,
                                  const float* __restrict__ b1, float* Hout) {
    constexpr int LDA = lda_of(H);
    for (int idx = threadIdx.x; idx < TM * H; idx += MLP_THREADS) {
        const int m = idx / H, j = idx % H;
        float acc = __ldg(b1 + j);
        for (int i = 0; i < O; i++) acc = fmaf(Xs[m * O + i], __ldg(W1 + j * O + i), acc);
        Hout[m * LDA + j] = tanhf(acc);
    }
}

// ---- hidden layer forward: Hout = tanh(Hin W2^T + b2) ---------------------------------------
template <int TM, int H>
__device__ inline void layer2_fwd(const float* Hin, const float* __restrict__ W2, const float* __restrict__ b2,
                                  float* Hout, float* Ws) {
    constexpr int LDA = lda_of(H);
    constexpr int RM = TM / 16;  // rows per thread
    static_assert(TM % 16 == 0 && H % 64 == 0, "tile shape");
    const int g = threadIdx.x & 15, mg = threadIdx.x >> 4;
    for (int n0 = 0; n0 < H; n0 += 64) {
        float acc[RM][4];
#pragma unroll
        for (int i = 0; i < RM; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
        for (int k0 = 0; k0 < H; k0 += 64) {
            __syncthreads();  // previous users of Ws are done
            stage_w(Ws, W2, H, n0, k0);
            __syncthreads();
#pragma unroll 4
            for (int k = 0; k < 64; k += 4) {
                float4 a[RM], w[4];
#pragma unroll
                for (int i = 0; i < RM; i++) a[i] = ld4(Hin + (mg + 16 * i) * LDA + k0 + k);
#pragma unroll
                for (int j = 0; j < 4; j++) w[j] = ld4(Ws + (g + 16 * j) * WS_LD + k);
#pragma unroll
                for (int i = 0; i < RM; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        acc[i][j] = fmaf(a[i].x, w[j].x, acc[i][j]);
                        acc[i][j] = fmaf(a[i].y, w[j].y, acc[i][j]);
                        acc[i][j] = fmaf(a[i].z, w[j].z, acc[i][j]);
                        acc[i][j] = fmaf(a[i].w, w[j].w, acc[i][j]);
                    }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int n = n0 + g + 16 * j;
            const float b = __ldg(b2 + n);
#pragma unroll
            for (int i = 0; i < RM; i++) Hout[(mg + 16 * i) * LDA + n] = tanhf(acc[i][j] + b);
        }
    }
    __syncthreads();
}

// ---- output layer: out[m][q] = b3[q] + sum_k Hin[m][k] W3[q][k];  out smem [TM][ldo] ----------
// a group of 4 lanes shares one output (k split 4 ways), reduced with two shuffles
template <int TM, int H>
__device__ inline void layer3_fwd(const float* Hin, const float* __restrict__ W3, const float* __restrict__ b3,
                                  int K3, float* out, int ldo) {
    constexpr int LDA = lda_of(H);
    const int part = threadIdx.x & 3, grp = threadIdx.x >> 2;
    const int total = TM * K3;
    for (int base = 0; base < total; base += MLP_THREADS / 4) {
        const int idx = base + grp;
        const bool on = idx < total;
        const int m = on ? idx / K3 : 0, q = on ? idx % K3 : 0;
        float acc = 0.0f;
        for (int k = part * 4; k < H; k += 16) {
            const float4 a = ld4(Hin + m * LDA + k);
            const float4 w = ldg4_any(W3 + q * H + k);
            acc = fmaf(a.x, w.x, acc); acc = fmaf(a.y, w.y, acc);
            acc = fmaf(a.z, w.z, acc); acc = fmaf(a.w, w.w, acc);
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (on && part == 0) out[m * ldo + q] = acc + __ldg(b3 + q);
    }
    __syncthreads();
}

// ---- backward of the output layer -----------------------------------------------------------
// dZ2 = (dOut W3) * (1 - H2^2), written in place over H2.  dOut smem [TM][ldo].
template <int TM, int H>
__device__ inline void layer3_bwd_dx(float* H2, const float* __restrict__ W3, int K3, const float* dOut, int ldo) {
    constexpr int LDA = lda_of(H);
    for (int idx = threadIdx.x; idx < TM * H; idx += MLP_THREADS) {
        const int m = idx / H, k = idx % H;
        float acc = 0.0f;
        for (int q = 0; q < K3; q++) acc = fmaf(dOut[m * ldo + q], __ldg(W3 + q * H + k), acc);
        const float h = H2[m * LDA + k];
        H2[m * LDA + k] = acc * (1.0f - h * h);
    }
}
// gW3[q][k] += sum_m dOut[m][q] H2[m][k];  gb3[q] += sum_m dOut[m][q]   (H2 still holds activations)
template <int TM, int H>
__device__ inline void layer3_bwd_dw(const float* H2, int K3, const float* dOut, int ldo, int rows,
                                     float* __restrict__ gW3, float* __restrict__ gb3) {
    constexpr int LDA = lda_of(H);
    for (int idx = threadIdx.x; idx < K3 * H; idx += MLP_THREADS) {
        const int q = idx / H, k = idx % H;
        float acc = 0.0f;
        for (int m = 0; m < rows; m++) acc = fmaf(dOut[m * ldo + q], H2[m * LDA + k], acc);
        gW3[idx] += acc;
    }
    for (int q = threadIdx.x; q < K3; q += MLP_THREADS) {
        float acc = 0.0f;
        for (int m = 0; m < rows; m++) acc += dOut[m * ldo + q];
        gb3[q] += acc;
    }
}

// ---- hidden layer backward ---------------------------------------------------------------
// gW2[n][k] += sum_m dZ2[m][n] H1[m][k];  gb2[n] += sum_m dZ2[m][n]   (rows beyond `rows` hold zeros)
template <int TM, int H>
__device__ inline void layer2_bwd_dw(const float* dZ2, const float* H1, float* __restrict__ gW2,
                                     float* __restrict__ gb2) {
    constexpr int LDA = lda_of(H);
    const int gk = threadIdx.x & 15, gn = threadIdx.x >> 4;
    for (int n0 = 0; n0 < H; n0 += 64)
        for (int k0 = 0; k0 < H; k0 += 64) {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
#pragma unroll 4
            for (int m = 0; m < TM; m++) {
                const float4 d = ld4(dZ2 + m * LDA + n0 + gn * 4);
                const float4 a = ld4(H1 + m * LDA + k0 + gk * 4);
                const float dd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    acc[i][0] = fmaf(dd[i], a.x, acc[i][0]);
                    acc[i][1] = fmaf(dd[i], a.y, acc[i][1]);
                    acc[i][2] = fmaf(dd[i], a.z, acc[i][2]);
                    acc[i][3] = fmaf(dd[i], a.w, acc[i][3]);
                }
            }
            const bool al = (reinterpret_cast<uintptr_t>(gW2) & 15u) == 0;  // actor slabs sit 1 float off (see ldg4_any)
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float* q = gW2 + (size_t)(n0 + gn * 4 + i) * H + k0 + gk * 4;
                if (al) {
                    float4* p = reinterpret_cast<float4*>(q);
                    float4 v = *p;
                    v.x += acc[i][0]; v.y += acc[i][1]; v.z += acc[i][2]; v.w += acc[i][3];
                    *p = v;
                } else {
                    q[0] += acc[i][0]; q[1] += acc[i][1]; q[2] += acc[i][2]; q[3] += acc[i][3];
                }
            }
        }
    for (int n = threadIdx.x; n < H; n += MLP_THREADS) {
        float acc = 0.0f;
        for (int m = 0; m < TM; m++) acc += dZ2[m * LDA + n];
        gb2[n] += acc;
    }
}
// dZ1 = (dZ2 W2) * (1 - H1^2), written in place over H1
template <int TM, int H>
__device__ inline void layer2_bwd_dx(const float* dZ2, const float* __restrict__ W2, float* H1, float* Ws) {
    constexpr int LDA = lda_of(H);
    constexpr int RM = TM / 16;
    const int g = threadIdx.x & 15, mg = threadIdx.x >> 4;
    for (int k0 = 0; k0 < H; k0 += 64) {
        float acc[RM][4];
#pragma unroll
        for (int i = 0; i < RM; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
        for (int n0 = 0; n0 < H; n0 += 64) {
            __syncthreads();
            stage_w(Ws, W2, H, n0, k0);
            __syncthreads();
#pragma unroll 4
            for (int n = 0; n < 64; n += 4) {
                float4 d[RM], w[4];
#pragma unroll
                for (int i = 0; i < RM; i++) d[i] = ld4(dZ2 + (mg + 16 * i) * LDA + n0 + n);
#pragma unroll
                for (int j = 0; j < 4; j++) w[j] = ld4(Ws + (n + j) * WS_LD + g * 4);
#pragma unroll
                for (int i = 0; i < RM; i++) {
                    const float dd[4] = {d[i].x, d[i].y, d[i].z, d[i].w};
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        acc[i][0] = fmaf(dd[j], w[j].x, acc[i][0]);
                        acc[i][1] = fmaf(dd[j], w[j].y, acc[i][1]);
                        acc[i][2] = fmaf(dd[j], w[j].z, acc[i][2]);
                        acc[i][3] = fmaf(dd[j], w[j].w, acc[i][3]);
                    }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < RM; i++) {
            float4* p = reinterpret_cast<float4*>(H1 + (mg + 16 * i) * LDA + k0 + g * 4);
            const float4 h = *p;
            float4 o;
            o.x = acc[i][0] * (1.0f - h.x * h.x); o.y = acc[i][1] * (1.0f - h.y * h.y);
            o.z = acc[i][2] * (1.0f - h.z * h.z); o.w = acc[i][3] * (1.0f - h.w * h.w);
            *p = o;
        }
    }
    __syncthreads();
}

// ---- layer 1 backward: gW1[j][i] += sum_m dZ1[m][j] X[m][i];  gb1[j] += sum_m dZ1[m][j] --------
template <int TM, int H>
__device__ inline void layer1_bwd_dw(const float* dZ1, const float* Xs, int O, float* __restrict__ gW1,
                                     float* __restrict__ gb1) {
    constexpr int LDA = lda_of(H);
    for (int idx = threadIdx.x; idx < H * O; idx += MLP_THREADS) {
        const int j = idx / O, i = idx % O;
        float acc = 0.0f;
        for (int m = 0; m < TM; m++) acc = fmaf(dZ1[m * LDA + j], Xs[m * O + i], acc);
        gW1[idx] += acc;
    }
    for (int j = threadIdx.x; j < H; j += MLP_THREADS) {
        float acc = 0.0f;
        for (int m = 0; m < TM; m++) acc += dZ1[m * LDA + j];
        gb1[j] += acc;
    }
}

}  // namespace ppo
