// h256.cuh — building blocks shared by the hidden = 256 tensor-core kernels (update_h256.cu, rollout_h256.cu):
// the 64-column "panel" form of a 256-wide 16-bit tile in shared memory and the tcgen05 descriptors over it.
//
//   activation tile X  [128 rows][256]  = 4 panels of [128 rows][128 B] (16 KB each), 128-byte swizzle
//   weight tile     W2 [256 rows][256]  = 4 panels of [256 rows][128 B] (32 KB each), 128-byte swizzle
// Element (r, c) of a tile lives in panel c / 64 at  r * 128 + ((((c % 64) / 8) ^ (r % 8)) * 16) + (c % 8) * 2.
// K-major view (K along the row): one descriptor per panel and 16-element K step (umma.cuh: act_desc_kmajor).
// MN-major view (K down the rows; MN spans panels): leading byte offset = panel stride (mn_desc below).
#pragma once
#include "umma.cuh"

namespace ppo {

constexpr int HH = 256;                 // hidden width served here
constexpr int HR = 128;                 // rows per tile
constexpr int H_PANEL_X = HR * 128;     // one 64-column panel of the activation tile (16 KB)
constexpr int H_PANEL_W = HH * 128;     // one 64-column panel of W2 (32 KB)
constexpr int H_NCOMP = 512;            // compute threads (4 warpgroups)
constexpr int H_NTHR = 640;             // + the control warpgroup
constexpr int H_MAX_K = 8;


__device__ __forceinline__ float tanh_fast256(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// MN-major view across 64-column panels `panel_stride` bytes apart: K step ks = rows 16 ks .. 16 ks + 15
__device__ __forceinline__ uint64_t mn_desc(uint32_t base, uint32_t panel_stride, int ks) {
    return umma::smem_desc(base + (uint32_t)ks * 2048u, panel_stride, 1024, umma::LAYOUT_SW128);
}
template <int FMT>
__device__ __forceinline__ void st8(uint8_t* dst, const float* v) {
    uint4 q;
    q.x = umma::pack2<FMT>(v[0], v[1]); q.y = umma::pack2<FMT>(v[2], v[3]);
    q.z = umma::pack2<FMT>(v[4], v[5]); q.w = umma::pack2<FMT>(v[6], v[7]);
    *reinterpret_cast<uint4*>(dst) = q;
}
// W2 [j][i] (fp32, row-major, global) -> 16-bit panels at `dst`: chunk task e = row j (e >> 5), 8 columns (e & 31)
template <int FMT>
__device__ __forceinline__ void stage_w2_panels(uint8_t* dst, const float* __restrict__ W2, int tid, int nthreads) {
    for (int e = tid; e < HH * 32; e += nthreads) {
        const int j = e >> 5, c = e & 31;
        const float* src = W2 + j * HH + c * 8;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = __ldg(src + i);
        st8<FMT>(dst + (c >> 3) * H_PANEL_W + umma::act_chunk_off(j, c & 7), v);
    }
}
// layer 1 on the CUDA cores for 8 consecutive hidden units: h = tanh(P1[j] . xe), P1 row = {W1[j][0..O), b1[j], 0..}
template <int OP>
__device__ __forceinline__ void layer1_8(const float* P1rows, const float (&xe)[OP], float* h) {
#pragma unroll
    for (int jj = 0; jj < 8; jj++) {
        const float* pr = P1rows + jj * OP;
        const float4 w0 = *reinterpret_cast<const float4*>(pr);
        float acc = xe[0] * w0.x;
        acc = fmaf(xe[1], w0.y, acc); acc = fmaf(xe[2], w0.z, acc); acc = fmaf(xe[3], w0.w, acc);
        if (OP == 8) {
            const float4 w1 = *reinterpret_cast<const float4*>(pr + 4);
            acc = fmaf(xe[OP - 4], w1.x, acc); acc = fmaf(xe[OP - 3], w1.y, acc);
            acc = fmaf(xe[OP - 2], w1.z, acc); acc = fmaf(xe[OP - 1], w1.w, acc);
        }
        h[jj] = tanh_fast256(acc);
    }
}

}  // namespace ppo
