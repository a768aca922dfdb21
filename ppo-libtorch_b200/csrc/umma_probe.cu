// umma_probe.cu — known-answer probe of the tcgen05 building blocks in umma.cuh (descriptor bit layouts,
// the dual-use swizzled activation tile, the narrow tile, the M = 64 tensor-memory layout).  One CTA, one GEMM
// per launch, operands rounded to fp16 (fmt 0) or bf16 (fmt 1), fp32 accumulation:
//   mode 0: D[128,64] = A[128,64] * B[64,64]^T   A K-major, B K-major        (forward  y = x W^T)
//   mode 1: D[128,64] = A[128,64] * B[64,64]     A K-major, B MN-major       (backward dX = dY W)
//   mode 2: D[64,64]  = A[128,64]^T * B[128,64]  A MN-major (M = 64), B MN-major (backward dW = dY^T X)
//   mode 3: D[64,16]  = A[128,64]^T * B[128,16]  B = narrow tile of 2 chunks, N = 16   (dW1 / db / dW3)
//   mode 4: mode 2 with the accumulator placed at lane offset 16 (two M = 64 accumulators interleaved in lanes)
// D comes back as the raw tensor-memory image out[lane][col] (128 x 64); tests/test_gpu_tensorcore.py maps
// M = 64 rows to lanes (m % 16) + 32 * (m / 16).
#include "common.cuh"
#include "umma.cuh"

namespace ppo {
using namespace umma;

template <int FMT>
__device__ inline void store_row_chunk(uint8_t* dst, const float* src) {  // 8 floats -> one 16-byte chunk
    uint4 v;
    v.x = pack2<FMT>(src[0], src[1]); v.y = pack2<FMT>(src[2], src[3]);
    v.z = pack2<FMT>(src[4], src[5]); v.w = pack2<FMT>(src[6], src[7]);
    *reinterpret_cast<uint4*>(dst) = v;
}

template <int FMT>
__global__ void __launch_bounds__(128) umma_probe_kernel(int mode, const float* __restrict__ A,
                                                         const float* __restrict__ B, float* __restrict__ out, int reps,
                                                         long long* __restrict__ cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // swizzle atoms need 1024 B alignment
    uint8_t* sA = smem;                 // activation tile, 128 rows x 128 B = 16 KB
    uint8_t* sB = smem + 16384;         // activation tile (64 or 128 rows) or narrow tile
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int t = threadIdx.x, warp = t >> 5;

    if (warp == 0) tmem_alloc(&tmem_base_s, 64);
    if (t == 0) {
        mbar_init(&bar, 1);
        mbar_init_fence();
    }
    // stage operands: thread = row
    for (int c = 0; c < 8; c++) store_row_chunk<FMT>(sA + act_chunk_off(t, c), A + t * 64 + c * 8);
    if (mode == 0 || mode == 1) {
        if (t < 64)
            for (int c = 0; c < 8; c++) store_row_chunk<FMT>(sB + act_chunk_off(t, c), B + t * 64 + c * 8);
    } else if (mode == 2 || mode == 4) {
        for (int c = 0; c < 8; c++) store_row_chunk<FMT>(sB + act_chunk_off(t, c), B + t * 64 + c * 8);
    } else {
        for (int c = 0; c < 2; c++) store_row_chunk<FMT>(sB + c * (128 * 16) + t * 16, B + t * 16 + c * 8);
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;

    long long t0 = 0;
    if (t == 0) {
        const uint32_t a = smem_u32(sA), b = smem_u32(sB);
        t0 = clock64();
        for (int rep = 0; rep < reps; rep++) {  // reps > 1: timing only (the accumulator keeps the last repetition)
            if (mode == 0) {
                const uint32_t id = idesc_f16(FMT, 128, 64, 0, 0);
                for (int ks = 0; ks < 4; ks++) mma_f16(tm, act_desc_kmajor(a, ks), act_desc_kmajor(b, ks), id, ks > 0);
            } else if (mode == 1) {
                const uint32_t id = idesc_f16(FMT, 128, 64, 0, 1);
                for (int ks = 0; ks < 4; ks++) mma_f16(tm, act_desc_kmajor(a, ks), act_desc_mnmajor(b, ks), id, ks > 0);
            } else if (mode == 2 || mode == 4) {  // mode 4: the accumulator in the upper half-subpartitions (lanes +16)
                const uint32_t id = idesc_f16(FMT, 64, 64, 1, 1);
                const uint32_t dst = mode == 4 ? tm + (16u << 16) : tm;
                for (int ks = 0; ks < 8; ks++) mma_f16(dst, act_desc_mnmajor(a, ks), act_desc_mnmajor(b, ks), id, ks > 0);
            } else {
                const uint32_t id = idesc_f16(FMT, 64, 16, 1, 1);
                for (int ks = 0; ks < 8; ks++)
                    mma_f16(tm, act_desc_mnmajor(a, ks), narrow_desc_mnmajor(b, 128, ks), id, ks > 0);
            }
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    if (t == 0 && cycles) cycles[0] = clock64() - t0;
    tc_fence_after();
    float v[64];
#pragma unroll
    for (int c = 0; c < 4; c++) tmem_ld16(tmem_addr(tm, warp * 32, c * 16), v + c * 16);
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 64; c++) out[t * 64 + c] = v[c];
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 64);
}


// Throughput / concurrency probe: `nw` warps (one per scheduler and beyond) each issue `reps` x 8 tcgen05.mma of one shape
// from their own shared-memory tiles into their own tensor-memory columns, descriptors warp-uniform as in update_tc.cu.
// cycles[0] = first issue -> last completion seen by warp 0's lane 0 after every warp's commit has arrived.
//   shape 0: M128 N64, A K-major, B K-major      (G1)        1: M64 N64, A MN-major, B MN-major (dW2)
//   shape 2: M64 N8, A MN-major, B narrow        (G3b/G4/G5)  3: M128 N8, A MN-major over two adjacent tiles, B narrow
//   shape 4: M128 N64, A K-major, B MN-major     (G2)
template <int FMT>
__global__ void __launch_bounds__(256) umma_conc_kernel(int shape, int nw, int reps, long long* __restrict__ cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ uint64_t bar[8];
    __shared__ uint32_t tmem_base_s;
    const int t = threadIdx.x, warp = t >> 5;
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    if (t == 0) {
        for (int i = 0; i < 8; i++) mbar_init(bar + i, 1);
        mbar_init_fence();
    }
    for (int i = t; i < 8 * 24576 / 4; i += 256) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;   // fp16 ones
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    long long t0 = clock64();
    if (warp < nw) {
        const uint32_t a = __shfl_sync(0xffffffffu, smem_u32(smem + warp * 24576), 0);   // 16 KB A tile (+ 8 KB B)
        const uint32_t b = a + 16384;
        const uint32_t d = tm + (uint32_t)warp * 64u;
        const bool lead = elect_one();
        for (int rep = 0; rep < reps; rep++) {
            if (shape == 0) {
                constexpr uint32_t id = idesc_f16(FMT, 128, 64, 0, 0);
                const uint64_t da = act_desc_kmajor(a, 0), db = act_desc_kmajor(b, 0);
                if (lead)
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) mma_f16(d, da + 2u * (ks & 3), db + 2u * (ks & 3), id, 1u);
            } else if (shape == 1) {
                constexpr uint32_t id = idesc_f16(FMT, 64, 64, 1, 1);
                const uint64_t da = act_desc_mnmajor(a, 0), db = act_desc_mnmajor(b, 0);
                if (lead)
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) mma_f16(d, da + 128u * (ks & 3), db + 128u * (ks & 3), id, 1u);
            } else if (shape == 2) {
                constexpr uint32_t id = idesc_f16(FMT, 64, 8, 1, 1);
                const uint64_t da = act_desc_mnmajor(a, 0), db = narrow_desc_mnmajor(b, 128, 0);
                if (lead)
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) mma_f16(d, da + 128u * ks, db + 16u * ks, id, 1u);
            } else if (shape == 3) {
                constexpr uint32_t id = idesc_f16(FMT, 128, 8, 1, 1);
                const uint64_t da = smem_desc(a, 8192, 1024, LAYOUT_SW128), db = narrow_desc_mnmajor(b, 128, 0);
                if (lead)
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) mma_f16(d, da + 128u * (ks & 3), db + 16u * ks, id, 1u);
            } else {
                constexpr uint32_t id = idesc_f16(FMT, 128, 64, 0, 1);
                const uint64_t da = act_desc_kmajor(a, 0), db = act_desc_mnmajor(b, 0);
                if (lead)
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) mma_f16(d, da + 2u * (ks & 3), db + 128u * (ks & 3), id, 1u);
            }
            __syncwarp();
        }
        if (lead) mma_commit(bar + warp);
        __syncwarp();
    }
    if (warp == 0) {
        for (int w = 0; w < nw; w++) mbar_wait(bar + w, 0);
        if (t == 0 && cycles) cycles[0] = clock64() - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}

}  // namespace ppo

static int probe_launch(int mode, int fmt, const float* A, const float* B, float* D, int reps, long long* cycles,
                        ppo_stream_t stream);
extern "C" int ppo_debug_umma_probe(int mode, int fmt, const float* A, const float* B, float* D, ppo_stream_t stream) {
    return probe_launch(mode, fmt, A, B, D, 1, nullptr, stream);
}
extern "C" int ppo_debug_umma_time(int mode, int fmt, const float* A, const float* B, float* D, int reps,
                                   long long* d_cycles, ppo_stream_t stream) {
    return probe_launch(mode, fmt, A, B, D, reps, d_cycles, stream);
}
static int probe_launch(int mode, int fmt, const float* A, const float* B, float* D, int reps, long long* cycles,
                        ppo_stream_t stream) {
    using namespace ppo;
    PPO_REQUIRE(mode >= 0 && mode <= 4 && (fmt == 0 || fmt == 1) && A && B && D, "ppo_debug_umma_probe: bad argument");
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(umma_probe_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40960));
        PPO_CUDA(cudaFuncSetAttribute(umma_probe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40960));
        attr_done = true;
    }
    if (fmt == 0) umma_probe_kernel<0><<<1, 128, 40960, as_stream(stream)>>>(mode, A, B, D, reps, cycles);
    else umma_probe_kernel<1><<<1, 128, 40960, as_stream(stream)>>>(mode, A, B, D, reps, cycles);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int ppo_debug_umma_conc(int shape, int nwarps, int reps, long long* d_cycles, ppo_stream_t stream) {
    using namespace ppo;
    PPO_REQUIRE(shape >= 0 && shape <= 4 && nwarps >= 1 && nwarps <= 8 && reps >= 1, "ppo_debug_umma_conc: bad argument");
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(umma_conc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 24576 + 1024));
        attr_done = true;
    }
    umma_conc_kernel<0><<<1, 256, 8 * 24576 + 1024, as_stream(stream)>>>(shape, nwarps, reps, d_cycles);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
