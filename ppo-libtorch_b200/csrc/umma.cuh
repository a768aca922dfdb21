// umma.cuh — thin inline-PTX layer over the sm_100a tensor-core path used by the fused PPO kernels:
// tcgen05.mma (kind::f16: bf16 or fp16 operands in shared memory, fp32 accumulators in tensor memory), TMEM
// allocation / loads, mbarriers and the proxy fences between generic shared-memory writes and the
// tensor core's reads.  No CUTLASS: descriptors are built by hand (bit layouts below).
//
// Why 16-bit operands and not TF32: a TF32 operand may only be MN-major in the 128B-swizzle-with-32B-base
// layout, a K-major one only in the 16B-based swizzles, so one stored activation could not serve both the
// forward / dX GEMM (K along the row) and the dW GEMM (K down the rows).  With 16-bit elements both views
// accept the same 128-byte swizzle, so every activation is stored ONCE (measured with ppo_debug_umma_probe).
//
// Shared-memory operand layouts used by this repo (16-bit elements):
//
//  * "activation tile" [R rows][64 elements] = one 128-byte line per row, 128-byte swizzle: element (r, c) at
//        r * 128 + (((c / 8) ^ (r % 8)) * 16) + (c % 8) * 2.
//    The SAME bytes are a valid K-major operand  [MN = r][K = c]   (G = X W^T: K runs along the row)
//    and a valid MN-major operand               [MN = c][K = r]   (G = X^T Y: K runs down the rows).
//  * "narrow tile" [R rows][8*C elements], C column chunks of 8 elements, no swizzle: element (r, c) at
//        (c / 8) * (R * 16) + r * 16 + (c % 8) * 2   — an MN-major operand with N = 8*C, K = r.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ppo {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_test(bar, parity)) {
    }
}

// ------------------------------------------------------------------ fences
// generic-proxy shared-memory writes -> visible to the async proxy (tensor core operand reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// named barrier over `count` threads (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void bar_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// ------------------------------------------------------------------ tensor memory
// One full warp allocates `cols` (power of two >= 32) columns; the base address lands in *dst (shared).
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// TMEM address = (lane << 16) | column.  A warp may only touch lanes 32*(warp % 4) .. +31.
__device__ __forceinline__ uint32_t tmem_addr(uint32_t base, uint32_t lane, uint32_t col) {
    return base + (lane << 16) + col;
}
// 32 lanes x 16 consecutive columns: thread i of the warp receives lane (lane0 + i), columns col .. col+15
__device__ __forceinline__ void tmem_ld16(uint32_t addr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
        "[%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(addr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld8(uint32_t addr, float* v) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(addr)
                 : "memory");
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ descriptors
enum : uint64_t { LAYOUT_NONE = 0, LAYOUT_SW128 = 2 };
// Shared-memory matrix descriptor (sm_100 format): start address [0,14) (>>4), leading byte offset [16,30) (>>4),
// stride byte offset [32,46) (>>4), version [46,48) = 1, layout type [61,64).
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint64_t layout) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46) | (layout << 61);
}
// Instruction descriptor, kind::f16, fp32 accumulate: c_format [4,6) = 1 (F32), a/b format [7,10)/[10,13)
// (0 = F16, 1 = BF16), a_major bit 15, b_major bit 16 (0 = K-major, 1 = MN-major), N >> 3 at [17,23), M >> 4 at [24,29).
enum { FMT_F16 = 0, FMT_BF16 = 1 };
__host__ __device__ constexpr uint32_t idesc_f16(int fmt, int M, int N, int a_mn_major, int b_mn_major) {
    return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)a_mn_major << 15) |
           ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem], K = 16 per instruction; issued by ONE thread.  accumulate == 0 overwrites D.
__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// true in exactly one (elected) lane of a converged warp
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// mbarrier arrive once every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// ------------------------------------------------------------------ activation-tile addressing
// byte offset of the 16-byte chunk holding columns 8*chunk .. 8*chunk+7 of row r
__device__ __forceinline__ uint32_t act_chunk_off(int r, int chunk) {
    return (uint32_t)(r * 128 + ((chunk ^ (r & 7)) << 4));
}
// K-major view  [MN = rows][K = 64 cols]: K-step ks (16 elements = 32 B), ks = 0..3
__device__ __forceinline__ uint64_t act_desc_kmajor(uint32_t base, int ks) {
    return smem_desc(base + (uint32_t)ks * 32u, 16, 1024, LAYOUT_SW128);
}
// MN-major view [MN = 64 cols][K = rows]: K-step ks covers rows 16*ks .. 16*ks+15 (two 1024-byte swizzle atoms)
__device__ __forceinline__ uint64_t act_desc_mnmajor(uint32_t base, int ks) {
    return smem_desc(base + (uint32_t)ks * 2048u, 8192, 1024, LAYOUT_SW128);
}
// narrow tile, MN-major [MN = 8*C cols][K = rows], no swizzle: core matrix = 8 rows x 16 B; R rows per chunk
__device__ __forceinline__ uint64_t narrow_desc_mnmajor(uint32_t base, int R, int ks) {
    return smem_desc(base + (uint32_t)ks * 256u, 128, (uint32_t)(R * 16), LAYOUT_NONE);
}

// narrow tile viewed K-major [MN = rows][K = 16], no swizzle: K chunk 0 at `base` (rows 16 B apart, 8-row core matrices
// 128 B apart), K chunk 1 `chunk1_off` bytes further on (a second chunk of the same shape)
__device__ __forceinline__ uint64_t narrow_desc_kmajor(uint32_t base, uint32_t chunk1_off) {
    return smem_desc(base, chunk1_off, 128, LAYOUT_NONE);
}

// two fp32 -> one packed 16-bit pair (lo = first element)
template <int FMT>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    uint32_t r;
    if constexpr (FMT == FMT_BF16) asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    else asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));  // saturate, never inf
    return r;
}
template <int FMT>
__device__ __forceinline__ float unpack_lo(uint32_t v) {
    if constexpr (FMT == FMT_BF16) return __uint_as_float(v << 16);
    else { float f; asm("{.reg .b16 l, h; mov.b32 {l, h}, %1; cvt.f32.f16 %0, l;}" : "=f"(f) : "r"(v)); return f; }
}
template <int FMT>
__device__ __forceinline__ float unpack_hi(uint32_t v) {
    if constexpr (FMT == FMT_BF16) return __uint_as_float(v & 0xFFFF0000u);
    else { float f; asm("{.reg .b16 l, h; mov.b32 {l, h}, %1; cvt.f32.f16 %0, h;}" : "=f"(f) : "r"(v)); return f; }
}

// tanh derivative on packed 16-bit pairs: d * (1 - h^2) for two elements in three instructions (pack d, fused
// -h*h+1, multiply) instead of unpacking h, two fp32 FMAs, two multiplies and the pack.  Each result carries the same
// single 16-bit rounding it gets when it is stored as a GEMM operand anyway.
template <int FMT>
__device__ __forceinline__ uint32_t dtanh2(float d_lo, float d_hi, uint32_t h) {
    const uint32_t d = pack2<FMT>(d_lo, d_hi);
    uint32_t nh, g, r;   // neg (not an xor) so that ptxas folds it into the HFMA2 operand modifier
    if constexpr (FMT == FMT_BF16) {
        asm("neg.bf16x2 %0, %1;" : "=r"(nh) : "r"(h));
        asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(g) : "r"(nh), "r"(h), "r"(0x3F803F80u));
        asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(d), "r"(g));
    } else {
        asm("neg.f16x2 %0, %1;" : "=r"(nh) : "r"(h));
        asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(g) : "r"(nh), "r"(h), "r"(0x3C003C00u));
        asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(d), "r"(g));
    }
    return r;
}

// packed 16-bit pair arithmetic (one instruction for two elements)
template <int FMT>
__device__ __forceinline__ uint32_t hmul2(uint32_t a, uint32_t b) {
    uint32_t r;
    if constexpr (FMT == FMT_BF16) asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    else asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
template <int FMT>
__device__ __forceinline__ uint32_t hfma2(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    if constexpr (FMT == FMT_BF16) asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    else asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
// dtanh2 for an upstream gradient that is already a packed pair: d * (1 - h^2)
template <int FMT>
__device__ __forceinline__ uint32_t dtanh2p(uint32_t d, uint32_t h) {
    uint32_t nh;
    if constexpr (FMT == FMT_BF16) asm("neg.bf16x2 %0, %1;" : "=r"(nh) : "r"(h));
    else asm("neg.f16x2 %0, %1;" : "=r"(nh) : "r"(h));
    const uint32_t one = FMT == FMT_BF16 ? 0x3F803F80u : 0x3C003C00u;
    return hmul2<FMT>(d, hfma2<FMT>(nh, h, one));
}

}  // namespace umma
}  // namespace ppo
