// update_tc.cu — K5 on the sm_100a tensor cores: fused minibatch forward + PPO loss + backward for the
// 2 x 64 actor / critic MLPs (PPO/Agent.cpp:25-59), replacing PPO/PPO_Discrete.cpp:579-638.
//
// The three 64 x 64 GEMMs of each net (forward y = x W2^T, backward dX = dY W2, backward dW2 = dY^T X) and the
// skinny weight-gradient GEMMs (dW1|db1, db2, dW3^T) run as tcgen05.mma (kind::f16: bf16 or fp16 operands,
// fp32 accumulation in tensor memory); everything per-sample (layer 1, bias + tanh, the output layer, softmax /
// ratio / clipping / value loss, the tanh derivatives) runs one-thread-per-row on the CUDA cores.
//
// CTA = NWG warpgroups (128 threads = 128 minibatch rows each).  A warpgroup walks its 128-row tiles; per tile and
// net it writes four operand stages to shared memory, and after a warpgroup-wide named barrier its thread 0 issues
// the stage's MMAs (measured: one issuer warp serving all warpgroups was itself the bottleneck, ~1100 dependent
// instructions per tile).  Completions come back through two mbarriers per warpgroup ("doneZ": an accumulator it must
// read; "doneW": operands it may overwrite):
//   S1  h1 = tanh([x,1] P1^T)            -> A1          G1 : Z   = A1 W2^T                        (doneZ)
//   S2  h2 = tanh(Z + b2) -> A2, logits/value, loss, dOut -> DO
//                                                        G5 : dW3^T += A2^T DO                     (doneW)
//   S3  dZ2 = (dOut W3) (1 - h2^2)       -> A3          G2 : Z   = A3 W2                          (doneZ)
//                                                        G3 : dW2 += A3^T A1 ; d[.|b2] += A3^T XE   (doneW)
//   S4  dZ1 = Z (1 - h1^2)               -> A2          G4 : d[W1|b1] += A2^T XE                  (doneW)
// While one warpgroup waits on the tensor core another computes, so the two pipes overlap without a software
// pipeline inside a warpgroup.  All weight-gradient accumulators stay in tensor memory for the whole kernel (private
// to a warpgroup, one issuing thread each => in-order accumulation) and are drained once into the CTA's slab; reduce_kernel (update.cu) sums
// the slabs in fixed order, so results are run-to-run identical.  Gradients are kept un-normalised (no 1/n) until
// the drain so that fp16 operands do not underflow.
#include "ppo_loss.cuh"
#include "umma.cuh"
#include "update.cuh"

namespace ppo {
using namespace umma;

constexpr int TC_H = 64;
constexpr int TC_ROWS = 128;        // rows per warpgroup tile
constexpr int TC_MAX_K = 24;        // sum of logits: 2 x (168 + pad8(K)) + 128 tensor-memory columns must fit in 512
constexpr int ACT_BYTES = TC_ROWS * 128;  // one activation tile

struct TcLayout {   // byte offsets into the 1024-aligned dynamic shared memory
    int w2[2];      // bf16/fp16 W2 of actor (0) and critic (1), 64 rows x 128 B each
    int wg0, wg_stride;                  // per-warpgroup block
    int a1, a2, a3, xe, dout, lg, b3;    // offsets inside a warpgroup block
    int p1[2], b2[2], w3[2], b3p[2];     // fp32 small parameters per net
    int bars, tmem_slot, end;
    int ldo, n3a;                        // logits row stride (floats), padded actor logit count
};
__host__ __device__ inline TcLayout tc_layout(int K, int nwg) {
    TcLayout L;
    L.ldo = K | 1;
    L.n3a = (K + 7) & ~7;
    L.w2[0] = 0;
    L.w2[1] = 8192;
    L.a1 = 0;
    L.a2 = ACT_BYTES;
    L.a3 = 2 * ACT_BYTES;
    L.xe = 3 * ACT_BYTES;
    L.dout = L.xe + TC_ROWS * 16;
    L.lg = L.dout + (L.n3a / 8) * TC_ROWS * 16;
    L.b3 = L.lg + ((TC_ROWS * L.ldo * 4 + 15) & ~15);
    L.wg_stride = (L.b3 + ((TC_ROWS * L.ldo * 4 + 15) & ~15) + 1023) & ~1023;
    L.wg0 = 16384;
    int o = L.wg0 + nwg * L.wg_stride;
    for (int n = 0; n < 2; n++) {
        const int k3 = n == 0 ? K : 1;
        L.p1[n] = o; o += TC_H * 8 * 4;
        L.b2[n] = o; o += TC_H * 4;
        L.w3[n] = o; o += k3 * TC_H * 4;
        L.b3p[n] = o; o += ((k3 + 3) & ~3) * 4;
    }
    L.bars = o; o += 3 * 4 * 8;
    L.tmem_slot = o; o += 16;
    L.end = o;
    return L;
}

__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <int FMT>
__device__ __forceinline__ void store8(uint8_t* dst, const float* v) {
    uint4 u;
    u.x = pack2<FMT>(v[0], v[1]); u.y = pack2<FMT>(v[2], v[3]);
    u.z = pack2<FMT>(v[4], v[5]); u.w = pack2<FMT>(v[6], v[7]);
    *reinterpret_cast<uint4*>(dst) = u;
}
template <int FMT>
__device__ __forceinline__ void load8(const uint8_t* src, float* v) {
    const uint4 u = *reinterpret_cast<const uint4*>(src);
    v[0] = unpack_lo<FMT>(u.x); v[1] = unpack_hi<FMT>(u.x); v[2] = unpack_lo<FMT>(u.y); v[3] = unpack_hi<FMT>(u.y);
    v[4] = unpack_lo<FMT>(u.z); v[5] = unpack_hi<FMT>(u.z); v[6] = unpack_lo<FMT>(u.w); v[7] = unpack_hi<FMT>(u.w);
}

// tensor-memory column map: Z of warpgroup g at 64 g; then one block of gradient accumulators (M = 64 layout) PER
// warpgroup, so that every accumulator sees its MMAs in a fixed order (run-to-run identical sums):
//   actor {dW2 64, d[.|b2] 8, d[W1|b1] 8, dW3^T n3a}, critic {64, 8, 8, 8}  =>  168 + n3a columns per warpgroup
struct TcCols {
    uint32_t w2, xe2, xe1, w3;
};
__device__ __forceinline__ TcCols tc_cols(int nwg, int g, int net, int n3a) {
    const uint32_t base = (uint32_t)nwg * 64u + (uint32_t)g * (168u + (uint32_t)n3a) + (net == 0 ? 0u : 80u + (uint32_t)n3a);
    return TcCols{base, base + 64u, base + 72u, base + 80u};
}

// ---- MMA issue: thread 0 of a warpgroup issues that warpgroup's stage (its accumulators are private, so no ordering
// between the issuing threads of different warpgroups is needed).  Descriptors are precomputed; a K-step only adds a
// constant to the start-address field (16-byte units).
struct WgDescs {
    uint64_t a1k, a1m, a2m, a3k, a3m, xe, dout;  // activation tiles (K-major / MN-major views), narrow tiles
    uint64_t w2k[2], w2m[2];                      // W2 of actor / critic
    uint32_t Z;                                   // tensor-memory address of this warpgroup's Z block
    TcCols c[2];
    uint32_t id_w3[2];
    uint32_t fresh;                               // bit (net * 4 + accumulator): next MMA into it must overwrite
};
template <int FMT, int ST>
__device__ __forceinline__ void issue_stage(WgDescs& d, int net, uint32_t tmem, uint64_t* doneZ, uint64_t* doneW) {
    constexpr uint32_t id_fwd = idesc_f16(FMT, 128, 64, 0, 0);   // Z = A1 W2^T
    constexpr uint32_t id_dx = idesc_f16(FMT, 128, 64, 0, 1);    // Z = A3 W2
    constexpr uint32_t id_dw = idesc_f16(FMT, 64, 64, 1, 1);     // dW2 += A3^T A1
    constexpr uint32_t id_n8 = idesc_f16(FMT, 64, 8, 1, 1);      // skinny, N = 8
    const uint32_t fb = (uint32_t)net * 4u;
    const TcCols c = d.c[net];
    tc_fence_after();
    if (ST == 0) {
#pragma unroll
        for (int ks = 0; ks < 4; ks++) mma_f16(d.Z, d.a1k + 2u * ks, d.w2k[net] + 2u * ks, id_fwd, ks > 0);
        mma_commit(doneZ);
    } else if (ST == 1) {
        const uint32_t acc0 = (d.fresh >> (fb + 3)) & 1u ? 0u : 1u;
#pragma unroll
        for (int ks = 0; ks < 8; ks++)
            mma_f16(tmem + c.w3, d.a2m + 128u * ks, d.dout + 16u * ks, d.id_w3[net], ks > 0 ? 1u : acc0);
        d.fresh &= ~(1u << (fb + 3));
        mma_commit(doneW);
    } else if (ST == 2) {
#pragma unroll
        for (int ks = 0; ks < 4; ks++) mma_f16(d.Z, d.a3k + 2u * ks, d.w2m[net] + 128u * ks, id_dx, ks > 0);
        mma_commit(doneZ);
        const uint32_t acc0 = (d.fresh >> fb) & 1u ? 0u : 1u;
#pragma unroll
        for (int ks = 0; ks < 8; ks++) {
            mma_f16(tmem + c.w2, d.a3m + 128u * ks, d.a1m + 128u * ks, id_dw, ks > 0 ? 1u : acc0);
            mma_f16(tmem + c.xe2, d.a3m + 128u * ks, d.xe + 16u * ks, id_n8, ks > 0 ? 1u : acc0);
        }
        d.fresh &= ~(3u << fb);
        mma_commit(doneW);
    } else {
        const uint32_t acc0 = (d.fresh >> (fb + 2)) & 1u ? 0u : 1u;
#pragma unroll
        for (int ks = 0; ks < 8; ks++) mma_f16(tmem + c.xe1, d.a2m + 128u * ks, d.xe + 16u * ks, id_n8, ks > 0 ? 1u : acc0);
        d.fresh &= ~(1u << (fb + 2));
        mma_commit(doneW);
    }
}

// ---- per-row pieces of a compute warpgroup --------------------------------------------------------------------
// layer 1 for 8 consecutive hidden units j0..j0+7: h = tanh(sum_i xe[i] P1[j][i]); P1 row = {W1[j][0..O), b1[j], 0..}
__device__ __forceinline__ void layer1_chunk(const float* P1, const float (&xe)[8], bool wide, int j0, float* h) {
#pragma unroll
    for (int jj = 0; jj < 8; jj++) {
        const float4 w0 = *reinterpret_cast<const float4*>(P1 + (j0 + jj) * 8);
        float acc = xe[0] * w0.x;
        acc = fmaf(xe[1], w0.y, acc); acc = fmaf(xe[2], w0.z, acc); acc = fmaf(xe[3], w0.w, acc);
        if (wide) {
            const float4 w1 = *reinterpret_cast<const float4*>(P1 + (j0 + jj) * 8 + 4);
            acc = fmaf(xe[4], w1.x, acc); acc = fmaf(xe[5], w1.y, acc);
            acc = fmaf(xe[6], w1.z, acc); acc = fmaf(xe[7], w1.w, acc);
        }
        h[jj] = tanh_fast(acc);
    }
}

template <int FMT, int NWG>
__global__ void __launch_bounds__(NWG * 128, 1) update_tc_kernel(const UpdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int O = a.d.O, K = a.d.K, heads = a.d.heads;
    const TcLayout L = tc_layout(K, NWG);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const NetOff offs[2] = {actor_off(O, TC_H, K), critic_off(O, TC_H)};
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);  // doneZ[NWG + g], doneW[2 NWG + g]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const int ntiles = (a.mb_size + TC_ROWS - 1) / TC_ROWS;
    const int nthreads = NWG * 128;

    // ---------------- prologue: tensor memory, barriers, parameters -> shared memory -------------------------
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    if (tid == 0) {
        for (int g = 0; g < NWG; g++) {
            mbar_init(bars + NWG + g, 1);
            mbar_init(bars + 2 * NWG + g, 1);
        }
        mbar_init_fence();
    }
    {   // W2 of both nets as 16-bit activation-layout tiles: 1024 chunk tasks, all loads of a thread in flight together
        constexpr int W2_IT = (2 * 64 * 8 + NWG * 128 - 1) / (NWG * 128);
        float v[W2_IT][8];
#pragma unroll
        for (int it = 0; it < W2_IT; it++) {
            const int e = tid + it * nthreads;
            if (e < 2 * 64 * 8) {
                const float* src = a.params + offs[e >> 9].W2 + ((e >> 3) & 63) * TC_H + (e & 7) * 8;
#pragma unroll
                for (int i = 0; i < 8; i++) v[it][i] = __ldg(src + i);
            }
        }
#pragma unroll
        for (int it = 0; it < W2_IT; it++) {
            const int e = tid + it * nthreads;
            if (e < 2 * 64 * 8) store8<FMT>(sm + L.w2[e >> 9] + act_chunk_off((e >> 3) & 63, e & 7), v[it]);
        }
    }
    {   // fp32 small parameters: P1 = {W1 row, b1, 0..} [64][8], b2 [64], W3 [k3][64], b3 [k3] per net, one flat task list
        const int S1 = 2 * 512, S2 = S1 + 128, S3 = S2 + 64 * (K + 1), S4 = S3 + (K + 1);
        for (int e0 = tid; e0 < S4; e0 += 4 * nthreads) {
            float v[4];
            float* dst[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int e = e0 + u * nthreads;
                dst[u] = nullptr;
                v[u] = 0.0f;
                if (e >= S4) continue;
                const float* src = nullptr;
                if (e < S1) {
                    const int net = e >> 9, j = (e >> 3) & 63, i = e & 7;
                    dst[u] = reinterpret_cast<float*>(sm + L.p1[net]) + (e & 511);
                    if (i < O) src = a.params + offs[net].W1 + j * O + i;
                    else if (i == O) src = a.params + offs[net].b1 + j;
                } else if (e < S2) {
                    const int t = e - S1, net = t >> 6;
                    dst[u] = reinterpret_cast<float*>(sm + L.b2[net]) + (t & 63);
                    src = a.params + offs[net].b2 + (t & 63);
                } else if (e < S3) {
                    const int t = e - S2, net = t < 64 * K ? 0 : 1, idx = net ? t - 64 * K : t;
                    dst[u] = reinterpret_cast<float*>(sm + L.w3[net]) + idx;
                    src = a.params + offs[net].W3 + idx;
                } else {
                    const int t = e - S3, net = t < K ? 0 : 1, idx = net ? t - K : t;
                    dst[u] = reinterpret_cast<float*>(sm + L.b3p[net]) + idx;
                    src = a.params + offs[net].b3 + idx;
                }
                if (src) v[u] = __ldg(src);
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (dst[u]) *dst[u] = v[u];
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    RowSums rs;
    float gb3c = 0.0f;  // critic output-bias gradient partial of this thread
    {
        // ---------------- compute warpgroup ------------------------------------------------------------------
        const int g = warp >> 2, wt = tid & 127, w4 = warp & 3;
        uint8_t* wg = sm + L.wg0 + g * L.wg_stride;
        uint8_t *A1 = wg + L.a1, *A2 = wg + L.a2, *A3 = wg + L.a3, *XE = wg + L.xe, *DO = wg + L.dout;
        float* x = reinterpret_cast<float*>(wg + L.lg) + wt * L.ldo;     // this row's logits / dL/dlogits
        float* b3acc = reinterpret_cast<float*>(wg + L.b3) + wt * L.ldo; // this thread's actor db3 partials
        for (int q = 0; q < K; q++) b3acc[q] = 0.0f;
        uint64_t *doneZ = bars + NWG + g, *doneW = bars + 2 * NWG + g;
        WgDescs D;
        {
            const uint32_t wb = smem_u32(wg);
            D.a1k = act_desc_kmajor(wb + L.a1, 0); D.a1m = act_desc_mnmajor(wb + L.a1, 0);
            D.a2m = act_desc_mnmajor(wb + L.a2, 0);
            D.a3k = act_desc_kmajor(wb + L.a3, 0); D.a3m = act_desc_mnmajor(wb + L.a3, 0);
            D.xe = narrow_desc_mnmajor(wb + L.xe, TC_ROWS, 0); D.dout = narrow_desc_mnmajor(wb + L.dout, TC_ROWS, 0);
            for (int n = 0; n < 2; n++) {
                D.w2k[n] = act_desc_kmajor(smem_u32(sm) + L.w2[n], 0);
                D.w2m[n] = act_desc_mnmajor(smem_u32(sm) + L.w2[n], 0);
                D.c[n] = tc_cols(NWG, g, n, L.n3a);
            }
            D.Z = tmem + (uint32_t)g * 64u;
            D.id_w3[0] = idesc_f16(FMT, 64, L.n3a, 1, 1);
            D.id_w3[1] = idesc_f16(FMT, 64, 8, 1, 1);
            D.fresh = 0xFFu;
        }
        const int barid = 1 + g;
        uint32_t pz = 0, pw = 0;
        bool pendingW = false;
        const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)g * 64u);
        const bool masked = a.mask != nullptr;
        const bool wide = O + 1 > 4;
        float meanf = 0.0f, stdf = 1.0f;
        if (a.hp.norm_adv) { meanf = a.adv_stats[0]; stdf = a.adv_stats[1]; }
        const int nwg_total = gridDim.x * NWG;

        // per-row inputs of a tile, gathered through the permutation one tile ahead (the loads of tile t+1 fly while
        // tile t computes)
        struct RowIn {
            int64_t r;
            float o[4];      // obs (O <= 4: prefetched; wider observations are loaded at use)
            float old_logp, adv, ret, old_val;
            bool on;
        };
        auto gather = [&](int tile) {
            RowIn in;
            const int rloc = tile * TC_ROWS + wt;
            in.on = tile < ntiles && rloc < a.mb_size;
            in.r = 0;
            in.o[0] = in.o[1] = in.o[2] = in.o[3] = 0.0f;
            in.old_logp = in.adv = in.ret = in.old_val = 0.0f;
            if (in.on) {
                const int64_t r = a.perm ? (int64_t)__ldg(a.perm + a.mb_start + rloc) : a.mb_start + rloc;
                in.r = r;
                if (O == 4) {
                    const float4 o4 = __ldg(reinterpret_cast<const float4*>(a.obs + r * 4));
                    in.o[0] = o4.x; in.o[1] = o4.y; in.o[2] = o4.z; in.o[3] = o4.w;
                } else if (O < 4) {
#pragma unroll
                    for (int i = 0; i < 3; i++)
                        if (i < O) in.o[i] = __ldg(a.obs + r * O + i);
                }
                in.old_logp = __ldg(a.old_logprob + r);
                in.adv = __ldg(a.adv + r);
                in.ret = __ldg(a.ret + r);
                if (a.hp.clip_vloss) in.old_val = __ldg(a.old_val + r);
            }
            return in;
        };
        RowIn nxt = gather(blockIdx.x * NWG + g);
        for (int tile = blockIdx.x * NWG + g; tile < ntiles; tile += nwg_total) {
            const RowIn cur = nxt;
            nxt = gather(tile + nwg_total);
            const bool on = cur.on;
            const int64_t r = cur.r;
            float xe[8];
#pragma unroll
            for (int i = 0; i < 8; i++) xe[i] = 0.0f;
            if (O <= 4) {
#pragma unroll
                for (int i = 0; i < 4; i++) xe[i] = cur.o[i];
            } else if (on) {
#pragma unroll
                for (int i = 0; i < 7; i++)
                    if (i < O) xe[i] = __ldg(a.obs + r * O + i);
            }
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (i == O) xe[i] = 1.0f;

            for (int net = 0; net < 2; net++) {
                const int k3 = net == 0 ? K : 1;
                const float* P1 = reinterpret_cast<const float*>(sm + L.p1[net]);
                const float* B2 = reinterpret_cast<const float*>(sm + L.b2[net]);
                const float* W3 = reinterpret_cast<const float*>(sm + L.w3[net]);
                const float* B3 = reinterpret_cast<const float*>(sm + L.b3p[net]);
                if (pendingW) {  // G4 of the previous pass: every operand buffer of this warpgroup is free again
                    mbar_wait(doneW, pw); pw ^= 1u; pendingW = false;
                }
                if (net == 0) store8<FMT>(XE + wt * 16, xe);
                // ---- S1: layer 1 -> A1
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    float h[8];
                    layer1_chunk(P1, xe, wide, c * 8, h);
                    store8<FMT>(A1 + act_chunk_off(wt, c), h);
                }
                fence_async_smem();
                bar_sync(barid, 128);
                if (wt == 0) issue_stage<FMT, 0>(D, net, tmem, doneZ, doneW);
                mbar_wait(doneZ, pz); pz ^= 1u;
                tc_fence_after();
                // ---- S2: h2 = tanh(Z + b2) -> A2; output layer; loss; dOut -> DO
                float h2[64];
#pragma unroll
                for (int c = 0; c < 4; c++) tmem_ld16(zaddr + c * 16, h2 + c * 16);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 16; c++) {
                    const float4 b = *reinterpret_cast<const float4*>(B2 + c * 4);
                    h2[c * 4 + 0] = tanh_fast(h2[c * 4 + 0] + b.x); h2[c * 4 + 1] = tanh_fast(h2[c * 4 + 1] + b.y);
                    h2[c * 4 + 2] = tanh_fast(h2[c * 4 + 2] + b.z); h2[c * 4 + 3] = tanh_fast(h2[c * 4 + 3] + b.w);
                }
#pragma unroll
                for (int c = 0; c < 8; c++) store8<FMT>(A2 + act_chunk_off(wt, c), h2 + c * 8);
                float gv = 0.0f;
                for (int q = 0; q < k3; q++) {
                    float acc = B3[q];
#pragma unroll
                    for (int c = 0; c < 16; c++) {
                        const float4 w = *reinterpret_cast<const float4*>(W3 + q * TC_H + c * 4);
                        acc = fmaf(h2[c * 4 + 0], w.x, acc); acc = fmaf(h2[c * 4 + 1], w.y, acc);
                        acc = fmaf(h2[c * 4 + 2], w.z, acc); acc = fmaf(h2[c * 4 + 3], w.w, acc);
                    }
                    if (net == 0) x[q] = acc; else gv = acc;
                }
                if (net == 0) {
                    if (on) {
                        const uint64_t mbits = masked ? pack_mask(a.mask + r * K, K) : ~0ull;
                        actor_row_loss(x, a.d, masked, mbits, a.actions + r * heads, cur.old_logp, cur.adv, a.hp, meanf, stdf,
                                       1.0f, rs);
                        for (int q = 0; q < K; q++) b3acc[q] += x[q];
                    } else {
                        for (int q = 0; q < K; q++) x[q] = 0.0f;
                    }
                    for (int c = 0; c < L.n3a / 8; c++) {
                        float v[8];
#pragma unroll
                        for (int i = 0; i < 8; i++) v[i] = (c * 8 + i < K) ? x[c * 8 + i] : 0.0f;
                        store8<FMT>(DO + c * (TC_ROWS * 16) + wt * 16, v);
                    }
                } else {
                    gv = on ? critic_row_loss(gv, cur.ret, cur.old_val, a.hp, 1.0f, rs) : 0.0f;
                    gb3c += gv;
                    float v[8] = {gv, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    store8<FMT>(DO + wt * 16, v);
                }
                tc_fence_before();
                fence_async_smem();
                bar_sync(barid, 128);
                if (wt == 0) issue_stage<FMT, 1>(D, net, tmem, doneZ, doneW);
                // ---- S3: dZ2 = (dOut W3) * (1 - h2^2) -> A3
                float d[64];
#pragma unroll
                for (int k = 0; k < 64; k++) d[k] = 0.0f;
                for (int q = 0; q < k3; q++) {
                    const float dq = net == 0 ? x[q] : gv;
#pragma unroll
                    for (int c = 0; c < 16; c++) {
                        const float4 w = *reinterpret_cast<const float4*>(W3 + q * TC_H + c * 4);
                        d[c * 4 + 0] = fmaf(dq, w.x, d[c * 4 + 0]); d[c * 4 + 1] = fmaf(dq, w.y, d[c * 4 + 1]);
                        d[c * 4 + 2] = fmaf(dq, w.z, d[c * 4 + 2]); d[c * 4 + 3] = fmaf(dq, w.w, d[c * 4 + 3]);
                    }
                }
#pragma unroll
                for (int k = 0; k < 64; k++) d[k] *= fmaf(-h2[k], h2[k], 1.0f);
                mbar_wait(doneW, pw); pw ^= 1u;  // G5 has read A2 / DO
#pragma unroll
                for (int c = 0; c < 8; c++) store8<FMT>(A3 + act_chunk_off(wt, c), d + c * 8);
                fence_async_smem();
                bar_sync(barid, 128);
                if (wt == 0) issue_stage<FMT, 2>(D, net, tmem, doneZ, doneW);
                mbar_wait(doneZ, pz); pz ^= 1u;
                tc_fence_after();
                // ---- S4: dZ1 = Z * (1 - h1^2) -> A2 (h2's buffer; G5 is done with it)
#pragma unroll
                for (int c = 0; c < 4; c++) tmem_ld16(zaddr + c * 16, d + c * 16);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    float h1[8];
                    load8<FMT>(A1 + act_chunk_off(wt, c), h1);
#pragma unroll
                    for (int i = 0; i < 8; i++) d[c * 8 + i] *= fmaf(-h1[i], h1[i], 1.0f);
                    store8<FMT>(A2 + act_chunk_off(wt, c), d + c * 8);
                }
                mbar_wait(doneW, pw); pw ^= 1u;  // G3 has read A1 / A3 (keeps doneW in lock-step)
                tc_fence_before();
                fence_async_smem();
                bar_sync(barid, 128);
                if (wt == 0) issue_stage<FMT, 3>(D, net, tmem, doneZ, doneW);
                pendingW = true;
            }
        }
        if (pendingW) { mbar_wait(doneW, pw); pw ^= 1u; }
    }

    // ---------------- drain: tensor-memory accumulators -> this CTA's slab (x 1/n), diagnostics ----------------
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const int P = num_params(O, TC_H, K);
    float* slab = a.slabs + (size_t)blockIdx.x * slab_stride(P);
    const float sc = a.inv_n;
    if (warp < 4) {
        // M = 64 accumulators: row m sits in lane (m % 16) + 32 (m / 16)  =>  lanes 0..15 of warp w hold rows 16 w + lane
        const int n = 16 * warp + lane;
        const bool has = lane < 16;
        const uint32_t la = tmem_addr(tmem, (uint32_t)warp * 32u, 0);
        // warpgroups that owned no tile never wrote their accumulators: skip them (same test as the issuer's)
        int ng = 0;
        for (int g = 0; g < NWG; g++)
            if (blockIdx.x * NWG + g < ntiles) ng = g + 1;
        for (int net = 0; net < 2; net++) {
            const NetOff& of = offs[net];
            const int k3 = net == 0 ? K : 1;
            float v[64], w[64];
#pragma unroll
            for (int g = 0; g < NWG; g++) {
                if (g >= ng) break;
                const TcCols c = tc_cols(NWG, g, net, L.n3a);
#pragma unroll
                for (int cc = 0; cc < 4; cc++) tmem_ld16(la + c.w2 + cc * 16, (g == 0 ? v : w) + cc * 16);
                tmem_ld_wait();
                if (g > 0)
#pragma unroll
                    for (int k = 0; k < 64; k++) v[k] += w[k];
            }
            if (has)
#pragma unroll
                for (int k = 0; k < 64; k++) slab[of.W2 + n * TC_H + k] = v[k] * sc;
            float e[8], f[8];
#pragma unroll
            for (int g = 0; g < NWG; g++) {
                if (g >= ng) break;
                tmem_ld8(la + tc_cols(NWG, g, net, L.n3a).xe2, g == 0 ? e : f);
                tmem_ld_wait();
                if (g > 0)
#pragma unroll
                    for (int i = 0; i < 8; i++) e[i] += f[i];
            }
            if (has)
#pragma unroll
                for (int i = 0; i < 8; i++)
                    if (i == O) slab[of.b2 + n] = e[i] * sc;
#pragma unroll
            for (int g = 0; g < NWG; g++) {
                if (g >= ng) break;
                tmem_ld8(la + tc_cols(NWG, g, net, L.n3a).xe1, g == 0 ? e : f);
                tmem_ld_wait();
                if (g > 0)
#pragma unroll
                    for (int i = 0; i < 8; i++) e[i] += f[i];
            }
            if (has)
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (i < O) slab[of.W1 + n * O + i] = e[i] * sc;
                    if (i == O) slab[of.b1 + n] = e[i] * sc;
                }
            for (int cc = 0; cc < ((k3 + 7) >> 3); cc++) {
#pragma unroll
                for (int g = 0; g < NWG; g++) {
                    if (g >= ng) break;
                    tmem_ld8(la + tc_cols(NWG, g, net, L.n3a).w3 + cc * 8, g == 0 ? e : f);
                    tmem_ld_wait();
                    if (g > 0)
#pragma unroll
                        for (int i = 0; i < 8; i++) e[i] += f[i];
                }
                if (has)
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        if (cc * 8 + i < k3) slab[of.W3 + (cc * 8 + i) * TC_H + n] = e[i] * sc;
            }
        }
    }
    constexpr int NT = NWG * 128;
    double* red = reinterpret_cast<double*>(sm + L.wg0);                    // [5][NT]   (A1/A2 of warpgroup 0)
    float* gbc = reinterpret_cast<float*>(sm + L.wg0 + 5 * NT * 8);         // [NT]
    {
        red[0 * NT + tid] = rs.pg; red[1 * NT + tid] = rs.v; red[2 * NT + tid] = rs.ent;
        red[3 * NT + tid] = rs.kl; red[4 * NT + tid] = rs.clip;
        gbc[tid] = gb3c;
    }
    __syncthreads();
    // one warp per quantity: lane-strided partial sums, then a fixed xor tree (deterministic)
    if (warp < 5) {
        double acc = 0.0;
        for (int m = lane; m < NT; m += 32) acc += red[warp * NT + m];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) a.dpart[(size_t)blockIdx.x * UPD_NDIAG + warp] = acc;
    } else if (warp == 5) {
        float acc = 0.0f;
        for (int m = lane; m < NT; m += 32) acc += gbc[m];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) slab[offs[1].b3] = acc * sc;
    } else {
        for (int q = warp - 6; q < K; q += NWG * 4 - 6) {
            float acc = 0.0f;
            for (int g = 0; g < NWG; g++) {
                const float* b = reinterpret_cast<const float*>(sm + L.wg0 + g * L.wg_stride + L.b3);
                for (int m = lane; m < TC_ROWS; m += 32) acc += b[m * L.ldo + q];
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) slab[offs[0].b3 + q] = acc * sc;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int FMT, int NWG>
static int launch_tc(const UpdArgs& a, int* grid_out, cudaStream_t st) {
    const TcLayout L = tc_layout(a.d.K, NWG);
    const size_t smem = (size_t)L.end + 1024;
    PPO_REQUIRE(smem <= 227 * 1024, "update_tc: needs %zu B of shared memory", smem);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(update_tc_kernel<FMT, NWG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done = true;
    }
    const int ntiles = (a.mb_size + TC_ROWS - 1) / TC_ROWS;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int grid = (ntiles + NWG - 1) / NWG;
    if (grid > sms) grid = sms;
    if (grid > UPD_MAX_GRID) grid = UPD_MAX_GRID;
    update_tc_kernel<FMT, NWG><<<grid, NWG * 128, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    *grid_out = grid;
    return PPO_OK;
}

int launch_update_tc(const UpdArgs& a, int fmt, int* grid_out, cudaStream_t st) {
    if (a.d.H != TC_H || a.d.O + 1 > 8 || a.d.K > TC_MAX_K) return PPO_ERR_UNSUPPORTED;
    if (fmt == FMT_BF16) return launch_tc<FMT_BF16, 2>(a, grid_out, st);
    return launch_tc<FMT_F16, 2>(a, grid_out, st);
}

}  // namespace ppo
