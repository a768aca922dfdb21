// update_tc.cu — K5 on the sm_100a tensor cores: fused minibatch forward + PPO loss + backward for the
// 2 x 64 actor / critic MLPs (PPO/Agent.cpp:25-59), replacing PPO/PPO_Discrete.cpp:579-638.
//
// The three 64 x 64 GEMMs of each net (forward y = x W2^T, backward dX = dY W2, backward dW2 = dY^T X) and the
// skinny weight-gradient GEMMs (dW1|db1, db2, dW3^T) run as tcgen05.mma (kind::f16: bf16 or fp16 operands,
// fp32 accumulation in tensor memory); everything per-sample (layer 1, bias + tanh, the output layer, softmax /
// ratio / clipping / value loss, the tanh derivatives) runs one-thread-per-row on the CUDA cores.
//
// CTA = NWG warpgroups (128 threads = 128 minibatch rows each).  A warpgroup walks its 128-row tiles; per tile it
// runs four phases, each over BOTH nets (actor, then critic), so that while the tensor core works on one net's
// stage the same threads compute the other net's (the nets are independent graphs) — a tensor round trip is
// ~800 cycles, measured, and four of them per net were the critical path before the nets were interleaved:
//   P1  h1 = tanh([x,1] P1^T)               -> A1[net]      G1 : Z[net] = A1 W2^T                      (doneZ)
//   P2  h2 = tanh(Z + b2) -> A2[net], logits / value, loss, dOut -> DO[net]
//                                                            G5 : dW3^T += A2^T DO                      (doneW)
//   P3  dZ2 = (dOut W3)(1 - h2^2)           -> A2[net]      G2 : Z[net] = A2 W2                        (doneZ)
//                                                            G3 : dW2 += A2^T A1 ; d[.|b2] += A2^T XE   (doneW)
//   P4  dZ1 = Z (1 - h1^2)                  -> A2[net]      G4 : d[W1|b1] += A2^T XE                   (doneW)
// After a phase's shared-memory writes and a warpgroup-wide named barrier, thread 0 of the warpgroup issues the
// MMAs (one issuer warp for the whole CTA was measured to be a bottleneck).  Completions come back through two
// mbarriers per (warpgroup, net): "doneZ" (an accumulator to read) and "doneW" (operands free to overwrite).
// All weight-gradient accumulators stay in tensor memory for the whole kernel, private to a warpgroup (one issuing
// thread each => fixed accumulation order) and are drained once into the CTA's slab; reduce_kernel (update.cu) sums
// the slabs in fixed order, so results are run-to-run identical.  Gradients are kept un-normalised (no 1/n) until
// the drain so that fp16 operands do not underflow.
#include "ppo_loss.cuh"
#include "umma.cuh"
#include "update.cuh"

namespace ppo {
using namespace umma;

constexpr int TC_H = 64;
constexpr int TC_ROWS = 128;        // rows per warpgroup tile
constexpr int TC_MAX_K = 21;        // sum of logits (shared-memory budget of the per-row logit / db3 scratch)
constexpr int ACT_BYTES = TC_ROWS * 128;  // one activation tile

// Two-element members are stored as scalar pairs and read through selects: indexing a member ARRAY with a runtime `net`
// put the whole struct in local memory (194 LDL/STL in the kernel, a dependent ~30-cycle load in front of every phase).
#define TC_PAIR(name)                                                                     \
    int name##_0, name##_1;                                                               \
    __host__ __device__ __forceinline__ int name(int i) const { return i ? name##_1 : name##_0; }
struct TcLayout {   // byte offsets into the 1024-aligned dynamic shared memory
    TC_PAIR(w2)      // 16-bit W2 of actor (0) and critic (1), 64 rows x 128 B each
    int wg0, wg_stride;                  // per-warpgroup block
    TC_PAIR(a1) TC_PAIR(a2) TC_PAIR(xe) TC_PAIR(dout)   // offsets inside a warpgroup block ((net), XE: (tile parity))
    int lg, b3;
    TC_PAIR(p1h)     // 16-bit P1 = {W1|b1|0} as K chunks {hi, hi, lo, 0} [4][64 units][8]: B of the layer-1 MMAs
    TC_PAIR(p1) TC_PAIR(b2) TC_PAIR(w3) TC_PAIR(b3p)    // fp32 small parameters per net
    int bars, tmem_slot, end;
    int ldo, n3a;                        // logits row stride (floats), padded actor logit count
};
#undef TC_PAIR
__host__ __device__ inline TcLayout tc_layout(int K, int nwg) {
    TcLayout L;
    L.ldo = K | 1;
    L.n3a = (K + 7) & ~7;
    L.w2_0 = 0;
    L.w2_1 = 8192;
    L.a1_0 = 0; L.a1_1 = ACT_BYTES; L.a2_0 = 2 * ACT_BYTES; L.a2_1 = 3 * ACT_BYTES;
    L.xe_0 = 4 * ACT_BYTES; L.xe_1 = L.xe_0 + 2 * TC_ROWS * 16;   // per parity: 16-bit x (hi), then its residual (lo)
    L.dout_0 = L.xe_1 + 2 * TC_ROWS * 16;
    L.dout_1 = L.dout_0 + (L.n3a / 8) * TC_ROWS * 16;
    L.lg = L.dout_1 + TC_ROWS * 16;
    L.b3 = L.lg + ((TC_ROWS * L.ldo * 4 + 15) & ~15);
    L.wg_stride = (L.b3 + ((TC_ROWS * L.ldo * 4 + 15) & ~15) + 1023) & ~1023;
    L.wg0 = 16384;
    int o = L.wg0 + nwg * L.wg_stride;
    L.p1_0 = o; o += TC_H * 8 * 4;
    L.b2_0 = o; o += TC_H * 4;
    L.w3_0 = o; o += K * TC_H * 4;
    L.b3p_0 = o; o += ((K + 3) & ~3) * 4;
    L.p1_1 = o; o += TC_H * 8 * 4;
    L.b2_1 = o; o += TC_H * 4;
    L.w3_1 = o; o += TC_H * 4;
    L.b3p_1 = o; o += 4 * 4;
    o = (o + 127) & ~127;
    L.p1h_0 = o; o += 4 * TC_H * 16;
    L.p1h_1 = o; o += 4 * TC_H * 16;
    L.bars = o; o += 4 * 4 * 8;
    L.tmem_slot = o; o += 16;
    L.end = o;
    return L;
}

// optional in-kernel timeline (ppo_debug_tc_timeline): SM clock stamps of CTA 0, thread `who`, slot-indexed
__device__ long long* g_tc_ts = nullptr;
__device__ int g_tc_ts_who = 0;
// compiled in with -DPPO_TC_TIMELINE only (make TIMELINE=1): ~40 clock reads + predicated stores per tile were 4 % of the
// kernel's executed instructions (CS2R 144 per row in profiles/r02b_update_tc_E_ncu.txt)
#ifdef PPO_TC_TIMELINE
#define TC_TS(slot)                                                                       \
    do {                                                                                  \
        if (ts_buf && (slot) < 256) ts_buf[(slot)] = clock64();                           \
    } while (0)
#else
#define TC_TS(slot) do { } while (0)
#endif

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

// tensor-memory map: Z[g][net] at columns 64 (2 g + net) (M = 128: all lanes); then one block of gradient
// accumulators PER warpgroup (fixed accumulation order => run-to-run identical sums), M = 64 layout, the actor's in
// the lower half of every 32-lane subpartition and the critic's in the upper half (lane + 16):
//   {dW2 64, d[.|b2] 8, d[W1|b1] 8, dW3^T n3a}  =>  80 + n3a columns per warpgroup
struct TcCols {
    uint32_t w2, xe2, xe1, w3;   // tensor-memory address offsets (lane << 16 | column)
};
__device__ __forceinline__ TcCols tc_cols(int nwg, int g, int net, int n3a) {
    const uint32_t base = (uint32_t)nwg * 128u + (uint32_t)g * (80u + (uint32_t)n3a) + (net ? (16u << 16) : 0u);
    return TcCols{base, base + 64u, base + 72u, base + 80u};
}

// ---- MMA issue: warp 0 of a warpgroup issues that warpgroup's stage.  The whole warp walks the code with
// warp-uniform operands (every input is broadcast from lane 0 with a shuffle, which lets the compiler keep the
// descriptors in uniform registers instead of converting per-thread registers before every tcgen05.mma — that
// conversion chain made one MMA cost ~120 issue cycles); one elected lane executes the instructions.
struct WgIssue {
    uint32_t wb, sb, tmem;    // shared-memory address of the warpgroup block / of the CTA block, tensor-memory base
    int g;
    uint32_t fresh;           // bit (net * 4 + accumulator): the next MMA into it must overwrite
};
// ST 0..3 = G1, G5, G2+G3, G4 (see the header); ST 4 = G0 alone: Z[net] = XE[par] P1h^T (layer-1 pre-activations, K = 16:
// the 8 real inputs {obs, 1, 0..} + a zero chunk).  g0_next (ST 3 only): also issue G0 of the NEXT tile (XE[par ^ 1])
// in front of G4 - Z[net] has just been read by P4 and is free.
template <int FMT, int NWG, int ST>
__device__ __forceinline__ void issue_stage(WgIssue& w, const TcLayout& L, int net_, int par_, uint64_t* doneZ,
                                            uint64_t* doneW, bool g0_next = false) {
    constexpr uint32_t id_fwd = idesc_f16(FMT, 128, 64, 0, 0);   // Z = A1 W2^T
    constexpr uint32_t id_dx = idesc_f16(FMT, 128, 64, 0, 1);    // Z = A2 W2
    constexpr uint32_t id_dw = idesc_f16(FMT, 64, 64, 1, 1);     // dW2 += A2^T A1
    constexpr uint32_t id_n8 = idesc_f16(FMT, 64, 8, 1, 1);      // skinny, N = 8
    const uint32_t wb = __shfl_sync(0xffffffffu, w.wb, 0), sb = __shfl_sync(0xffffffffu, w.sb, 0);
    const uint32_t tmem = __shfl_sync(0xffffffffu, w.tmem, 0);
    const int g = __shfl_sync(0xffffffffu, w.g, 0), net = __shfl_sync(0xffffffffu, net_, 0);
    const int par = __shfl_sync(0xffffffffu, par_, 0);
    const uint32_t fresh = __shfl_sync(0xffffffffu, w.fresh, 0);
    const uint32_t fb = (uint32_t)net * 4u;
    const TcCols c = tc_cols(NWG, g, net, L.n3a);
    const uint32_t Z = tmem + (uint32_t)(g * 2 + net) * 64u;
    const uint32_t A1 = wb + L.a1(net), A2 = wb + L.a2(net), W2 = sb + L.w2(net);
    tc_fence_after();
    const bool lead = elect_one();
    if (ST == 0) {
        const uint64_t da = act_desc_kmajor(A1, 0), db = act_desc_kmajor(W2, 0);
        if (lead) {
#pragma unroll
            for (int ks = 0; ks < 4; ks++) mma_f16(Z, da + 2u * ks, db + 2u * ks, id_fwd, ks > 0);
            mma_commit(doneZ);
        }
    } else if (ST == 1) {
        const uint32_t acc0 = (fresh >> (fb + 3)) & 1u ? 0u : 1u;
        const uint64_t da = act_desc_mnmajor(A2, 0), db = narrow_desc_mnmajor(wb + L.dout(net), TC_ROWS, 0);
        const uint32_t id = net == 0 ? idesc_f16(FMT, 64, L.n3a, 1, 1) : id_n8;
        if (lead) {
#pragma unroll
            for (int ks = 0; ks < 8; ks++) mma_f16(tmem + c.w3, da + 128u * ks, db + 16u * ks, id, ks > 0 ? 1u : acc0);
            mma_commit(doneW);
        }
        w.fresh &= ~(1u << (fb + 3));
    } else if (ST == 2) {
        const uint64_t dak = act_desc_kmajor(A2, 0), dam = act_desc_mnmajor(A2, 0);
        const uint64_t dw = act_desc_mnmajor(W2, 0), dh = act_desc_mnmajor(A1, 0);
        const uint64_t dx = narrow_desc_mnmajor(wb + L.xe(par), TC_ROWS, 0);
        const uint32_t acc0 = (fresh >> fb) & 1u ? 0u : 1u;
        if (lead) {
#pragma unroll
            for (int ks = 0; ks < 4; ks++) mma_f16(Z, dak + 2u * ks, dw + 128u * ks, id_dx, ks > 0);
            mma_commit(doneZ);
#pragma unroll
            for (int ks = 0; ks < 8; ks++) {
                mma_f16(tmem + c.w2, dam + 128u * ks, dh + 128u * ks, id_dw, ks > 0 ? 1u : acc0);
                mma_f16(tmem + c.xe2, dam + 128u * ks, dx + 16u * ks, id_n8, ks > 0 ? 1u : acc0);
            }
            mma_commit(doneW);
        }
        w.fresh &= ~(3u << fb);
    } else if (ST == 4) {
        const uint64_t da = narrow_desc_kmajor(wb + L.xe(par), TC_ROWS * 16);          // {x.hi, x.lo}
        const uint64_t db = narrow_desc_kmajor(sb + L.p1h(net), TC_H * 16);             // {P.hi, P.hi}
        const uint64_t db2 = narrow_desc_kmajor(sb + L.p1h(net) + 2 * TC_H * 16, TC_H * 16);   // {P.lo, 0}
        if (lead) {
            mma_f16(Z, da, db, id_fwd, 0u);
            mma_f16(Z, da, db2, id_fwd, 1u);
            mma_commit(doneZ);
        }
    } else {
        const uint64_t dam = act_desc_mnmajor(A2, 0), dx = narrow_desc_mnmajor(wb + L.xe(par), TC_ROWS, 0);
        const uint32_t acc0 = (fresh >> (fb + 2)) & 1u ? 0u : 1u;
        const uint64_t da0 = narrow_desc_kmajor(wb + L.xe(par ^ 1), TC_ROWS * 16);
        const uint64_t db0 = narrow_desc_kmajor(sb + L.p1h(net), TC_H * 16);
        const uint64_t db1 = narrow_desc_kmajor(sb + L.p1h(net) + 2 * TC_H * 16, TC_H * 16);
        const bool g0 = __shfl_sync(0xffffffffu, (int)g0_next, 0) != 0;
        if (lead) {
            if (g0) {
                mma_f16(Z, da0, db0, id_fwd, 0u);
                mma_f16(Z, da0, db1, id_fwd, 1u);
                mma_commit(doneZ);
            }
#pragma unroll
            for (int ks = 0; ks < 8; ks++) mma_f16(tmem + c.xe1, dam + 128u * ks, dx + 16u * ks, id_n8, ks > 0 ? 1u : acc0);
            mma_commit(doneW);
        }
        w.fresh &= ~(1u << (fb + 2));
    }
    __syncwarp();
}

// ---- per-row pieces of a compute warpgroup --------------------------------------------------------------------
// layer 1 for 8 consecutive hidden units j0..j0+7: h = tanh(sum_i xe[i] P1[j][i]); P1 row = {W1[j][0..O), b1[j], 0..}
template <bool WIDE>
__device__ __forceinline__ void layer1_chunk(const float* P1, const float (&xe)[8], int j0, float* h) {
#pragma unroll
    for (int jj = 0; jj < 8; jj++) {
        const float4 w0 = *reinterpret_cast<const float4*>(P1 + (j0 + jj) * 8);
        float acc = xe[0] * w0.x;
        acc = fmaf(xe[1], w0.y, acc); acc = fmaf(xe[2], w0.z, acc); acc = fmaf(xe[3], w0.w, acc);
        if (WIDE) {
            const float4 w1 = *reinterpret_cast<const float4*>(P1 + (j0 + jj) * 8 + 4);
            acc = fmaf(xe[4], w1.x, acc); acc = fmaf(xe[5], w1.y, acc);
            acc = fmaf(xe[6], w1.z, acc); acc = fmaf(xe[7], w1.w, acc);
        }
        h[jj] = tanh_fast(acc);
    }
}

// ppo_debug_fuse_tail.  OFF by default: measured on B200 (config B, 148 CTAs) the in-kernel grid barriers cost as much as
// the launch they save - 3.30-3.32 ms per iteration with the tail in the kernel against 3.26 ms with reduce_apply_kernel as
// a second graph node - because the tail can only start after the slowest CTA either way.
int g_fuse_tail = 0;

// ---- fused tail: slab reduction + clip_grad_norm_ + AdamW behind a grid-wide barrier ------------------------------------
// Every CTA of the launch is resident (grid <= SM count, one CTA per SM; the launcher checks), so the CTAs can wait for
// each other on a global counter.  The spin is bounded: a CTA that never sees its peers falls through (wrong numbers, no
// hang).  Same arithmetic and summation order as reduce_apply_kernel (update.cu), which stays the path for small grids.
template <int NT>
__device__ __forceinline__ void grid_arrive_wait(unsigned* ctr, int tid) {
    __threadfence();   // this thread's global stores are visible device-wide before the CTA arrives
    bar_sync(3, NT);
    if (tid == 0) {
        atomicAdd(ctr, 1u);
        const long long t0 = clock64();
        while (*reinterpret_cast<volatile unsigned*>(ctr) < gridDim.x) {
            if (clock64() - t0 > 2000000000ll) break;
        }
        __threadfence();
    }
    bar_sync(3, NT);
}

template <int NT>
__device__ void fused_reduce_apply(const UpdArgs& a, uint8_t* scratch, int tid) {
    static_assert(NT == RED_COLS * RED_PARTS, "the tail is written for 256 threads");
    const TailArgs& t = a.tail;
    const FuseAdam& f = t.f;
    double(*part)[RED_COLS] = reinterpret_cast<double(*)[RED_COLS]>(scratch);
    double* shn = reinterpret_cast<double*>(scratch + RED_PARTS * RED_COLS * 8);
    const int P = t.P, nslabs = (int)gridDim.x, nblocks = (P + RED_COLS - 1) / RED_COLS;
    const int64_t stride = t.stride;
    grid_arrive_wait<NT>(t.ticket + 2, tid);   // every slab and every dpart row of this launch is written
    const int col = tid & (RED_COLS - 1), pt = tid / RED_COLS;
    const int blk = (int)blockIdx.x;           // gridDim.x >= nblocks: at most one column block per CTA
    const int i = blk * RED_COLS + col;
    double acc = 0.0;
    if (blk < nblocks && i < P) {
        const int per = (nslabs + RED_PARTS - 1) / RED_PARTS;
        const int s0 = pt * per, s1 = min(nslabs, s0 + per);
        const float* p = a.slabs + (size_t)s0 * stride + i;
        int sl = s0;
        for (; sl + 4 <= s1; sl += 4, p += 4 * stride) {
            const float a0 = __ldcg(p), a1 = __ldcg(p + stride), a2 = __ldcg(p + 2 * stride), a3 = __ldcg(p + 3 * stride);
            acc += (double)a0; acc += (double)a1; acc += (double)a2; acc += (double)a3;
        }
        for (; sl < s1; sl++, p += stride) acc += (double)__ldcg(p);
    }
    part[pt][col] = acc;
    bar_sync(3, NT);
    double sq = 0.0;
    if (pt == 0 && blk < nblocks && i < P) {
        double tt = part[0][col];
#pragma unroll
        for (int q = 1; q < RED_PARTS; q++) tt += part[q][col];
        const float g = (float)tt;
        t.grads_out[i] = g;
        sq = (double)g * (double)g;
    }
    if (tid < RED_COLS) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if ((tid & 31) == 0) shn[tid >> 5] = sq;
    }
    if (blk == 0 && tid >= 64 && tid < 96) {   // diagnostics (warp 2 of block 0)
        const int l = tid - 64;
        double v = 0.0;
        if (l < 5)
            for (int sl = 0; sl < nslabs; sl++) v += __ldcg(a.dpart + (size_t)sl * UPD_NDIAG + l);
        const double n_inv = (double)a.inv_n;
        const float pg = (float)(__shfl_sync(0xffffffffu, v, 0) * n_inv);
        const float vl = 0.5f * (float)(__shfl_sync(0xffffffffu, v, 1) * n_inv);
        const float en = (float)(__shfl_sync(0xffffffffu, v, 2) * n_inv);
        const float kl = (float)(__shfl_sync(0xffffffffu, v, 3) * n_inv);
        const float cf = (float)(__shfl_sync(0xffffffffu, v, 4) * n_inv);
        if (l < 8) {
            const float d8[8] = {pg, vl, en, pg - t.ent_coef * en + vl * t.vf_coef, kl, cf, 0.0f, 0.0f};
            float dv = 0.0f;
#pragma unroll
            for (int q = 0; q < 8; q++)
                if (q == l) dv = d8[q];
            if (t.diag_out) t.diag_out[l] = dv;
            if (f.diag_hist) f.diag_hist[l] = dv;
        }
    }
    bar_sync(3, NT);
    if (tid == 0 && blk < nblocks) t.sq_part[blk] = shn[0] + shn[1];
    grid_arrive_wait<NT>(t.ticket + 3, tid);   // every block's squared-norm partial is published
    double ss = 0.0;
    for (int b = tid; b < nblocks; b += NT) ss += __ldcg(t.sq_part + b);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((tid & 31) == 0) shn[tid >> 5] = ss;
    bar_sync(3, NT);
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < NT / 32; w++) tot += shn[w];
    const double norm = sqrt(tot);
    if (tid == 0 && blk == 0 && f.norm_out) *f.norm_out = (float)norm;
    const double coef = (double)f.max_grad_norm / (norm + 1e-6);
    const float cf = (f.max_grad_norm > 0.0f && coef < 1.0) ? (float)coef : 1.0f;
    const float lr = *f.d_lr;
    const int64_t step = f.step_host + (f.d_step ? *f.d_step : 0);
    const double b1 = (double)f.beta1, b2 = (double)f.beta2;
    const double bc1 = 1.0 - pow(b1, (double)step);
    const double bc2 = 1.0 - pow(b2, (double)step);
    const float decay = (float)(1.0 - (double)lr * (double)f.weight_decay);
    const float omb1 = (float)(1.0 - b1), omb2 = (float)(1.0 - b2);
    const float sq_bc2 = (float)sqrt(bc2);
    const float neg_step = (float)(-((double)lr / bc1));
    if (blk < nblocks && tid < RED_COLS && i < P) {   // this CTA steps its own 64 columns
        const int j = i;
        float gg = __ldcg(t.grads_out + j);
        if (cf != 1.0f) gg = __fmul_rn(gg, cf);
        const float pv = __fmul_rn(f.params[j], decay);
        const float mi = __fadd_rn(__fmul_rn(f.exp_avg[j], f.beta1), __fmul_rn(gg, omb1));
        const float vi = __fadd_rn(__fmul_rn(f.exp_avg_sq[j], f.beta2), __fmul_rn(__fmul_rn(gg, gg), omb2));
        f.exp_avg[j] = mi;
        f.exp_avg_sq[j] = vi;
        const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), sq_bc2), f.eps);
        f.params[j] = __fadd_rn(pv, __fmul_rn(neg_step, __fdiv_rn(mi, denom)));
    }
    // re-arm the counters for the next launch: the last CTA to leave has seen everybody pass both barriers
    bar_sync(3, NT);
    if (tid == 0 && atomicAdd(t.ticket + 4, 1u) == gridDim.x - 1) {
        t.ticket[2] = 0u; t.ticket[3] = 0u; t.ticket[4] = 0u;
    }
}

template <int FMT, int NWG, bool WIDE>
__global__ void __launch_bounds__((NWG + 1) * 128, 1) update_tc_kernel(const UpdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int O = a.d.O, K = a.d.K, heads = a.d.heads;
    const TcLayout L = tc_layout(K, NWG);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const NetOff offs[2] = {actor_off(O, TC_H, K), critic_off(O, TC_H)};
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);  // [(g * 2 + net) * 2 + {doneZ, doneW}]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const int ntiles = (a.mb_size + TC_ROWS - 1) / TC_ROWS;
    const int nthreads = (NWG + 1) * 128;   // NWG compute warpgroups + the MMA-issue warpgroup
    long long* ts_buf = (blockIdx.x == 0 && tid == (g_tc_ts_who & 0xFFFF)) ? g_tc_ts : nullptr;
    (void)ts_buf;
    // who | 0x40000000: thread 0 of EVERY CTA also stamps the global timer (ns) at entry / exit into slots 256 + 2 b + {0, 1}
    long long* cta_ts = (g_tc_ts && (g_tc_ts_who & 0x40000000) && tid == 0) ? g_tc_ts + 256 + 2 * blockIdx.x : nullptr;
    if (cta_ts) { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); cta_ts[0] = t; }
    TC_TS(0);

    // ---------------- prologue: tensor memory, barriers, parameters -> shared memory -------------------------
    pdl_trigger();   // the reduction / optimiser kernel behind this one may be scheduled; it waits for our completion
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    if (tid == 0) {
        for (int i = 0; i < NWG * 4; i++) mbar_init(bars + i, 1);   // [(g * 2 + net) * 2 + {doneZ, doneW}]
        for (int i = 0; i < NWG * 2; i++) mbar_init(bars + NWG * 4 + i, 128);   // ready[g * 2 + net]: operands staged
        mbar_init_fence();
    }
    pdl_wait();      // the previous optimiser step's parameters (and everything older) are visible from here on
    {   // W2 of both nets as 16-bit activation-layout tiles: 1024 chunk tasks, all loads of a thread in flight together
        constexpr int W2_IT = (2 * 64 * 8 + (NWG + 1) * 128 - 1) / ((NWG + 1) * 128);
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
            if (e < 2 * 64 * 8) store8<FMT>(sm + L.w2(e >> 9) + act_chunk_off((e >> 3) & 63, e & 7), v[it]);
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
                    dst[u] = reinterpret_cast<float*>(sm + L.p1(net)) + (e & 511);
                    if (i < O) src = a.params + offs[net].W1 + j * O + i;
                    else if (i == O) src = a.params + offs[net].b1 + j;
                } else if (e < S2) {
                    const int t = e - S1, net = t >> 6;
                    dst[u] = reinterpret_cast<float*>(sm + L.b2(net)) + (t & 63);
                    src = a.params + offs[net].b2 + (t & 63);
                } else if (e < S3) {
                    const int t = e - S2, net = t < 64 * K ? 0 : 1, idx = net ? t - 64 * K : t;
                    dst[u] = reinterpret_cast<float*>(sm + L.w3(net)) + idx;
                    src = a.params + offs[net].W3 + idx;
                } else {
                    const int t = e - S3, net = t < K ? 0 : 1, idx = net ? t - K : t;
                    dst[u] = reinterpret_cast<float*>(sm + L.b3p(net)) + idx;
                    src = a.params + offs[net].b3 + idx;
                }
                if (src) v[u] = __ldg(src);
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (dst[u]) *dst[u] = v[u];
        }
    }
    {   // 16-bit layer-1 weights P1[j] = {W1[j][0..O), b1[j], 0..}, K-major no-swizzle chunks {hi, hi, lo, 0}: with the
        // input split the same way (x = hi + lo) two K = 16 MMAs give x.hi P.hi + x.lo P.hi + x.hi P.lo, i.e. layer 1 to
        // ~2^-21 although every operand is 16 bits wide
        for (int e = tid; e < 2 * TC_H; e += nthreads) {
            const int net = e >> 6, j = e & 63;
            float v[8], lo[8], z[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                v[i] = 0.0f;
                z[i] = 0.0f;
                if (i < O) v[i] = __ldg(a.params + offs[net].W1 + j * O + i);
                if (i == O) v[i] = __ldg(a.params + offs[net].b1 + j);
            }
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                const uint32_t ph = pack2<FMT>(v[i], v[i + 1]);
                lo[i] = v[i] - unpack_lo<FMT>(ph);
                lo[i + 1] = v[i + 1] - unpack_hi<FMT>(ph);
            }
            uint8_t* dst = sm + L.p1h(net) + j * 16;
            store8<FMT>(dst, v);
            store8<FMT>(dst + TC_H * 16, v);
            store8<FMT>(dst + 2 * TC_H * 16, lo);
            store8<FMT>(dst + 3 * TC_H * 16, z);
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    TC_TS(1);

    // ---------------- MMA-issue warpgroup: warp i of it serves compute warpgroup i ---------------------------------
    // Every tcgen05.mma of these shapes occupies the issue slot for ~58 cycles (measured, tools/umma_time.py), 84 of them
    // per tile: issued from a compute warp that was ~4900 cycles per tile during which its warpgroup waited at the next
    // barrier.  Here the compute threads only arrive on ready[net] when a stage's operands are in shared memory; the
    // issuer replays the fixed stage order.  Registers move from this warpgroup to the compute ones (setmaxnreg).
    if (warp >= NWG * 4) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        const int g = warp - NWG * 4;
        if (g < NWG) {
            WgIssue D;
            D.wb = smem_u32(sm + L.wg0 + g * L.wg_stride); D.sb = smem_u32(sm); D.tmem = tmem; D.g = g; D.fresh = 0xFFu;
            uint64_t* wbars = bars + g * 4;
            uint64_t* ready = bars + NWG * 4 + g * 2;
            uint32_t pr = 0u;   // bit net = phase parity of ready[net] (a bit mask, not an array: runtime-indexed arrays live in local memory)
            const int nwg_total = gridDim.x * NWG;
            int tiles_done = 0;
            if (g * (int)gridDim.x + (int)blockIdx.x < ntiles) {
#pragma unroll 1
                for (int net = 0; net < 2; net++) {
                    mbar_wait(ready + net, (pr >> net) & 1u); pr ^= 1u << net;
                    issue_stage<FMT, NWG, 4>(D, L, net, 0, wbars + net * 2, wbars + net * 2 + 1);
                }
            }
            for (int tile = g * (int)gridDim.x + (int)blockIdx.x; tile < ntiles; tile += nwg_total) {
                const bool has_next = tile + nwg_total < ntiles;
                const int par = tiles_done & 1;
                tiles_done++;
#pragma unroll 1
                for (int net = 0; net < 2; net++) {
                    mbar_wait(ready + net, (pr >> net) & 1u); pr ^= 1u << net;
                    issue_stage<FMT, NWG, 0>(D, L, net, par, wbars + net * 2, wbars + net * 2 + 1);
                }
#pragma unroll 1
                for (int net = 0; net < 2; net++) {
                    mbar_wait(ready + net, (pr >> net) & 1u); pr ^= 1u << net;
                    issue_stage<FMT, NWG, 1>(D, L, net, par, wbars + net * 2, wbars + net * 2 + 1);
                }
#pragma unroll 1
                for (int net = 0; net < 2; net++) {
                    mbar_wait(ready + net, (pr >> net) & 1u); pr ^= 1u << net;
                    issue_stage<FMT, NWG, 2>(D, L, net, par, wbars + net * 2, wbars + net * 2 + 1);
                }
#pragma unroll 1
                for (int net = 0; net < 2; net++) {
                    mbar_wait(ready + net, (pr >> net) & 1u); pr ^= 1u << net;
                    issue_stage<FMT, NWG, 3>(D, L, net, par, wbars + net * 2, wbars + net * 2 + 1, has_next);
                }
            }
        }
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");

    RowSumsF rs;
    float gb3c = 0.0f;  // critic output-bias gradient partial of this thread
    {
        // ---------------- compute warpgroup ------------------------------------------------------------------
        const int g = warp >> 2, wt = tid & 127, w4 = warp & 3;
        uint64_t* ready = bars + NWG * 4 + g * 2;   // [net]: this stage's operands are staged (128 arrivals)
        uint8_t* wg = sm + L.wg0 + g * L.wg_stride;
        float* x = reinterpret_cast<float*>(wg + L.lg) + wt * L.ldo;     // this row's logits / dL/dlogits
        float* b3acc = reinterpret_cast<float*>(wg + L.b3) + wt * L.ldo; // this thread's actor db3 partials
        for (int q = 0; q < K; q++) b3acc[q] = 0.0f;
        uint64_t* wbars = bars + g * 4;          // [net * 2 + {doneZ, doneW}]
        uint32_t pz = 0u, pw = 0u;   // bit net = phase parity of doneZ / doneW of that net
        bool pendingW = false;                   // G4 of the previous tile not yet waited for (both nets)
        const bool masked = a.mask != nullptr;
        float meanf = 0.0f, stdf = 1.0f;
        if (a.hp.norm_adv) { meanf = a.adv_stats[0]; stdf = a.adv_stats[1]; }
        const int nwg_total = gridDim.x * NWG;

        // per-row inputs of a tile, gathered through the permutation one tile ahead (the loads of tile t+1 fly while
        // tile t computes)
        struct RowIn {
            int64_t r;
            float o[4];      // obs (O <= 4: prefetched; wider observations are loaded at use)
            float old_logp, adv, ret, old_val;
            int act0;        // action of head 0 (single-head fast path)
            bool on;
        };
        uint32_t pkey[4] = {0u, 0u, 0u, 0u};
        if (!a.perm && a.pf.inline_on) perm_keys(a.pf.seed, a.pf.epoch + (a.pf.d_epoch ? *a.pf.d_epoch : 0ull), pkey);
        auto gather = [&](int tile) {
            RowIn in;
            const int rloc = tile * TC_ROWS + wt;
            in.on = tile < ntiles && rloc < a.mb_size;
            in.r = 0;
            in.o[0] = in.o[1] = in.o[2] = in.o[3] = 0.0f;
            in.old_logp = in.adv = in.ret = in.old_val = 0.0f;
            in.act0 = 0;
            if (in.on) {
                const int64_t r = a.perm ? (int64_t)__ldg(a.perm + a.mb_start + rloc)
                                         : (a.pf.inline_on ? perm_index(a.mb_start + rloc, a.pf.batch, a.pf.half_bits, pkey)
                                                           : a.mb_start + rloc);
                in.r = r;
                if (a.rec) {   // packed row: {obs[4]} {old_logp, adv, ret, old_val} {action, -, -, -} {-}
                    const float4 q0 = __ldg(a.rec + r * 4), q1 = __ldg(a.rec + r * 4 + 1), q2 = __ldg(a.rec + r * 4 + 2);
                    in.o[0] = q0.x; in.o[1] = q0.y; in.o[2] = q0.z; in.o[3] = q0.w;
                    in.old_logp = q1.x; in.adv = q1.y; in.ret = q1.z; in.old_val = q1.w;
                    in.act0 = __float_as_int(q2.x);
                    return in;
                }
                if (O == 4) {
                    const float4 o4 = __ldg(reinterpret_cast<const float4*>(a.obs + r * 4));
                    in.o[0] = o4.x; in.o[1] = o4.y; in.o[2] = o4.z; in.o[3] = o4.w;
                } else if (O < 4) {
#pragma unroll
                    for (int i = 0; i < 3; i++)
                        if (i < O) in.o[i] = __ldg(a.obs + r * O + i);
                }
                in.old_logp = __ldg(a.old_logprob + r);
                in.act0 = __ldg(a.actions + r * heads);
                in.adv = __ldg(a.adv + r);
                in.ret = __ldg(a.ret + r);
                if (a.hp.clip_vloss) in.old_val = __ldg(a.old_val + r);
            }
            return in;
        };
        // XE row of a tile = {obs, 1, 0..} in 16 bits: A operand of the layer-1 MMA (G0) and B operand of d[W1|b1] (G3b/G4)
        auto stage_xe = [&](const RowIn& in, int par_) {
            float xe[8];
#pragma unroll
            for (int i = 0; i < 8; i++) xe[i] = 0.0f;
            if (O <= 4) {
#pragma unroll
                for (int i = 0; i < 4; i++) xe[i] = in.o[i];
            } else if (in.on) {
#pragma unroll
                for (int i = 0; i < 7; i++)
                    if (i < O) xe[i] = __ldg(a.obs + in.r * O + i);
            }
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (i == O) xe[i] = 1.0f;
            float lo[8];
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                const uint32_t ph = pack2<FMT>(xe[i], xe[i + 1]);
                lo[i] = xe[i] - unpack_lo<FMT>(ph);
                lo[i + 1] = xe[i + 1] - unpack_hi<FMT>(ph);
            }
            store8<FMT>(wg + L.xe(par_) + wt * 16, xe);
            store8<FMT>(wg + L.xe(par_) + TC_ROWS * 16 + wt * 16, lo);
        };
        RowIn nxt = gather(g * (int)gridDim.x + (int)blockIdx.x);
        int tiles_done = 0;
        if (g * (int)gridDim.x + (int)blockIdx.x < ntiles) {   // G0 of the first tile
            stage_xe(nxt, 0);
            fence_async_smem();
            mbar_arrive(ready + 0);
            mbar_arrive(ready + 1);
        }
        for (int tile = g * (int)gridDim.x + (int)blockIdx.x; tile < ntiles; tile += nwg_total) {
            const RowIn cur = nxt;
            nxt = gather(tile + nwg_total);
            const bool has_next = tile + nwg_total < ntiles;
            const bool on = cur.on;
            const int64_t r = cur.r;
            const int tsb = 8 + tiles_done * 40;
            (void)tsb;
            TC_TS(tsb);
            const int par = tiles_done & 1;      // XE is double-buffered: G3b / G4 of the previous tile may still read it
            tiles_done++;
            float gvc = 0.0f;                     // critic dL/dvalue of this row (lives from P2 to P3)

            // ---- P1: h1 = tanh(Z) (G0 = layer-1 pre-activations incl. bias, issued one tile ahead) -> A1[net]; G1
#pragma unroll 1
            for (int net = 0; net < 2; net++) {
                uint8_t* A1 = wg + L.a1(net);
                uint64_t* doneZ = wbars + net * 2;
                mbar_wait(doneZ, (pz >> net) & 1u); pz ^= 1u << net;
                tc_fence_after();
                const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)(g * 2 + net) * 64u);
#pragma unroll
                for (int hlf = 0; hlf < 2; hlf++) {
                    float h[32];
                    tmem_ld16(zaddr + hlf * 32, h);
                    tmem_ld16(zaddr + hlf * 32 + 16, h + 16);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; i++) h[i] = tanh_fast(h[i]);
#pragma unroll
                    for (int c = 0; c < 4; c++) store8<FMT>(A1 + act_chunk_off(wt, hlf * 4 + c), h + c * 8);
                }
                tc_fence_before();
                fence_async_smem();
                TC_TS(tsb + 1 + net * 3);
                mbar_arrive(ready + net);
                TC_TS(tsb + 2 + net * 3);
                TC_TS(tsb + 3 + net * 3);
            }
            // ---- P2: h2 = tanh(Z + b2) -> A2[net]; output layer; loss; dOut -> DO[net]; G5
#pragma unroll 1
            for (int net = 0; net < 2; net++) {
                const int k3 = net == 0 ? K : 1;
                const float* B2 = reinterpret_cast<const float*>(sm + L.b2(net));
                const float* W3 = reinterpret_cast<const float*>(sm + L.w3(net));
                const float* B3 = reinterpret_cast<const float*>(sm + L.b3p(net));
                uint8_t *A2 = wg + L.a2(net), *DO = wg + L.dout(net);
                uint64_t *doneZ = wbars + net * 2, *doneW = doneZ + 1;
                if (pendingW) {  // G4 of the previous tile has read A2[net]
                    mbar_wait(doneW, (pw >> net) & 1u); pw ^= 1u << net;
                }
                mbar_wait(doneZ, (pz >> net) & 1u); pz ^= 1u << net;
                tc_fence_after();
                TC_TS(tsb + 7 + net * 4);
                float h2[64];
                const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)(g * 2 + net) * 64u);
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
                    float a0 = B3[q], a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;   // four chains instead of one 64-deep one
#pragma unroll
                    for (int c = 0; c < 16; c++) {
                        const float4 w = *reinterpret_cast<const float4*>(W3 + q * TC_H + c * 4);
                        a0 = fmaf(h2[c * 4 + 0], w.x, a0); a1 = fmaf(h2[c * 4 + 1], w.y, a1);
                        a2 = fmaf(h2[c * 4 + 2], w.z, a2); a3 = fmaf(h2[c * 4 + 3], w.w, a3);
                    }
                    gv = (a0 + a1) + (a2 + a3);
                    if (net == 0) x[q] = gv;
                }
                if (net == 0) {
                    if (on) {
                        if (heads == 1 && !masked) {
                            actor_row_loss_fast1(x, K, cur.act0, cur.old_logp, cur.adv, a.hp, meanf, stdf, rs);
                        } else {
                            const uint64_t mbits = masked ? pack_mask(a.mask + r * K, K) : ~0ull;
                            actor_row_loss(x, a.d, masked, mbits, a.actions + r * heads, cur.old_logp, cur.adv, a.hp, meanf,
                                           stdf, 1.0f, rs);
                        }
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
                    gvc = on ? critic_row_loss(gv, cur.ret, cur.old_val, a.hp, 1.0f, rs) : 0.0f;
                    gb3c += gvc;
                    float v[8] = {gvc, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    store8<FMT>(DO + wt * 16, v);
                }
                tc_fence_before();
                fence_async_smem();
                TC_TS(tsb + 8 + net * 4);
                mbar_arrive(ready + net);
                TC_TS(tsb + 9 + net * 4);
                TC_TS(tsb + 10 + net * 4);
            }
            pendingW = false;
            // ---- P3: dZ2 = (dOut W3) * (1 - h2^2) -> A2[net] (over h2, once G5 has read it); G2, G3
#pragma unroll 1
            for (int net = 0; net < 2; net++) {
                const int k3 = net == 0 ? K : 1;
                const float* W3 = reinterpret_cast<const float*>(sm + L.w3(net));
                uint8_t* A2 = wg + L.a2(net);
                uint64_t *doneZ = wbars + net * 2, *doneW = doneZ + 1;
                float d[64];
#pragma unroll
                for (int k = 0; k < 64; k++) d[k] = 0.0f;
                for (int q = 0; q < k3; q++) {
                    const float dq = net == 0 ? x[q] : gvc;
#pragma unroll
                    for (int c = 0; c < 16; c++) {
                        const float4 w = *reinterpret_cast<const float4*>(W3 + q * TC_H + c * 4);
                        d[c * 4 + 0] = fmaf(dq, w.x, d[c * 4 + 0]); d[c * 4 + 1] = fmaf(dq, w.y, d[c * 4 + 1]);
                        d[c * 4 + 2] = fmaf(dq, w.z, d[c * 4 + 2]); d[c * 4 + 3] = fmaf(dq, w.w, d[c * 4 + 3]);
                    }
                }
                uint4 dz[8];
#pragma unroll
                for (int c = 0; c < 8; c++) {   // h2 comes back from its 16-bit copy (this thread wrote it in P2); the
                    // derivative factor and the product are formed on packed pairs (umma.cuh: dtanh2)
                    const uint4 hq = *reinterpret_cast<const uint4*>(A2 + act_chunk_off(wt, c));
                    dz[c].x = dtanh2<FMT>(d[c * 8 + 0], d[c * 8 + 1], hq.x); dz[c].y = dtanh2<FMT>(d[c * 8 + 2], d[c * 8 + 3], hq.y);
                    dz[c].z = dtanh2<FMT>(d[c * 8 + 4], d[c * 8 + 5], hq.z); dz[c].w = dtanh2<FMT>(d[c * 8 + 6], d[c * 8 + 7], hq.w);
                }
                TC_TS(tsb + 15 + net * 4);
                mbar_wait(doneW, (pw >> net) & 1u); pw ^= 1u << net;  // G5 has read A2 / DO
                TC_TS(tsb + 16 + net * 4);
#pragma unroll
                for (int c = 0; c < 8; c++) *reinterpret_cast<uint4*>(A2 + act_chunk_off(wt, c)) = dz[c];
                fence_async_smem();
                mbar_arrive(ready + net);
                TC_TS(tsb + 17 + net * 4);
                TC_TS(tsb + 18 + net * 4);
            }
            // ---- P4: dZ1 = Z * (1 - h1^2) -> A2[net] (once G2 / G3 have read dZ2); G4
#pragma unroll 1
            for (int net = 0; net < 2; net++) {
                uint8_t *A1 = wg + L.a1(net), *A2 = wg + L.a2(net);
                uint64_t *doneZ = wbars + net * 2, *doneW = doneZ + 1;
                if (net == 0 && has_next) stage_xe(nxt, par ^ 1);   // XE[par ^ 1]: last read by the previous tile's G4
                TC_TS(tsb + 23 + net * 6);
                mbar_wait(doneZ, (pz >> net) & 1u); pz ^= 1u << net;
                tc_fence_after();
                TC_TS(tsb + 24 + net * 6);
                float d[64];
                const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)(g * 2 + net) * 64u);
#pragma unroll
                for (int c = 0; c < 4; c++) tmem_ld16(zaddr + c * 16, d + c * 16);
                tmem_ld_wait();
                uint4 dz[8];
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    const uint4 hq = *reinterpret_cast<const uint4*>(A1 + act_chunk_off(wt, c));
                    dz[c].x = dtanh2<FMT>(d[c * 8 + 0], d[c * 8 + 1], hq.x); dz[c].y = dtanh2<FMT>(d[c * 8 + 2], d[c * 8 + 3], hq.y);
                    dz[c].z = dtanh2<FMT>(d[c * 8 + 4], d[c * 8 + 5], hq.z); dz[c].w = dtanh2<FMT>(d[c * 8 + 6], d[c * 8 + 7], hq.w);
                }
                TC_TS(tsb + 25 + net * 6);
                mbar_wait(doneW, (pw >> net) & 1u); pw ^= 1u << net;  // G3 has read A1 / A2
                TC_TS(tsb + 26 + net * 6);
#pragma unroll
                for (int c = 0; c < 8; c++) *reinterpret_cast<uint4*>(A2 + act_chunk_off(wt, c)) = dz[c];
                tc_fence_before();
                fence_async_smem();
                mbar_arrive(ready + net);
                TC_TS(tsb + 27 + net * 6);
                TC_TS(tsb + 28 + net * 6);
            }
            pendingW = true;
        }
        if (pendingW) {
            mbar_wait(wbars + 1, pw & 1u);
            mbar_wait(wbars + 3, (pw >> 1) & 1u);
        }
    }

    // ---------------- drain: tensor-memory accumulators -> this CTA's slab (x 1/n), diagnostics ----------------
    TC_TS(2);
    tc_fence_before();
    bar_sync(3, NWG * 128);
    TC_TS(3);
    tc_fence_after();
    const int P = num_params(O, TC_H, K);
    float* slab = a.slabs + (size_t)blockIdx.x * slab_stride(P);
    const float sc = a.inv_n;
    if (warp < 4) {
        // M = 64 accumulators: row m sits in lane (m % 16) + 32 (m / 16); the actor's use lanes 0..15 of each
        // subpartition and the critic's lanes 16..31, so one load serves both nets: lane l of warp w holds row
        // 16 w + (l & 15) of net (l >> 4).
        const int n = 16 * warp + (lane & 15);
        const int net = lane >> 4;
        const NetOff& of = offs[net];
        const int k3 = net == 0 ? K : 1;
        const uint32_t la = tmem_addr(tmem, (uint32_t)warp * 32u, 0);
        // warpgroups that owned no tile never wrote their accumulators: skip them
        int ng = 0;
        for (int g = 0; g < NWG; g++)
            if (g * (int)gridDim.x + (int)blockIdx.x < ntiles) ng = g + 1;
        float v[64], w[64];
#pragma unroll
        for (int g = 0; g < NWG; g++) {
            if (g >= ng) break;
            const uint32_t base = la + tc_cols(NWG, g, 0, L.n3a).w2;
#pragma unroll
            for (int cc = 0; cc < 4; cc++) tmem_ld16(base + cc * 16, (g == 0 ? v : w) + cc * 16);
            tmem_ld_wait();
            if (g > 0)
#pragma unroll
                for (int k = 0; k < 64; k++) v[k] += w[k];
        }
        {   // dW2 rows -> slab through a per-warp shared-memory transpose: a thread owns a whole 64-float row, so direct
            // stores would hit 32 sectors per instruction; staged, every store instruction covers one row segment
            float* buf = reinterpret_cast<float*>(sm + L.wg0 + (NWG > 1 ? L.wg_stride : 8192)) + warp * (32 * 65);
#pragma unroll
            for (int k = 0; k < 64; k++) buf[lane * 65 + k] = v[k] * sc;
            __syncwarp();
#pragma unroll 4
            for (int r = 0; r < 32; r++) {
                float* dst = slab + offs[r >> 4].W2 + (16 * warp + (r & 15)) * TC_H;
                dst[lane] = buf[r * 65 + lane];
                dst[lane + 32] = buf[r * 65 + lane + 32];
            }
        }
        float e[8], f[8];
#pragma unroll
        for (int g = 0; g < NWG; g++) {
            if (g >= ng) break;
            tmem_ld8(la + tc_cols(NWG, g, 0, L.n3a).xe2, g == 0 ? e : f);
            tmem_ld_wait();
            if (g > 0)
#pragma unroll
                for (int i = 0; i < 8; i++) e[i] += f[i];
        }
#pragma unroll
        for (int i = 0; i < 8; i++)
            if (i == O) slab[of.b2 + n] = e[i] * sc;
#pragma unroll
        for (int g = 0; g < NWG; g++) {
            if (g >= ng) break;
            tmem_ld8(la + tc_cols(NWG, g, 0, L.n3a).xe1, g == 0 ? e : f);
            tmem_ld_wait();
            if (g > 0)
#pragma unroll
                for (int i = 0; i < 8; i++) e[i] += f[i];
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (i < O) slab[of.W1 + n * O + i] = e[i] * sc;
            if (i == O) slab[of.b1 + n] = e[i] * sc;
        }
        for (int cc = 0; cc < L.n3a / 8; cc++) {   // critic lanes only own chunk 0 (the actor's other chunks map to junk there)
#pragma unroll
            for (int g = 0; g < NWG; g++) {
                if (g >= ng) break;
                tmem_ld8(la + tc_cols(NWG, g, 0, L.n3a).w3 + cc * 8, g == 0 ? e : f);
                tmem_ld_wait();
                if (g > 0)
#pragma unroll
                    for (int i = 0; i < 8; i++) e[i] += f[i];
            }
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (cc * 8 + i < k3) slab[of.W3 + (cc * 8 + i) * TC_H + n] = e[i] * sc;
        }
    }
    constexpr int NT = NWG * 128;
    double* red = reinterpret_cast<double*>(sm + L.wg0);                    // [5][NT]   (A1/A2 of warpgroup 0)
    float* gbc = reinterpret_cast<float*>(sm + L.wg0 + 5 * NT * 8);         // [NT]
    {
        red[0 * NT + tid] = (double)rs.pg; red[1 * NT + tid] = (double)rs.v; red[2 * NT + tid] = (double)rs.ent;
        red[3 * NT + tid] = (double)rs.kl; red[4 * NT + tid] = (double)rs.clip;
        gbc[tid] = gb3c;
    }
    bar_sync(3, NWG * 128);
    // one warp per quantity (5 diagnostics, the critic's output-bias gradient, K actor output-bias gradients), dealt round
    // robin over the NWG * 4 compute warps: lane-strided partial sums, then a fixed xor tree (deterministic)
    for (int item = warp; item < 6 + K; item += NWG * 4) {
        if (item < 5) {
            double acc = 0.0;
            for (int m = lane; m < NT; m += 32) acc += red[item * NT + m];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) a.dpart[(size_t)blockIdx.x * UPD_NDIAG + item] = acc;
        } else if (item == 5) {
            float acc = 0.0f;
            for (int m = lane; m < NT; m += 32) acc += gbc[m];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) slab[offs[1].b3] = acc * sc;
        } else {
            const int q = item - 6;
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
    bar_sync(3, NWG * 128);
    TC_TS(4);
    if (cta_ts) { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); cta_ts[1] = t; }
    if (warp == 0) tmem_dealloc(tmem, 512);
    if constexpr (NWG * 128 == RED_COLS * RED_PARTS) {
        if (a.tail.on) fused_reduce_apply<NWG * 128>(a, sm + L.wg0, tid);
    }
}

template <int FMT, int NWG, bool WIDE>
static int launch_tc(UpdArgs& a, int* grid_out, cudaStream_t st, const FuseAdam* fuse, float* grads_out, float* diag_out,
                     bool* tail_done) {
    const TcLayout L = tc_layout(a.d.K, NWG);
    const size_t smem = (size_t)L.end + 1024;
    PPO_REQUIRE(smem <= 227 * 1024, "update_tc: needs %zu B of shared memory", smem);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(update_tc_kernel<FMT, NWG, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      227 * 1024));
        attr_done = true;
    }
    const int ntiles = (a.mb_size + TC_ROWS - 1) / TC_ROWS;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    // tiles are dealt CTA-first (warpgroup g of CTA b starts at tile g * grid + b): a partial last round leaves ONE
    // busy warpgroup on many SMs instead of two on half as many
    int grid = ntiles;
    if (grid > sms) grid = sms;
    if (grid > UPD_MAX_GRID) grid = UPD_MAX_GRID;
    a.tail.on = 0;
    const int P = num_params(a.d.O, a.d.H, a.d.K);
    if (fuse && g_fuse_tail && NWG * 128 == RED_COLS * RED_PARTS && grid <= sms && grid >= (P + RED_COLS - 1) / RED_COLS) {
        char* tail = reinterpret_cast<char*>(a.dpart) + (size_t)UPD_MAX_GRID * UPD_NDIAG * sizeof(double);
        a.tail.on = 1; a.tail.P = P; a.tail.stride = slab_stride(P);
        a.tail.ent_coef = a.hp.ent_coef; a.tail.vf_coef = a.hp.vf_coef;
        a.tail.grads_out = grads_out; a.tail.diag_out = diag_out; a.tail.f = *fuse;
        a.tail.ticket = reinterpret_cast<unsigned*>(tail);
        a.tail.sq_part = reinterpret_cast<double*>(tail + 64);
    }
    if (tail_done) *tail_done = a.tail.on != 0;
    launch_pdl(a.pdl != 0, update_tc_kernel<FMT, NWG, WIDE>, dim3(grid), dim3((NWG + 1) * 128), smem, st, a);
    PPO_LAUNCH_CHECK();
    *grid_out = grid;
    return PPO_OK;
}

}  // namespace ppo
// d_ts: device buffer of 256 int64 (or NULL to switch off): later update_tc launches stamp the SM clock of CTA 0,
// thread `who` at fixed points (slot map in tools/tc_timeline.py)
extern "C" int ppo_debug_tc_timeline(long long* d_ts, int who) {
    PPO_CUDA(cudaMemcpyToSymbol(ppo::g_tc_ts, &d_ts, sizeof(d_ts)));
    PPO_CUDA(cudaMemcpyToSymbol(ppo::g_tc_ts_who, &who, sizeof(who)));
    return PPO_OK;
}
namespace ppo {
static bool tc_fits(int K, int nwg) { return (size_t)tc_layout(K, nwg).end + 1024 <= 227 * 1024; }
bool update_tc_supported(const Dims& d) { return d.H == TC_H && d.O + 1 <= 8 && d.K <= TC_MAX_K && tc_fits(d.K, 1); }
int launch_update_tc(UpdArgs& a, int fmt, int* grid_out, cudaStream_t st, const FuseAdam* fuse, float* grads_out,
                     float* diag_out, bool* tail_done) {
    if (tail_done) *tail_done = false;
    if (!update_tc_supported(a.d)) return PPO_ERR_UNSUPPORTED;
    const bool wide = a.d.O + 1 > 4;
    // two compute warpgroups per CTA when their per-row logit scratch fits next to the activation tiles (K <= ~16),
    // else one (the 21-logit MultiDiscrete shape nvec {3,3,3,3,3,2,2,2})
    if (!tc_fits(a.d.K, 2)) {
        if (fmt == FMT_BF16)
            return wide ? launch_tc<FMT_BF16, 1, true>(a, grid_out, st, fuse, grads_out, diag_out, tail_done)
                        : launch_tc<FMT_BF16, 1, false>(a, grid_out, st, fuse, grads_out, diag_out, tail_done);
        return wide ? launch_tc<FMT_F16, 1, true>(a, grid_out, st, fuse, grads_out, diag_out, tail_done)
                    : launch_tc<FMT_F16, 1, false>(a, grid_out, st, fuse, grads_out, diag_out, tail_done);
    }
    if (fmt == FMT_BF16)
        return wide ? launch_tc<FMT_BF16, 2, true>(a, grid_out, st, fuse, grads_out, diag_out, tail_done)
                    : launch_tc<FMT_BF16, 2, false>(a, grid_out, st, fuse, grads_out, diag_out, tail_done);
    return wide ? launch_tc<FMT_F16, 2, true>(a, grid_out, st, fuse, grads_out, diag_out, tail_done)
                : launch_tc<FMT_F16, 2, false>(a, grid_out, st, fuse, grads_out, diag_out, tail_done);
}
}  // namespace ppo

extern "C" int ppo_debug_fuse_tail(int on) {
    ppo::g_fuse_tail = on;
    return PPO_OK;
}
namespace ppo {

}  // namespace ppo
