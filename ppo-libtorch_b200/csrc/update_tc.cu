// update_tc.cu — K5 on the sm_100a tensor cores: fused minibatch forward + PPO loss + backward for the
// 2 x 64 actor / critic MLPs (PPO/Agent.cpp:25-59), replacing PPO/PPO_Discrete.cpp:579-638.
//
// The three 64 x 64 GEMMs of each net (forward y = x W2^T, backward dX = dY W2, backward dW2 = dY^T X) and the
// skinny weight-gradient GEMMs (dW1|db1, db2, dW3^T) run as tcgen05.mma (kind::f16: bf16 or fp16 operands,
// fp32 accumulation in tensor memory); everything per-sample (bias + tanh, the output layer, softmax /
// ratio / clipping / value loss, the tanh derivatives) runs one-thread-per-row on the CUDA cores.
//
// CTA = 2 NP compute warpgroups + one MMA-issue warpgroup.  Compute warpgroup c = 2 g + net walks the 128-row tiles of
// tile stream g (tile = g * grid + blockIdx.x, step grid * NP) for ONE net (0 = actor, 1 = critic): the two nets are
// independent graphs that only share the row inputs, so their warpgroups never wait for each other.  Rounds 1 / 2 had
// one warpgroup do both nets of a tile (interleaved, 232 registers): 8 compute warps per SM = 2 per scheduler, issue
// slots 43 % busy with every stall (tensor round trip ~800 cycles, MUFU, shared-memory round trips) exposed.  With a
// net per warpgroup the SAME shared-memory and tensor-memory budget feeds 16 compute warps (4 per scheduler, 112
// registers each): one warpgroup's tensor round trips are covered by the other three.  Per tile and net:
//   P1  h1 = tanh(Z)   (Z = layer-1 pre-activations, G0)      -> A1      G1 : Z = A1 W2^T                      (doneZ)
//   P2  h2 = tanh(Z + b2) -> A2, logits / value, loss, dOut -> DO         G5 : dW3^T += A2^T DO                 (doneW)
//   P3  dZ2 = (dOut W3)(1 - h2^2)           -> A2 (over h2)              G2 : Z = A2 W2                        (doneZ)
//                                                                         G3 : dW2 += A2^T A1 ; d[.|b2] += A2^T XE (doneW)
//   P4  dZ1 = Z (1 - h1^2)                  -> A2                         G0': Z = XE' P1h^T of the NEXT tile   (doneZ)
//                                                                         G4 : d[W1|b1] += A2^T XE              (doneW)
// Head layouts other than one unmasked head of 2 / 3 logits (KT = 0, up to 21 logits over any number of heads, optional
// mask) also run the actor's output layer on the tensor core: G7 : logits = A2 W3^T between the two halves of P2 (W3 as
// 16-bit hi + lo, so that only h2 carries a 16-bit rounding) and G6 : Z = DO W3 issued with G5, which P3 reads instead
// of forming dOut W3 row by row.
// Compute threads only stage operands and arrive on ready[c]; warp c of the issue warpgroup replays the fixed stage
// order for compute warpgroup c and issues the MMAs (every tcgen05.mma of these shapes holds the issuing warp ~58
// cycles).  Completions come back through two mbarriers per compute warpgroup: "doneZ" (an accumulator to read) and
// "doneW" (operands free to overwrite).  All weight-gradient accumulators stay in tensor memory for the whole kernel,
// private to a (stream, net) (one issuing warp each => fixed accumulation order) and are drained once into the CTA's
// slab; reduce_kernel (update.cu) sums the slabs in fixed order, so results are run-to-run identical.  Gradients are
// kept un-normalised (no 1/n) until the drain so that fp16 operands do not underflow.
#include "ppo_loss.cuh"
#include "umma.cuh"
#include "update.cuh"

namespace ppo {
using namespace umma;

constexpr int TC_H = 64;
constexpr int TC_ROWS = 128;        // rows per warpgroup tile
constexpr int TC_MAX_K = 21;        // sum of logits (shared-memory budget of the per-row logit / db3 scratch)
constexpr int ACT_BYTES = TC_ROWS * 128;  // one activation tile

// Two-element members are stored as scalar pairs and read through selects: indexing a member ARRAY with a runtime index
// put the whole struct in local memory (194 LDL/STL in the round-1 kernel).
#define TC_PAIR(name)                                                                     \
    int name##_0, name##_1;                                                               \
    __host__ __device__ constexpr int name(int i) const { return i ? name##_1 : name##_0; }
struct TcLayout {   // byte offsets into the 1024-aligned dynamic shared memory
    TC_PAIR(w2)      // 16-bit W2 of actor (0) and critic (1), 64 rows x 128 B each
    int wg0, wg_stride;                  // block of compute warpgroup c = 2 g + net
    int a1, a2, dout, lg, b3;            // offsets inside a block: activation tiles, dOut tile, logits / db3 scratch (actor)
    TC_PAIR(xe)      // (tile parity): 16-bit {obs, 1, 0..} (hi) then its residual (lo)
    TC_PAIR(p1h)     // 16-bit P1 = {W1|b1|0} as K chunks {hi, hi, lo, 0} [4][64 units][8]: B of the layer-1 MMAs
    TC_PAIR(b2) TC_PAIR(w3) TC_PAIR(b3p)    // fp32 small parameters per net
    TC_PAIR(w3h)     // W3 as packed 16-bit pairs [k3][32]: dh2 = dOut W3 of the single-head / critic path runs on pairs
    int bars, tmem_slot, end;
    int ldo, n3a;                        // logits row stride (floats), padded actor logit count
    // generic head path (KT = 0): the actor's output layer and dh2 = dOut W3 run on the tensor core
    int w3k;         // W3 as two 16-bit K-major swizzled tiles [n7 units][64] (hi, lo at + n7 * 128): B of the logits MMA (G7)
    int w3t;         // W3^T [64 hidden][16 kq units] 16-bit K-major narrow chunks: B of the dh2 MMA (G6)
    int n7, kq;      // N of G7 (K rounded up to 16), K-steps of G6 (n3a rounded up to 16, / 16)
    // single-head path (KT > 0): one 128B-swizzled [128 rows][64] tile whose logical 16-byte chunk c = 4 g + 2 net + parity
    // holds that chain's {obs, 1, 0..} (the hi halves of XE).  Addressed as the second 64-column group of the B operand of
    // dW2 += dZ2^T h1 (leading-dimension offset of the descriptor), it turns that MMA into dZ2^T [h1 | XE] with N = 72:
    // the separate d[.|b2] += dZ2^T XE series (8 MMAs that re-read the 16 KB dZ2 tile for 64 column sums) disappears.
    int xt;          // -1: not present (generic layouts have no room for it)
};
// tensor-memory columns of the logits accumulator of stream g (generic path), behind the gradient accumulators
__host__ __device__ constexpr int tc_lgcol(int np, int g, int n3a) { return np * 128 + np * (80 + n3a) + g * 32; }
#undef TC_PAIR
__host__ __device__ constexpr TcLayout tc_layout(int K, int np, bool gen = false) {
    TcLayout L{};
    L.ldo = K | 1;
    L.n3a = (K + 7) & ~7;
    L.w2_0 = 0;
    L.w2_1 = 8192;
    L.a1 = 0; L.a2 = ACT_BYTES;
    L.xe_0 = 2 * ACT_BYTES; L.xe_1 = L.xe_0 + 2 * TC_ROWS * 16;
    L.dout = L.xe_1 + 2 * TC_ROWS * 16;
    L.n7 = (K + 15) & ~15;
    L.kq = (L.n3a + 15) / 16;
    L.lg = L.dout + (gen ? 2 * L.kq : L.n3a / 8) * TC_ROWS * 16;   // generic: the dOut tile is padded to whole K = 16 steps
    L.b3 = L.lg + ((TC_ROWS * L.ldo * 4 + 15) & ~15);
    L.wg_stride = (L.b3 + ((TC_ROWS * L.ldo * 4 + 15) & ~15) + 1023) & ~1023;
    L.wg0 = 16384;
    int o = L.wg0 + 2 * np * L.wg_stride;
    L.b2_0 = o; o += TC_H * 4;
    L.w3_0 = o; o += K * TC_H * 4;
    L.b3p_0 = o; o += ((K + 3) & ~3) * 4;
    L.b2_1 = o; o += TC_H * 4;
    L.w3_1 = o; o += TC_H * 4;
    L.b3p_1 = o; o += 4 * 4;
    L.w3h_0 = o; o += K * TC_H * 2;
    L.w3h_1 = o; o += TC_H * 2;
    o = (o + 127) & ~127;
    L.p1h_0 = o; o += 4 * TC_H * 16;
    L.p1h_1 = o; o += 4 * TC_H * 16;
    L.w3k = 0; L.w3t = 0; L.xt = -1;
    if (!gen) {
        o = (o + 1023) & ~1023;
        L.xt = o; o += ACT_BYTES;
    }
    if (gen) {
        o = (o + 1023) & ~1023;
        L.w3k = o; o += 2 * L.n7 * 128;          // hi tile, lo tile at + n7 * 128 (2 or 4 KB: swizzle-atom aligned)
        L.w3t = o; o += 2 * L.kq * TC_H * 16;
    }
    L.bars = o; o += 4 * 3 * 8;          // per compute warpgroup c: doneZ, doneW at [2 c + {0, 1}], ready at [2 NC + c]
    L.tmem_slot = o; o += 16;
    L.end = o;
    return L;
}

// optional in-kernel timeline (ppo_debug_tc_timeline): SM clock stamps of CTA 0, thread `who`, slot-indexed
__device__ long long* g_tc_ts = nullptr;
__device__ int g_tc_ts_who = 0;
// compiled in with -DPPO_TC_TIMELINE only (make TIMELINE=1)
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

// tensor-memory map: Z[c] at columns 64 c (M = 128: all lanes), c = 2 g + net; then one block of gradient accumulators
// PER tile stream g (fixed accumulation order => run-to-run identical sums), M = 64 layout, the actor's in the lower half
// of every 32-lane subpartition and the critic's in the upper half (lane + 16):
//   {dW2 64, d[.|b2] 8, d[W1|b1] 8, dW3^T n3a}  =>  80 + n3a columns per stream
struct TcCols {
    uint32_t w2, xe2, xe1, w3;   // tensor-memory address offsets (lane << 16 | column)
};
__device__ __forceinline__ TcCols tc_cols(int np, int g, int net, int n3a) {
    const uint32_t base = (uint32_t)np * 128u + (uint32_t)g * (80u + (uint32_t)n3a) + (net ? (16u << 16) : 0u);
    return TcCols{base, base + 64u, base + 72u, base + 80u};
}

// ---- MMA issue: warp c of the issue warpgroup issues the stages of compute warpgroup c.  The whole warp walks the code
// with warp-uniform operands (every input is broadcast from lane 0 with a shuffle, which lets the compiler keep the
// descriptors in uniform registers instead of converting per-thread registers before every tcgen05.mma — that
// conversion chain made one MMA cost ~120 issue cycles); one elected lane executes the instructions.
struct WgIssue {
    uint32_t wb, sb, tmem;    // shared-memory address of the warpgroup block / of the CTA block, tensor-memory base
    int g, net;
    uint32_t fresh;           // bit (accumulator): the next MMA into it must overwrite
};
// ST 0..3 = G1, G5, G2+G3, G4 (see the header); ST 4 = G0 alone: Z = XE[par] P1h^T (layer-1 pre-activations, K = 16:
// the 8 real inputs {obs, 1, 0..} + a zero chunk).  g0_next (ST 3 only): also issue G0 of the NEXT tile (XE[par ^ 1])
// in front of G4 - Z has just been read by P4 and is free.
template <int FMT, int NP, int ST, bool GEN = false>
__device__ __forceinline__ void issue_stage(WgIssue& w, const TcLayout& L, int par_, uint64_t* doneZ, uint64_t* doneW,
                                            bool g0_next = false) {
    constexpr uint32_t id_fwd = idesc_f16(FMT, 128, 64, 0, 0);   // Z = A1 W2^T
    constexpr uint32_t id_dx = idesc_f16(FMT, 128, 64, 0, 1);    // Z = A2 W2
    constexpr uint32_t id_dw = idesc_f16(FMT, 64, 64, 1, 1);     // dW2 += A2^T A1
    constexpr uint32_t id_n8 = idesc_f16(FMT, 64, 8, 1, 1);      // skinny, N = 8
    const uint32_t wb = __shfl_sync(0xffffffffu, w.wb, 0), sb = __shfl_sync(0xffffffffu, w.sb, 0);
    const uint32_t tmem = __shfl_sync(0xffffffffu, w.tmem, 0);
    const int g = __shfl_sync(0xffffffffu, w.g, 0), net = __shfl_sync(0xffffffffu, w.net, 0);
    const int par = __shfl_sync(0xffffffffu, par_, 0);
    const uint32_t fresh = __shfl_sync(0xffffffffu, w.fresh, 0);
    const TcCols c = tc_cols(NP, g, net, L.n3a);
    const uint32_t Z = tmem + (uint32_t)(g * 2 + net) * 64u;
    const uint32_t A1 = wb + L.a1, A2 = wb + L.a2, W2 = sb + L.w2(net);
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
        const uint32_t acc0 = (fresh >> 3) & 1u ? 0u : 1u;
        const uint64_t da = act_desc_mnmajor(A2, 0), db = narrow_desc_mnmajor(wb + L.dout, TC_ROWS, 0);
        const uint32_t id = net == 0 ? idesc_f16(FMT, 64, L.n3a, 1, 1) : id_n8;
        if (lead) {
            if (GEN && net == 0) {   // G6: Z = dOut W3 (A = the dOut tile read K-major, B = W3^T), K = 16 kq
                for (int ks = 0; ks < L.kq; ks++)
                    mma_f16(Z, narrow_desc_kmajor(wb + L.dout + ks * 2 * (TC_ROWS * 16), TC_ROWS * 16),
                            narrow_desc_kmajor(sb + L.w3t + ks * 2 * (TC_H * 16), TC_H * 16), id_fwd, ks > 0);
                mma_commit(doneZ);
            }
#pragma unroll
            for (int ks = 0; ks < 8; ks++) mma_f16(tmem + c.w3, da + 128u * ks, db + 16u * ks, id, ks > 0 ? 1u : acc0);
            mma_commit(doneW);
        }
        w.fresh &= ~(1u << 3);
    } else if (ST == 2) {
        const uint64_t dak = act_desc_kmajor(A2, 0), dam = act_desc_mnmajor(A2, 0);
        const uint64_t dw = act_desc_mnmajor(W2, 0), dh = act_desc_mnmajor(A1, 0);
        const uint64_t dx = narrow_desc_mnmajor(wb + L.xe(par), TC_ROWS, 0);
        const uint32_t acc0 = fresh & 1u ? 0u : 1u;
        // B = [h1 | XE]: the second 64-column group of the MN-major operand lies LBO bytes behind the first (layout: xt)
        constexpr uint32_t id_dw72 = idesc_f16(FMT, 64, 72, 1, 1);
        const uint32_t xt_chunk = sb + (uint32_t)L.xt + 16u * (uint32_t)(4 * g + 2 * net + par);
        const uint64_t dh72 = smem_desc(A1, xt_chunk - A1, 1024, LAYOUT_SW128);
        if (lead) {
#pragma unroll
            for (int ks = 0; ks < 4; ks++) mma_f16(Z, dak + 2u * ks, dw + 128u * ks, id_dx, ks > 0);
            mma_commit(doneZ);
            if (!GEN) {
#pragma unroll
                for (int ks = 0; ks < 8; ks++) mma_f16(tmem + c.w2, dam + 128u * ks, dh72 + 128u * ks, id_dw72, ks > 0 ? 1u : acc0);
            } else {
#pragma unroll
                for (int ks = 0; ks < 8; ks++) {
                    mma_f16(tmem + c.w2, dam + 128u * ks, dh + 128u * ks, id_dw, ks > 0 ? 1u : acc0);
                    mma_f16(tmem + c.xe2, dam + 128u * ks, dx + 16u * ks, id_n8, ks > 0 ? 1u : acc0);
                }
            }
            mma_commit(doneW);
        }
        w.fresh &= ~3u;
    } else if (ST == 5) {   // G7 (generic actor path): logits = A2 W3^T, W3 as hi + lo so that only h2 carries a 16-bit rounding
        const uint32_t LG = tmem + (uint32_t)tc_lgcol(NP, g, L.n3a);
        const uint64_t da = act_desc_kmajor(A2, 0);
        const uint64_t dbh = act_desc_kmajor(sb + L.w3k, 0), dbl = act_desc_kmajor(sb + L.w3k + L.n7 * 128, 0);
        const uint32_t id7 = idesc_f16(FMT, 128, L.n7, 0, 0);
        if (lead) {
#pragma unroll
            for (int ks = 0; ks < 4; ks++) mma_f16(LG, da + 2u * ks, dbh + 2u * ks, id7, ks > 0);
#pragma unroll
            for (int ks = 0; ks < 4; ks++) mma_f16(LG, da + 2u * ks, dbl + 2u * ks, id7, 1u);
            mma_commit(doneZ);
        }
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
        const uint32_t acc0 = (fresh >> 2) & 1u ? 0u : 1u;
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
        w.fresh &= ~(1u << 2);
    }
    __syncwarp();
}

// ---- a compute warpgroup: thread wt owns row wt of every tile of its stream, for net NET ---------------------------------
// per-row inputs of a tile, gathered through the permutation during the previous tile's backward phases
struct TcRowIn {
    int64_t r;
    float o[4];      // obs (O <= 4: prefetched; wider observations are loaded at use)
    float f0, f1;    // actor: old_logp, adv;  critic: ret, old_val
    int act0;        // action of head 0 (single-head fast path)
    uint64_t mbits;  // generic actor path: the row's mask as bits and (heads <= 8) its actions as 8-bit fields, so that the
    uint64_t apack;  // loss does not wait for global loads in the middle of P2
    bool on;
};

// KT > 0: single unmasked head with exactly KT logits (PPO_Discrete on CartPole: 2, MountainCar: 3): K and with it the
// whole shared-memory layout are compile-time constants, and logits, dL/dlogits and the output-bias gradient partials
// stay in registers (ppo_loss.cuh: actor_row_loss_fast_reg).  KT = 0: any supported head layout, K at run time.
// Written for 112 registers per thread (NP = 2: 16 compute warps per SM): a spill costs a trip to L2 here, because
// 210 KB of the SM's 256 KB are shared memory and the ~20 KB of L1 left cannot hold 512 threads' stack frames (the
// first version of this kernel spilled ~60 words per tile and ran P2 at 7 800 cycles instead of 3 300).  Hence: every
// 64-wide row is processed in halves or quarters, and the next tile's rows are gathered after P2 instead of at the top.
template <int FMT, int NP, int NET, int KT>
__device__ __forceinline__ void tc_compute_wg(const UpdArgs& a, const TcLayout& Lrt, uint8_t* sm, uint32_t tmem, int g, int wt,
                                              int w4, int ntiles, RowSumsF& rs, float& gb3c, long long* ts_buf) {
    constexpr int NC = 2 * NP;
    constexpr bool FAST = KT > 0;
    constexpr int KR = FAST ? KT : 1;
    constexpr TcLayout LC = tc_layout(KR, NP);
    const TcLayout L = FAST ? LC : Lrt;
    const int O = a.d.O, K = FAST ? KT : a.d.K, heads = a.d.heads;
    const int c = 2 * g + NET;
    uint8_t* wg = sm + L.wg0 + c * L.wg_stride;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);
    uint64_t *doneZ = bars + 2 * c, *doneW = doneZ + 1, *ready = bars + 2 * NC + c;
    uint8_t *A1 = wg + L.a1, *A2 = wg + L.a2, *DO = wg + L.dout;
    float* x = reinterpret_cast<float*>(wg + L.lg) + wt * L.ldo;      // this row's logits / dL/dlogits (generic actor path)
    float* b3acc = reinterpret_cast<float*>(wg + L.b3) + wt * L.ldo;  // this thread's actor db3 partials
    const float* B2 = reinterpret_cast<const float*>(sm + L.b2(NET));
    const float* W3 = reinterpret_cast<const float*>(sm + L.w3(NET));
    const float* B3 = reinterpret_cast<const float*>(sm + L.b3p(NET));
    if (NET == 0 && !FAST)
        for (int q = 0; q < K; q++) b3acc[q] = 0.0f;
    float b3r[KR];
#pragma unroll
    for (int q = 0; q < KR; q++) b3r[q] = 0.0f;
    uint32_t pz = 0u, pw = 0u;       // phase parities of doneZ / doneW
    bool pendingW = false;           // G4 of the previous tile not yet waited for
    const bool masked = a.mask != nullptr;
    float meanf = 0.0f, stdf = 1.0f;
    if (NET == 0 && a.hp.norm_adv) { meanf = a.adv_stats[0]; stdf = a.adv_stats[1]; }
    const float rstd = __fdividef(1.0f, stdf + 1e-8f);
    const int stride = (int)gridDim.x * NP;
    const int tile0 = g * (int)gridDim.x + (int)blockIdx.x;
    const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)c * 64u);

    uint32_t pkey[4] = {0u, 0u, 0u, 0u};
    if (!a.perm && a.pf.inline_on) perm_keys(a.pf.seed, a.pf.epoch + (a.pf.d_epoch ? *a.pf.d_epoch : 0ull), pkey);
    auto gather = [&](int tile) {
        TcRowIn in;
        const int rloc = tile * TC_ROWS + wt;
        in.on = tile < ntiles && rloc < a.mb_size;
        in.r = 0;
        in.o[0] = in.o[1] = in.o[2] = in.o[3] = 0.0f;
        in.f0 = in.f1 = 0.0f;
        in.act0 = 0;
        in.mbits = ~0ull;
        in.apack = 0ull;
        if (in.on) {
            const int64_t r = a.perm ? (int64_t)__ldg(a.perm + a.mb_start + rloc)
                                     : (a.pf.inline_on ? perm_index(a.mb_start + rloc, a.pf.batch, a.pf.half_bits, pkey)
                                                       : a.mb_start + rloc);
            in.r = r;
            if (NET == 0 && !FAST) {
                if (masked) in.mbits = pack_mask(a.mask + r * K, K);
                if (heads <= 8)
                    for (int h = 0; h < heads; h++) in.apack |= (uint64_t)(__ldg(a.actions + r * heads + h) & 0xFF) << (8 * h);
            }
            if (a.rec) {   // packed row: {obs[4]} {old_logp, adv, ret, old_val} {action, -, -, -} {-}
                const float4 q0 = __ldg(a.rec + r * 4);
                in.o[0] = q0.x; in.o[1] = q0.y; in.o[2] = q0.z; in.o[3] = q0.w;
                const float2* q1 = reinterpret_cast<const float2*>(a.rec + r * 4 + 1);
                if (NET == 0) {
                    const float2 v = __ldg(q1);
                    in.f0 = v.x; in.f1 = v.y;
                    in.act0 = __float_as_int(__ldg(reinterpret_cast<const float*>(a.rec + r * 4 + 2)));
                } else {
                    const float2 v = __ldg(q1 + 1);
                    in.f0 = v.x; in.f1 = v.y;
                }
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
            if (NET == 0) {
                in.f0 = __ldg(a.old_logprob + r);
                in.f1 = __ldg(a.adv + r);
                in.act0 = __ldg(a.actions + r * heads);
            } else {
                in.f0 = __ldg(a.ret + r);
                if (a.hp.clip_vloss) in.f1 = __ldg(a.old_val + r);
            }
        }
        return in;
    };
    // XE row of a tile = {obs, 1, 0..} in 16 bits: A operand of the layer-1 MMA (G0) and B operand of d[W1|b1] (G3b / G4)
    auto stage_xe = [&](const TcRowIn& in, int par_) {
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
        if (FAST) store8<FMT>(sm + L.xt + act_chunk_off(wt, 4 * g + 2 * NET + par_), xe);   // the copy dW2's MMA reads (layout: xt)
    };
    TcRowIn cur = gather(tile0);
    int tiles_done = 0;
    if (tile0 < ntiles) {   // G0 of the first tile
        stage_xe(cur, 0);
        fence_async_smem();
        mbar_arrive(ready);
    }
    for (int tile = tile0; tile < ntiles; tile += stride) {
        const bool has_next = tile + stride < ntiles;
        const bool on = cur.on;
        const int tsb = 8 + tiles_done * 16;
        (void)tsb;
        TC_TS(tsb);
        const int par = tiles_done & 1;      // XE is double-buffered: G3b / G4 of the previous tile may still read it
        tiles_done++;

        // ---- P1: h1 = tanh(Z) (G0 = layer-1 pre-activations incl. bias, issued one tile ahead) -> A1; G1
        mbar_wait(doneZ, pz); pz ^= 1u;
        tc_fence_after();
        TC_TS(tsb + 1);
#pragma unroll
        for (int hlf = 0; hlf < 2; hlf++) {
            float h[32];
            tmem_ld16(zaddr + hlf * 32, h);
            tmem_ld16(zaddr + hlf * 32 + 16, h + 16);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i++) h[i] = tanh_fast(h[i]);
#pragma unroll
            for (int cc = 0; cc < 4; cc++) store8<FMT>(A1 + act_chunk_off(wt, hlf * 4 + cc), h + cc * 8);
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(ready);
        TC_TS(tsb + 2);

        // ---- P2: h2 = tanh(Z + b2) -> A2; output layer; loss; dOut -> DO; G5
        if (pendingW) {  // G4 of the previous tile has read A2
            mbar_wait(doneW, pw); pw ^= 1u;
        }
        mbar_wait(doneZ, pz); pz ^= 1u;
        tc_fence_after();
        TC_TS(tsb + 3);
        float xr[KR];        // FAST actor: logits, then dL/dlogits (lives into P3)
        float gvc = 0.0f;    // critic: value, then dL/dvalue of this row (lives into P3)
#pragma unroll
        for (int q = 0; q < KR; q++) xr[q] = 0.0f;
#pragma unroll
        for (int hlf = 0; hlf < 2; hlf++) {   // 32 hidden units at a time (register budget)
            float h2[32];
            tmem_ld16(zaddr + hlf * 32, h2);
            tmem_ld16(zaddr + hlf * 32 + 16, h2 + 16);
            tmem_ld_wait();
#pragma unroll
            for (int cc = 0; cc < 8; cc++) {
                const float4 b = *reinterpret_cast<const float4*>(B2 + hlf * 32 + cc * 4);
                h2[cc * 4 + 0] = tanh_fast(h2[cc * 4 + 0] + b.x); h2[cc * 4 + 1] = tanh_fast(h2[cc * 4 + 1] + b.y);
                h2[cc * 4 + 2] = tanh_fast(h2[cc * 4 + 2] + b.z); h2[cc * 4 + 3] = tanh_fast(h2[cc * 4 + 3] + b.w);
            }
#pragma unroll
            for (int cc = 0; cc < 4; cc++) store8<FMT>(A2 + act_chunk_off(wt, hlf * 4 + cc), h2 + cc * 8);
            // this half's share of one output unit: two chains
            auto out_half = [&](int q) {
                float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
                for (int cc = 0; cc < 8; cc++) {
                    const float4 w = *reinterpret_cast<const float4*>(W3 + q * TC_H + hlf * 32 + cc * 4);
                    a0 = fmaf(h2[cc * 4 + 0], w.x, a0); a1 = fmaf(h2[cc * 4 + 1], w.y, a1);
                    a0 = fmaf(h2[cc * 4 + 2], w.z, a0); a1 = fmaf(h2[cc * 4 + 3], w.w, a1);
                }
                return a0 + a1;
            };
            if (NET == 1) {
                gvc += out_half(0);
            } else if (FAST) {
#pragma unroll
                for (int q = 0; q < KR; q++) xr[q] += out_half(q);
            }
            // generic actor path: no per-row output layer here - the logits come from the tensor core (G7, below)
        }
        if (NET == 0 && !FAST) {
            tc_fence_before();
            fence_async_smem();
            mbar_arrive(ready);          // h2 is staged: G7 = logits
            mbar_wait(doneZ, pz); pz ^= 1u;
            tc_fence_after();
            const uint32_t lgaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)tc_lgcol(NP, g, L.n3a));
            for (int c16 = 0; c16 < L.n7 / 16; c16++) {
                float lg[16];
                tmem_ld16(lgaddr + c16 * 16, lg);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; i++)
                    if (c16 * 16 + i < K) x[c16 * 16 + i] = lg[i] + B3[c16 * 16 + i];
            }
        }
        if (NET == 1) {
            gvc = on ? critic_row_loss(gvc + B3[0], cur.f0, cur.f1, a.hp, 1.0f, rs) : 0.0f;
            gb3c += gvc;
            float v[8] = {gvc, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            store8<FMT>(DO + wt * 16, v);
        } else if (FAST) {
            if (on) {
#pragma unroll
                for (int q = 0; q < KR; q++) xr[q] += B3[q];
                actor_row_loss_fast_reg<KR>(xr, KR, cur.act0, cur.f0, cur.f1, a.hp, meanf, rstd, rs);
#pragma unroll
                for (int q = 0; q < KR; q++) b3r[q] += xr[q];
            } else {
#pragma unroll
                for (int q = 0; q < KR; q++) xr[q] = 0.0f;
            }
            float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int q = 0; q < KR; q++) v[q] = xr[q];
            store8<FMT>(DO + wt * 16, v);
        } else {
            if (on) {
                actor_row_loss_fast_multi(x, a.d, masked, cur.mbits, a.actions + cur.r * heads, cur.apack, cur.f0, cur.f1, a.hp, meanf,
                                          rstd, rs);
                for (int q = 0; q < K; q++) b3acc[q] += x[q];
            } else {
                for (int q = 0; q < K; q++) x[q] = 0.0f;
            }
            for (int cc = 0; cc < L.n3a / 8; cc++) {
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = (cc * 8 + i < K) ? x[cc * 8 + i] : 0.0f;
                store8<FMT>(DO + cc * (TC_ROWS * 16) + wt * 16, v);
            }
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(ready);
        TC_TS(tsb + 4);
        pendingW = false;

        // the next tile's rows: their loads fly during P3 / P4 and are first used at the end of P4
        const TcRowIn nxt = gather(tile + stride);

        // ---- P3: dZ2 = (dOut W3) * (1 - h2^2) -> A2 (over h2, once G5 has read it); G2, G3
        {
            uint4 dz[8];
            if (NET == 1 || FAST) {
                // dOut has at most three entries per row: dh2 is formed on packed 16-bit pairs (dOut carries the same
                // 16-bit rounding as the DO tile the tensor core reads for dW3), straight into the derivative
                const uint4* W3H = reinterpret_cast<const uint4*>(sm + L.w3h(NET));
                uint32_t g2[KR];
                if (NET == 1) g2[0] = pack2<FMT>(gvc, gvc);
                else
#pragma unroll
                    for (int q = 0; q < KR; q++) g2[q] = pack2<FMT>(xr[q], xr[q]);
#pragma unroll
                for (int cc = 0; cc < 8; cc++) {
                    uint4 w = W3H[cc];
                    uint4 d = {hmul2<FMT>(g2[0], w.x), hmul2<FMT>(g2[0], w.y), hmul2<FMT>(g2[0], w.z), hmul2<FMT>(g2[0], w.w)};
                    if (NET == 0) {
#pragma unroll
                        for (int q = 1; q < KR; q++) {
                            w = W3H[q * 8 + cc];
                            d.x = hfma2<FMT>(g2[q], w.x, d.x); d.y = hfma2<FMT>(g2[q], w.y, d.y);
                            d.z = hfma2<FMT>(g2[q], w.z, d.z); d.w = hfma2<FMT>(g2[q], w.w, d.w);
                        }
                    }
                    // h2 comes back from its 16-bit copy (this thread wrote it in P2)
                    const uint4 hq = *reinterpret_cast<const uint4*>(A2 + act_chunk_off(wt, cc));
                    dz[cc].x = dtanh2p<FMT>(d.x, hq.x); dz[cc].y = dtanh2p<FMT>(d.y, hq.y);
                    dz[cc].z = dtanh2p<FMT>(d.z, hq.z); dz[cc].w = dtanh2p<FMT>(d.w, hq.w);
                }
            } else {
                // generic actor path: dh2 = dOut W3 comes from the tensor core (G6, issued with G5 behind P2)
                mbar_wait(doneZ, pz); pz ^= 1u;
                tc_fence_after();
#pragma unroll
                for (int qt = 0; qt < 4; qt++) {
                    float d[16];
                    tmem_ld16(zaddr + qt * 16, d);
                    tmem_ld_wait();
#pragma unroll
                    for (int cc = 0; cc < 2; cc++) {
                        const uint4 hq = *reinterpret_cast<const uint4*>(A2 + act_chunk_off(wt, qt * 2 + cc));
                        uint4& o = dz[qt * 2 + cc];
                        o.x = dtanh2<FMT>(d[cc * 8 + 0], d[cc * 8 + 1], hq.x); o.y = dtanh2<FMT>(d[cc * 8 + 2], d[cc * 8 + 3], hq.y);
                        o.z = dtanh2<FMT>(d[cc * 8 + 4], d[cc * 8 + 5], hq.z); o.w = dtanh2<FMT>(d[cc * 8 + 6], d[cc * 8 + 7], hq.w);
                    }
                }
                tc_fence_before();
            }
            TC_TS(tsb + 5);
            mbar_wait(doneW, pw); pw ^= 1u;  // G5 has read A2 / DO
            TC_TS(tsb + 6);
#pragma unroll
            for (int cc = 0; cc < 8; cc++) *reinterpret_cast<uint4*>(A2 + act_chunk_off(wt, cc)) = dz[cc];
            fence_async_smem();
            mbar_arrive(ready);
        }
        TC_TS(tsb + 7);
        // ---- P4: dZ1 = Z * (1 - h1^2) -> A2 (once G2 / G3 have read dZ2); G0 of the next tile, G4
        {
            mbar_wait(doneZ, pz); pz ^= 1u;
            tc_fence_after();
            TC_TS(tsb + 8);
            uint4 dz[8];
#pragma unroll
            for (int qt = 0; qt < 4; qt++) {   // 16 hidden units at a time (register budget)
                float d[16];
                tmem_ld16(zaddr + qt * 16, d);
                tmem_ld_wait();
#pragma unroll
                for (int cc = 0; cc < 2; cc++) {
                    const uint4 hq = *reinterpret_cast<const uint4*>(A1 + act_chunk_off(wt, qt * 2 + cc));
                    uint4& o = dz[qt * 2 + cc];
                    o.x = dtanh2<FMT>(d[cc * 8 + 0], d[cc * 8 + 1], hq.x); o.y = dtanh2<FMT>(d[cc * 8 + 2], d[cc * 8 + 3], hq.y);
                    o.z = dtanh2<FMT>(d[cc * 8 + 4], d[cc * 8 + 5], hq.z); o.w = dtanh2<FMT>(d[cc * 8 + 6], d[cc * 8 + 7], hq.w);
                }
            }
            TC_TS(tsb + 9);
            mbar_wait(doneW, pw); pw ^= 1u;  // G3 has read A1 / A2
            TC_TS(tsb + 10);
#pragma unroll
            for (int cc = 0; cc < 8; cc++) *reinterpret_cast<uint4*>(A2 + act_chunk_off(wt, cc)) = dz[cc];
            if (has_next) stage_xe(nxt, par ^ 1);   // XE[par ^ 1]: last read by the previous tile's G4 (waited for in P2)
            tc_fence_before();
            fence_async_smem();
            mbar_arrive(ready);
            TC_TS(tsb + 11);
        }
        pendingW = true;
        cur = nxt;
    }
    if (pendingW) mbar_wait(doneW, pw);
    if (NET == 0 && FAST) {   // the drain sums the db3 partials from the per-row scratch
#pragma unroll
        for (int q = 0; q < KR; q++) b3acc[q] = b3r[q];
    }
}


// ppo_debug_fuse_tail: kept for ABI stability, no effect.  The round-1 experiment (slab reduction + clip + AdamW inside this
// kernel behind a grid barrier) measured 1-2 % slower than reduce_apply_kernel as a second graph node and was removed
// when the kernel went to 512 compute threads.
int g_fuse_tail = 0;

template <int FMT, int NP, int KT>
__global__ void __launch_bounds__((2 * NP + 1) * 128, 1) update_tc_kernel(const UpdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int NC = 2 * NP;                  // compute warpgroups
    constexpr int nthreads = (NC + 1) * 128;    // + the MMA-issue warpgroup
    constexpr TcLayout LC = tc_layout(KT > 0 ? KT : 1, NP);
    const int O = a.d.O, K = KT > 0 ? KT : a.d.K;
    const TcLayout L = KT > 0 ? LC : tc_layout(K, NP, true);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const NetOff offs[2] = {actor_off(O, TC_H, K), critic_off(O, TC_H)};
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);  // [2 c + {doneZ, doneW}], ready at [2 NC + c]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const int ntiles = (a.mb_size + TC_ROWS - 1) / TC_ROWS;
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
        for (int i = 0; i < NC * 2; i++) mbar_init(bars + i, 1);
        for (int i = 0; i < NC; i++) mbar_init(bars + NC * 2 + i, 128);   // ready[c]: operands staged
        mbar_init_fence();
    }
    pdl_wait();      // the previous optimiser step's parameters (and everything older) are visible from here on
    {   // W2 of both nets as 16-bit activation-layout tiles: 1024 chunk tasks, all loads of a thread in flight together
        constexpr int W2_IT = (2 * 64 * 8 + nthreads - 1) / nthreads;
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
    {   // fp32 small parameters: b2 [64], W3 [k3][64], b3 [k3] per net, one flat task list
        const int S2 = 128, S3 = S2 + 64 * (K + 1), S4 = S3 + (K + 1);
        for (int e = tid; e < S4; e += nthreads) {
            float* dst;
            const float* src;
            if (e < S2) {
                const int net = e >> 6;
                dst = reinterpret_cast<float*>(sm + L.b2(net)) + (e & 63);
                src = a.params + offs[net].b2 + (e & 63);
            } else if (e < S3) {
                const int t = e - S2, net = t < 64 * K ? 0 : 1, idx = net ? t - 64 * K : t;
                dst = reinterpret_cast<float*>(sm + L.w3(net)) + idx;
                src = a.params + offs[net].W3 + idx;
            } else {
                const int t = e - S3, net = t < K ? 0 : 1, idx = net ? t - K : t;
                dst = reinterpret_cast<float*>(sm + L.b3p(net)) + idx;
                src = a.params + offs[net].b3 + idx;
            }
            *dst = __ldg(src);
        }
    }
    if (KT == 0) {   // generic head path: operands of the logits MMA (G7) and of dh2 = dOut W3 (G6), zero padding of the dOut tiles
        for (int e = tid; e < L.n7 * 8; e += nthreads) {
            const int q = e >> 3, cc = e & 7;
            float v[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = q < K ? __ldg(a.params + offs[0].W3 + q * TC_H + cc * 8 + i) : 0.0f;
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                const uint32_t ph = pack2<FMT>(v[i], v[i + 1]);
                lo[i] = v[i] - unpack_lo<FMT>(ph);
                lo[i + 1] = v[i + 1] - unpack_hi<FMT>(ph);
            }
            store8<FMT>(sm + L.w3k + act_chunk_off(q, cc), v);
            store8<FMT>(sm + L.w3k + L.n7 * 128 + act_chunk_off(q, cc), lo);
        }
        for (int e = tid; e < 2 * L.kq * TC_H; e += nthreads) {
            const int cq = e >> 6, n = e & 63;
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = cq * 8 + i < K ? __ldg(a.params + offs[0].W3 + (cq * 8 + i) * TC_H + n) : 0.0f;
            store8<FMT>(sm + L.w3t + cq * (TC_H * 16) + n * 16, v);
        }
        const int pad0 = L.n3a / 8, npad = 2 * L.kq - pad0;   // dOut chunks beyond the real logits: read by G6, never written
        for (int e = tid; e < NP * npad * TC_ROWS; e += nthreads) {
            const int g = e / (npad * TC_ROWS), r = e % (npad * TC_ROWS);
            float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            store8<FMT>(sm + L.wg0 + (2 * g) * L.wg_stride + L.dout + pad0 * (TC_ROWS * 16) + r * 16, z);
        }
    }
    for (int e = tid; e < (K + 1) * 32; e += nthreads) {   // W3 as packed pairs, actor rows 0 .. K-1 then the critic's row
        const int q = e >> 5, j2 = e & 31, net = q < K ? 0 : 1;
        const float* src = a.params + offs[net].W3 + (net ? 0 : q * TC_H) + 2 * j2;
        reinterpret_cast<uint32_t*>(sm + L.w3h(net))[(net ? 0 : q * 32) + j2] = pack2<FMT>(__ldg(src), __ldg(src + 1));
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

    // ---------------- MMA-issue warpgroup: warp c of it serves compute warpgroup c ------------------------------------
    // Registers move from this warpgroup to the compute ones (setmaxnreg): the pool is what the launch allocated
    // (640 x 96 for NP = 2, 384 x 168 for NP = 1), so the release here must cover the compute warpgroups' request.
    if (warp >= NC * 4) {
        if constexpr (NP == 2) asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        const int c = warp - NC * 4;
        if (c < NC) {
            const int g = c >> 1;
            WgIssue D;
            D.wb = smem_u32(sm + L.wg0 + c * L.wg_stride); D.sb = smem_u32(sm); D.tmem = tmem; D.g = g; D.net = c & 1;
            D.fresh = 0xFu;
            uint64_t *doneZ = bars + 2 * c, *doneW = doneZ + 1, *ready = bars + 2 * NC + c;
            uint32_t pr = 0u;
            const int stride = (int)gridDim.x * NP;
            const int tile0 = g * (int)gridDim.x + (int)blockIdx.x;
            int tiles_done = 0;
            if (tile0 < ntiles) {
                mbar_wait(ready, pr); pr ^= 1u;
                issue_stage<FMT, NP, 4>(D, L, 0, doneZ, doneW);
            }
            for (int tile = tile0; tile < ntiles; tile += stride) {
                const bool has_next = tile + stride < ntiles;
                const int par = tiles_done & 1;
                tiles_done++;
                mbar_wait(ready, pr); pr ^= 1u;
                issue_stage<FMT, NP, 0>(D, L, par, doneZ, doneW);
                if (KT == 0 && (c & 1) == 0) {   // generic actor: the output layer's MMA between the two halves of P2
                    mbar_wait(ready, pr); pr ^= 1u;
                    issue_stage<FMT, NP, 5, true>(D, L, par, doneZ, doneW);
                }
                mbar_wait(ready, pr); pr ^= 1u;
                issue_stage<FMT, NP, 1, KT == 0>(D, L, par, doneZ, doneW);
                mbar_wait(ready, pr); pr ^= 1u;
                issue_stage<FMT, NP, 2, KT == 0>(D, L, par, doneZ, doneW);
                mbar_wait(ready, pr); pr ^= 1u;
                issue_stage<FMT, NP, 3>(D, L, par, doneZ, doneW, has_next);
            }
        }
        return;
    }
    if constexpr (NP == 2) asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");

    RowSumsF rs;
    float gb3c = 0.0f;  // critic output-bias gradient partial of this thread
    {
        const int c = warp >> 2, wt = tid & 127, w4 = warp & 3;
        if (c & 1) tc_compute_wg<FMT, NP, 1, KT>(a, L, sm, tmem, c >> 1, wt, w4, ntiles, rs, gb3c, ts_buf);
        else tc_compute_wg<FMT, NP, 0, KT>(a, L, sm, tmem, c >> 1, wt, w4, ntiles, rs, gb3c, ts_buf);
    }

    // ---------------- drain: tensor-memory accumulators -> this CTA's slab (x 1/n), diagnostics ----------------
    TC_TS(2);
    tc_fence_before();
    bar_sync(3, NC * 128);
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
        // tile streams that owned no tile never wrote their accumulators: skip them
        int ng = 0;
        for (int g = 0; g < NP; g++)
            if (g * (int)gridDim.x + (int)blockIdx.x < ntiles) ng = g + 1;
        {   // dW2 rows -> slab through a per-warp shared-memory transpose: a thread owns a whole 64-float row, so direct
            // stores would hit 32 sectors per instruction; staged, every store instruction covers one row segment.
            // Two 32-column halves keep the live registers inside the 112 of the NP = 2 build.
            float* buf = reinterpret_cast<float*>(sm + L.wg0 + L.wg_stride) + warp * (32 * 65);
#pragma unroll
            for (int hlf = 0; hlf < 2; hlf++) {
                float v[32], w[32];
#pragma unroll
                for (int g = 0; g < NP; g++) {
                    if (g >= ng) break;
                    const uint32_t base = la + tc_cols(NP, g, 0, L.n3a).w2 + hlf * 32;
                    tmem_ld16(base, g == 0 ? v : w);
                    tmem_ld16(base + 16, (g == 0 ? v : w) + 16);
                    tmem_ld_wait();
                    if (g > 0)
#pragma unroll
                        for (int k = 0; k < 32; k++) v[k] += w[k];
                }
#pragma unroll
                for (int k = 0; k < 32; k++) buf[lane * 65 + hlf * 32 + k] = v[k] * sc;
            }
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
        for (int g = 0; g < NP; g++) {
            if (g >= ng) break;
            tmem_ld8(la + tc_cols(NP, g, 0, L.n3a).xe2, g == 0 ? e : f);
            tmem_ld_wait();
            if (g > 0)
#pragma unroll
                for (int i = 0; i < 8; i++) e[i] += f[i];
        }
#pragma unroll
        for (int i = 0; i < 8; i++)
            if (i == O) slab[of.b2 + n] = e[i] * sc;
#pragma unroll
        for (int g = 0; g < NP; g++) {
            if (g >= ng) break;
            tmem_ld8(la + tc_cols(NP, g, 0, L.n3a).xe1, g == 0 ? e : f);
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
            for (int g = 0; g < NP; g++) {
                if (g >= ng) break;
                tmem_ld8(la + tc_cols(NP, g, 0, L.n3a).w3 + cc * 8, g == 0 ? e : f);
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
    constexpr int NT = NC * 128;
    double* red = reinterpret_cast<double*>(sm + L.wg0);                    // [5][NT]   (block of warpgroup 0)
    float* gbc = reinterpret_cast<float*>(sm + L.wg0 + 5 * NT * 8);         // [NT]
    static_assert(5 * NT * 8 + NT * 4 <= 2 * ACT_BYTES + 4 * TC_ROWS * 16, "diagnostic scratch overlaps the db3 partials");
    {
        red[0 * NT + tid] = (double)rs.pg; red[1 * NT + tid] = (double)rs.v; red[2 * NT + tid] = (double)rs.ent;
        red[3 * NT + tid] = (double)rs.kl; red[4 * NT + tid] = (double)rs.clip;
        gbc[tid] = gb3c;
    }
    bar_sync(3, NC * 128);
    // one warp per quantity (5 diagnostics, the critic's output-bias gradient, K actor output-bias gradients), dealt round
    // robin over the NC * 4 compute warps: lane-strided partial sums, then a fixed xor tree (deterministic)
    for (int item = warp; item < 6 + K; item += NC * 4) {
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
            for (int g = 0; g < NP; g++) {
                const float* b = reinterpret_cast<const float*>(sm + L.wg0 + (2 * g) * L.wg_stride + L.b3);
                for (int m = lane; m < TC_ROWS; m += 32) acc += b[m * L.ldo + q];
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) slab[offs[0].b3 + q] = acc * sc;
        }
    }
    tc_fence_before();
    bar_sync(3, NC * 128);
    TC_TS(4);
    if (cta_ts) { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); cta_ts[1] = t; }
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int FMT, int NP, int KT>
static int launch_tc(UpdArgs& a, int* grid_out, cudaStream_t st) {
    const TcLayout L = tc_layout(a.d.K, NP, KT == 0);
    const size_t smem = (size_t)L.end + 1024;
    PPO_REQUIRE(smem <= 227 * 1024, "update_tc: needs %zu B of shared memory", smem);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(update_tc_kernel<FMT, NP, KT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      227 * 1024));
        attr_done = true;
    }
    const int ntiles = (a.mb_size + TC_ROWS - 1) / TC_ROWS;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    // tiles are dealt CTA-first (stream g of CTA b starts at tile g * grid + b): a partial last round leaves ONE
    // busy stream on many SMs instead of two on half as many
    int grid = ntiles;
    if (grid > sms) grid = sms;
    if (grid > UPD_MAX_GRID) grid = UPD_MAX_GRID;
    a.tail.on = 0;
    launch_pdl(a.pdl != 0, update_tc_kernel<FMT, NP, KT>, dim3(grid), dim3((2 * NP + 1) * 128), smem, st, a);
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
// generic-path envelope: shared memory, and the tensor-memory columns (Z, gradient accumulators, logits) within 512
static bool tc_fits(int K, int np) {
    const TcLayout L = tc_layout(K, np, true);
    return (size_t)L.end + 1024 <= 227 * 1024 && tc_lgcol(np, np - 1, L.n3a) + 32 <= 512;
}
bool update_tc_supported(const Dims& d) { return d.H == TC_H && d.O + 1 <= 8 && d.K <= TC_MAX_K && tc_fits(d.K, 1); }
int launch_update_tc(UpdArgs& a, int fmt, int* grid_out, cudaStream_t st, const FuseAdam* fuse, float* grads_out,
                     float* diag_out, bool* tail_done) {
    (void)fuse; (void)grads_out; (void)diag_out;
    if (tail_done) *tail_done = false;   // the reduction + optimiser tail is always the caller's second launch
    if (!update_tc_supported(a.d)) return PPO_ERR_UNSUPPORTED;
    // single unmasked head of 2 or 3 logits (PPO_Discrete on CartPole / MountainCar): K is a template parameter, two tile
    // streams per CTA (its layout always fits: checked at compile time below)
    const int kt = (a.d.heads == 1 && a.mask == nullptr && (a.d.K == 2 || a.d.K == 3)) ? a.d.K : 0;
    static_assert(tc_layout(3, 2).end + 1024 <= 227 * 1024, "the single-head layout must fit two tile streams");
    if (kt == 2) return fmt == FMT_BF16 ? launch_tc<FMT_BF16, 2, 2>(a, grid_out, st) : launch_tc<FMT_F16, 2, 2>(a, grid_out, st);
    if (kt == 3) return fmt == FMT_BF16 ? launch_tc<FMT_BF16, 2, 3>(a, grid_out, st) : launch_tc<FMT_F16, 2, 3>(a, grid_out, st);
    // any other head layout: the generic path, with two tile streams per CTA (4 compute warpgroups) when the per-row logit
    // scratch and the tensor-core operands of the output layer fit next to the activation tiles, else one (the 21-logit
    // MultiDiscrete shape nvec {3,3,3,3,3,2,2,2})
    if (tc_fits(a.d.K, 2))
        return fmt == FMT_BF16 ? launch_tc<FMT_BF16, 2, 0>(a, grid_out, st) : launch_tc<FMT_F16, 2, 0>(a, grid_out, st);
    return fmt == FMT_BF16 ? launch_tc<FMT_BF16, 1, 0>(a, grid_out, st) : launch_tc<FMT_F16, 1, 0>(a, grid_out, st);
}
}  // namespace ppo

// which update_tc_kernel variant serves a shape: 10 * (tile streams per CTA) + (K as a template parameter, 0 = generic head
// path); -1 when the shape is outside the tensor-core envelope.  Host logic only (tests/test_host_facade.py pins the
// single-head shapes to the two-stream template: a dispatch slip once sent them to the generic one-stream kernel at half speed).
extern "C" int ppo_debug_update_tc_variant(const ppo_dims* dims, int masked) {
    ppo::Dims d;
    if (ppo::make_dims(dims, &d) != PPO_OK || !ppo::update_tc_supported(d)) return -1;
    if (d.heads == 1 && !masked && (d.K == 2 || d.K == 3)) return 20 + d.K;
    return ppo::tc_fits(d.K, 2) ? 20 : 10;
}

extern "C" int ppo_debug_fuse_tail(int on) {
    ppo::g_fuse_tail = on;
    return PPO_OK;
}
