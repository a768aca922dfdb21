// update_h256.cu — K5 for hidden = 256 (BASELINE configs[2]: MountainCar 65 536 x 256, 2 x 256 MLP) on the sm_100a
// tensor cores: minibatch forward + PPO loss + backward of PPO/PPO_Discrete.cpp:579-638 for the 256-wide actor /
// critic MLPs of PPO/Agent.cpp:25-59 (the widths the reference keeps as commented alternatives at :27,:32).
//
// A 256-wide layer does not fit update_tc.cu's "both nets in shared memory, every weight-gradient accumulator resident
// in tensor memory" layout: W2 alone is 128 KB of fp16 per net and dW2 is 256 x 256 fp32 = all 512 TMEM columns.  So:
//
//   * one net per CTA.  The actor and the critic are disjoint graphs that share only the observation, so CTA b works on
//     net (b & 1) and keeps that net's W2 (fp16, 128 KB, four 64-column panels in the 128-byte-swizzle layout) resident
//     in shared memory for the whole launch — it is the B operand of the forward GEMM (K-major view) AND of the dX GEMM
//     (MN-major view of the same bytes).
//   * update_h256_kernel walks 128-row tiles of the permuted minibatch through
//         P1  h1 = tanh([x,1] P1^T) (CUDA cores)                       -> X            G1 : Z  = X W2^T     (M128 N256 K256)
//         P2  h2 = tanh(Z + b2), logits / value, loss, dOut            -> X, NB        G5 : dW3^T += X^T NB
//             dZ2 = (dOut W3)(1 - h2^2)                                -> X            G2 : Z  = X W2       (dX)
//                                                                                      G3b: d[.|b2] += X^T NB
//         P4  dZ1 = Z (1 - h1^2)                                       -> X            G4 : d[W1|b1] += X^T NB
//     with ONE 64 KB activation tile X in shared memory (h1, h2, dZ2, dZ1 in turn; h1 / h2 stay in registers as packed
//     16-bit pairs for the derivative factors).  Four compute warpgroups split the 256 columns (thread = row x 64-column
//     panel), a fifth warpgroup's warp issues every tcgen05.mma and the TMA stores.
//   * dW2 = dZ2^T h1 needs both activation tiles at once and a full-TMEM accumulator, so update_h256_kernel writes h1 and
//     dZ2 (fp16) to an HBM scratch with TMA stores straight out of X, and dw2_h256_kernel — a pure TMA-producer /
//     MMA-issuer GEMM, K = minibatch rows, 3-stage mbarrier ring of 64-row slices — accumulates dW2 [256 x 256] in all
//     512 TMEM columns and drains it once per CTA.
//   Gradients land in per-CTA slabs (slab b: net b & 1, the other net's range zero-filled) and go through the same
//   fixed-order slab reduction + clip + AdamW as the other update kernels => run-to-run identical results.
#include <cuda.h>

#include "ppo_loss.cuh"
#include "h256.cuh"
#include "tma.cuh"
#include "umma.cuh"
#include "update.cuh"

namespace ppo {
using namespace umma;


struct HLayout {   // byte offsets into the 1024-aligned dynamic shared memory of update_h256_kernel
    int w2, x, nb, p1, b2, w3, b3, part, doutf, gb, w3h, bars, tmem_slot, end;
    int op, ldp;   // floats per P1 row (4 or 8), logits row stride
};
__host__ __device__ inline HLayout h_layout(int O, int K) {
    HLayout L;
    L.op = (O + 1 <= 4) ? 4 : 8;
    L.ldp = K | 1;
    L.w2 = 0;
    L.x = 4 * H_PANEL_W;
    L.nb = L.x + 4 * H_PANEL_X;                 // narrow tile, N = 16: chunk 0 = dOut, chunk 1 = {x, 1, 0..}
    int o = L.nb + 2 * HR * 16;
    L.p1 = o; o += HH * L.op * 4;
    L.b2 = o; o += HH * 4;
    L.w3 = o; o += (K > 1 ? K : 1) * HH * 4;
    L.b3 = o; o += 8 * 4;
    L.part = o; o += 4 * HR * L.ldp * 4;        // logit partial sums [quarter][row][ldp]
    L.doutf = o; o += HR * L.ldp * 4;           // logits -> dL/dlogits of a row (fp32)
    L.gb = o; o += (K > 1 ? K : 1) * HR * 4;    // per-row-slot running sums of dOut (output-bias gradient)
    o = (o + 15) & ~15;
    L.bars = o; o += 8 * 8;
    L.tmem_slot = o; o += 16;
    // 16-bit W3^T as a K-major narrow tile [2 chunks][256 units][16 B] (chunk 0 = W3[0..8)[j], chunk 1 = 0): B operand of
    // G6, dh2 = dOut W3 on the tensor core.  Only when it fits (-1: P2b forms dh2 with FMAs against the fp32 W3 instead).
    o = (o + 127) & ~127;
    L.w3h = -1;
    if ((size_t)o + 2 * HH * 16 + 1024 <= 227 * 1024) { L.w3h = o; o += 2 * HH * 16; }
    L.end = o;
    return L;
}

// scratch in HBM (one per process, grown outside graph capture): 16-bit [4 regions][rows_pad][256];
// region = net * 2 + {0: h1, 1: dZ2}
struct H256Scratch {
    void* ptr = nullptr;
    int64_t rows_pad = 0;   // rows per region the allocation holds
    int dev = -1;
};
static H256Scratch g_scratch;

struct H256Args {
    UpdArgs u;
    int64_t rows_pad;        // rows per scratch region
    int ntiles;
};

// optional in-kernel timeline (ppo_debug_h256_timeline): SM clock stamps of CTA 0, thread 0 at the phase edges of its
// first 4 tiles, 16 slots per tile (tools/h256_timeline.py prints the deltas)
__device__ long long* g_h256_ts = nullptr;
#ifdef PPO_TC_TIMELINE   // make TIMELINE=1
#define H_TS(slot)                                                                        \
    do {                                                                                  \
        if (ts_buf && tile_no < 4) ts_buf[tile_no * 16 + (slot)] = clock64();             \
    } while (0)
#else
#define H_TS(slot) do { } while (0)
#endif

// tensor-memory map of update_h256_kernel: Z (forward pre-activations, then dX) in columns 0..255; the skinny
// accumulators, 16 columns each, at 256 + 16 * (moment * 2 + jb): moment 0 = X holds h2 (cols 0..K-1 = dW3^T),
// 1 = X holds dZ2 (col 8 + O = db2), 2 = X holds dZ1 (cols 8..8+O-1 = dW1, col 8 + O = db1); jb = hidden units 128 jb ..
__device__ __forceinline__ uint32_t h_acc_col(int moment, int jb) { return 256u + 16u * (uint32_t)(moment * 2 + jb); }

// ---- MMA / TMA issue warp ------------------------------------------------------------------------------------------------
template <int FMT>
__device__ __forceinline__ void h256_issuer(const H256Args& a, const CUtensorMap* tmap, const HLayout& L, uint8_t* sm,
                                            uint64_t* bars, uint32_t tmem, int net, int lane) {
    constexpr uint32_t id_fwd = idesc_f16(FMT, 128, 256, 0, 0);   // Z = X W2^T
    constexpr uint32_t id_dx = idesc_f16(FMT, 128, 256, 0, 1);    // Z = X W2
    constexpr uint32_t id_sm = idesc_f16(FMT, 128, 16, 1, 1);     // acc += X^T NB
    const uint32_t sb = smem_u32(sm);
    const uint32_t X = sb + (uint32_t)L.x, W2 = sb + (uint32_t)L.w2, NB = sb + (uint32_t)L.nb;
    uint64_t *ready = bars, *doneZ = bars + 1, *doneW = bars + 2;
    const int cta = (int)blockIdx.x >> 1, ncta = (int)gridDim.x >> 1;
    const int reg_h1 = (int)((int64_t)(net * 2) * a.rows_pad), reg_dz = (int)((int64_t)(net * 2 + 1) * a.rows_pad);
    uint32_t ph = 0u, first = 1u;
    for (int tile = cta; tile < a.ntiles; tile += ncta) {
        const int row0 = tile * HR;
        // ---- P1 staged: X = h1, NB chunk 1 = {x, 1}.  h1 -> scratch, G1
        mbar_wait(ready, ph); ph ^= 1u;
        tc_fence_after();
        if (lane == 0) {
#pragma unroll
            for (int p = 0; p < 4; p++) tma::store_2d(tmap, X + (uint32_t)p * H_PANEL_X, 64 * p, reg_h1 + row0);
            tma::store_commit();
#pragma unroll
            for (int p = 0; p < 4; p++)
#pragma unroll
                for (int ks = 0; ks < 4; ks++)
                    mma_f16(tmem, act_desc_kmajor(X + (uint32_t)p * H_PANEL_X, ks),
                            act_desc_kmajor(W2 + (uint32_t)p * H_PANEL_W, ks), id_fwd, (p | ks) ? 1u : 0u);
            mma_commit(doneZ);
            tma::store_wait_read();
            mbar_arrive(doneZ);
        }
        __syncwarp();
        // ---- P2a staged: X = h2, NB chunk 0 = dOut.  G5
        mbar_wait(ready, ph); ph ^= 1u;
        tc_fence_after();
        if (lane == 0) {
#pragma unroll
            for (int jb = 0; jb < 2; jb++)
#pragma unroll
                for (int ks = 0; ks < 8; ks++)
                    mma_f16(tmem + h_acc_col(0, jb), mn_desc(X + (uint32_t)jb * 2u * H_PANEL_X, H_PANEL_X, ks),
                            narrow_desc_mnmajor(NB, HR, ks), id_sm, ks > 0 ? 1u : (first ^ 1u));
            mma_commit(doneW);
            if (L.w3h >= 0) {   // G6: Z = NB W3h^T = dOut W3 (K = 16: the {x, 1} chunk meets the zero chunk of W3h)
                mma_f16(tmem, narrow_desc_kmajor(NB, HR * 16), narrow_desc_kmajor(sb + (uint32_t)L.w3h, HH * 16), id_fwd, 0u);
                mma_commit(doneZ);
                mbar_arrive(doneZ);   // doneZ counts two arrivals per phase (the other phases pair the commit with a TMA store)
            }
        }
        __syncwarp();
        // ---- P2b staged: X = dZ2.  dZ2 -> scratch, G2 (dX), G3b (db2)
        mbar_wait(ready, ph); ph ^= 1u;
        tc_fence_after();
        if (lane == 0) {
#pragma unroll
            for (int p = 0; p < 4; p++) tma::store_2d(tmap, X + (uint32_t)p * H_PANEL_X, 64 * p, reg_dz + row0);
            tma::store_commit();
#pragma unroll
            for (int p = 0; p < 4; p++)
#pragma unroll
                for (int ks = 0; ks < 4; ks++)   // K = hidden unit j = 64 p + 16 ks ..: rows of every W2 panel
                    mma_f16(tmem, act_desc_kmajor(X + (uint32_t)p * H_PANEL_X, ks),
                            mn_desc(W2, H_PANEL_W, p * 4 + ks), id_dx, (p | ks) ? 1u : 0u);
            mma_commit(doneZ);
#pragma unroll
            for (int jb = 0; jb < 2; jb++)
#pragma unroll
                for (int ks = 0; ks < 8; ks++)
                    mma_f16(tmem + h_acc_col(1, jb), mn_desc(X + (uint32_t)jb * 2u * H_PANEL_X, H_PANEL_X, ks),
                            narrow_desc_mnmajor(NB, HR, ks), id_sm, ks > 0 ? 1u : (first ^ 1u));
            mma_commit(doneW);
            tma::store_wait_read();
            mbar_arrive(doneZ);
        }
        __syncwarp();
        // ---- P4 staged: X = dZ1.  G4
        mbar_wait(ready, ph); ph ^= 1u;
        tc_fence_after();
        if (lane == 0) {
#pragma unroll
            for (int jb = 0; jb < 2; jb++)
#pragma unroll
                for (int ks = 0; ks < 8; ks++)
                    mma_f16(tmem + h_acc_col(2, jb), mn_desc(X + (uint32_t)jb * 2u * H_PANEL_X, H_PANEL_X, ks),
                            narrow_desc_mnmajor(NB, HR, ks), id_sm, ks > 0 ? 1u : (first ^ 1u));
            mma_commit(doneW);
        }
        __syncwarp();
        first = 0u;
    }
    if (lane == 0) tma::store_wait_all();   // the scratch rows are in HBM before the grid ends
}

// ---- the fused forward / loss / backward kernel ------------------------------------------------------------------------------
// NQ = compute warpgroups (2 or 4): a thread owns row wt of the tile x CPT = 256 / NQ hidden units.  NQ = 4: 16 compute
// warps of 112 registers; NQ = 2: 8 compute warps of 232 registers (room for the compiler to keep whole 16-column groups
// of loads in flight).  ppo_debug_h256_variant picks; the default is what measured faster on B200.
template <int FMT, int OP, int NQ, int KMAX>
__global__ void __launch_bounds__(128 * (NQ + 1), 1)
    update_h256_kernel(const H256Args a, const __grid_constant__ CUtensorMap tmap) {
    constexpr int NCOMP = 128 * NQ, NTHR = 128 * (NQ + 1), CPT = HH / NQ, NCH = CPT / 8, PPT = CPT / 64;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const UpdArgs& u = a.u;
    const int O = u.d.O, K = u.d.K, heads = u.d.heads;
    const HLayout L = h_layout(O, K);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int net = (int)blockIdx.x & 1;                // 0 = actor, 1 = critic
    const int k3 = net == 0 ? K : 1;
    const NetOff of = net == 0 ? actor_off(O, HH, K) : critic_off(O, HH);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);   // ready (NCOMP arrivals), doneZ (2), doneW (1)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const int cta = (int)blockIdx.x >> 1, ncta = (int)gridDim.x >> 1;
    const bool dh_mma = L.w3h >= 0;

    // ---------------- prologue: tensor memory, barriers, this net's parameters -> shared memory ----------------
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    if (tid == 0) {
        mbar_init(bars + 0, NCOMP);
        mbar_init(bars + 1, 2);
        mbar_init(bars + 2, 1);
        mbar_init_fence();
    }
    if (warp == 4 * NQ && lane == 0) tma::prefetch_map(&tmap);
    stage_w2_panels<FMT>(sm + L.w2, u.params + of.W2, tid, NTHR);
    {
        float* P1 = reinterpret_cast<float*>(sm + L.p1);
        for (int e = tid; e < HH * OP; e += NTHR) {   // P1[j] = {W1[j][0..O), b1[j], 0..}
            const int j = e / OP, i = e - j * OP;
            float v = 0.0f;
            if (i < O) v = __ldg(u.params + of.W1 + j * O + i);
            else if (i == O) v = __ldg(u.params + of.b1 + j);
            P1[e] = v;
        }
        float* B2 = reinterpret_cast<float*>(sm + L.b2);
        for (int e = tid; e < HH; e += NTHR) B2[e] = __ldg(u.params + of.b2 + e);
        float* W3 = reinterpret_cast<float*>(sm + L.w3);
        for (int e = tid; e < k3 * HH; e += NTHR) W3[e] = __ldg(u.params + of.W3 + e);
        float* B3 = reinterpret_cast<float*>(sm + L.b3);
        if (tid < 8) B3[tid] = tid < k3 ? __ldg(u.params + of.b3 + tid) : 0.0f;
        if (dh_mma) {   // W3h: unit j's 16-byte chunk = {W3[0][j] .. W3[7][j]} in 16 bits, then a chunk of zeros
            for (int j = tid; j < HH; j += NTHR) {
                float v[8], z[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    v[k] = k < k3 ? __ldg(u.params + of.W3 + k * HH + j) : 0.0f;
                    z[k] = 0.0f;
                }
                st8<FMT>(sm + L.w3h + j * 16, v);
                st8<FMT>(sm + L.w3h + HH * 16 + j * 16, z);
            }
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp >= 4 * NQ) {   // control warpgroup: its first warp issues every MMA and TMA store, the others only give up registers
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        if (warp == 4 * NQ) h256_issuer<FMT>(a, &tmap, L, sm, bars, tmem, net, lane);
        return;
    }
    // registers released by the control warpgroup must cover the request (the pool is the CTA's launch allocation)
    if constexpr (NQ == 4) asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");

    // ---------------- compute threads: row wt of the tile x hidden units CPT q .. CPT q + CPT - 1 -----------------------
    const int wt = tid & 127, q = warp >> 2, w4 = warp & 3;
    const bool owner = q == 0;               // runs the per-row loss and owns the row's scalar inputs
    uint64_t *ready = bars, *doneZ = bars + 1, *doneW = bars + 2;
    uint32_t pz = 0u, pw = 0u;
    uint8_t* X0 = sm + L.x + q * PPT * H_PANEL_X;      // this thread's PPT panels of X; chunk c at panel c / 8, chunk c % 8
    uint8_t* NB = sm + L.nb;
    const float* P1 = reinterpret_cast<const float*>(sm + L.p1) + (q * CPT) * OP;
    const float* B2 = reinterpret_cast<const float*>(sm + L.b2) + q * CPT;
    const float* W3 = reinterpret_cast<const float*>(sm + L.w3) + q * CPT;
    const float* B3 = reinterpret_cast<const float*>(sm + L.b3);
    float* part = reinterpret_cast<float*>(sm + L.part);
    float* xrow = reinterpret_cast<float*>(sm + L.doutf) + wt * L.ldp;   // this row's logits -> dL/dlogits
    const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)(q * CPT));
    const bool masked = u.mask != nullptr;
    float meanf = 0.0f, stdf = 1.0f;
    if (u.hp.norm_adv) { meanf = u.adv_stats[0]; stdf = u.adv_stats[1]; }
    RowSumsF rs;
    float* gbrow = reinterpret_cast<float*>(sm + L.gb) + wt;   // [k][128]: this row slot's running sum of dOut[k]
    if (owner)
        for (int k = 0; k < k3; k++) gbrow[k * HR] = 0.0f;
    auto xchunk = [&](int c) -> uint8_t* { return X0 + (c >> 3) * H_PANEL_X + act_chunk_off(wt, c & 7); };

    // the observation of the NEXT tile's row is gathered one tile ahead (P1 needs it at once); the loss inputs of a row are
    // loaded by its owner thread at the top of the tile and are not needed before the loss, ~10 k cycles later
    struct RowIn {
        int64_t r;
        float o[4];
        bool on;
    };
    uint32_t pkey[4] = {0u, 0u, 0u, 0u};
    if (!u.perm && u.pf.inline_on) perm_keys(u.pf.seed, u.pf.epoch + (u.pf.d_epoch ? *u.pf.d_epoch : 0ull), pkey);
    auto gather = [&](int tile) {
        RowIn in;
        const int rloc = tile * HR + wt;
        in.on = tile < a.ntiles && rloc < u.mb_size;
        in.r = 0;
        in.o[0] = in.o[1] = in.o[2] = in.o[3] = 0.0f;
        if (in.on) {
            const int64_t r = u.perm ? (int64_t)__ldg(u.perm + u.mb_start + rloc)
                                     : (u.pf.inline_on ? perm_index(u.mb_start + rloc, u.pf.batch, u.pf.half_bits, pkey)
                                                       : u.mb_start + rloc);
            in.r = r;
            if (u.rec) {   // packed row: {obs[4]} {old_logp, adv, ret, old_val} {action, -, -, -} {-}
                const float4 q0 = __ldg(u.rec + r * 4);
                in.o[0] = q0.x; in.o[1] = q0.y; in.o[2] = q0.z; in.o[3] = q0.w;
            } else if (O <= 4) {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (i < O) in.o[i] = __ldg(u.obs + r * O + i);
            }
        }
        return in;
    };
    const float rstd = __fdividef(1.0f, stdf + 1e-8f);

    long long* ts_buf = (blockIdx.x == 0 && tid == 0) ? g_h256_ts : nullptr;
    (void)ts_buf;
    int tile_no = -1;
    RowIn nxt = gather(cta);
    bool pendingW = false;   // G4 of the previous tile still reads X / NB
    for (int tile = cta; tile < a.ntiles; tile += ncta) {
        const RowIn cur = nxt;
        nxt = gather(tile + ncta);
        tile_no++;
        H_TS(0);
        float c_old_logp = 0.0f, c_adv = 0.0f, c_ret = 0.0f, c_old_val = 0.0f;
        int c_act0 = 0;
        if (owner && cur.on) {
            if (u.rec) {
                const float4 q1 = __ldg(u.rec + cur.r * 4 + 1);
                c_old_logp = q1.x; c_adv = q1.y; c_ret = q1.z; c_old_val = q1.w;
                c_act0 = __float_as_int(__ldg(reinterpret_cast<const float*>(u.rec + cur.r * 4 + 2)));
            } else {
                c_old_logp = __ldg(u.old_logprob + cur.r);
                c_act0 = __ldg(u.actions + cur.r * heads);
                c_adv = __ldg(u.adv + cur.r);
                c_ret = __ldg(u.ret + cur.r);
                if (u.hp.clip_vloss) c_old_val = __ldg(u.old_val + cur.r);
            }
        }
        // ---- P1: h1 = tanh(P1 [x, 1]) for this thread's hidden units, kept as packed 16-bit pairs
        float xe[OP];
#pragma unroll
        for (int i = 0; i < OP; i++) xe[i] = 0.0f;
        if (O <= 4) {
#pragma unroll
            for (int i = 0; i < 4; i++)
                if (i < OP && i < O) xe[i] = cur.o[i];
        } else if (cur.on) {
#pragma unroll
            for (int i = 0; i < OP; i++)
                if (i < O) xe[i] = __ldg(u.obs + cur.r * O + i);
        }
#pragma unroll
        for (int i = 0; i < OP; i++)
            if (i == O) xe[i] = 1.0f;
        uint32_t h1p[CPT / 2];
#pragma unroll
        for (int c = 0; c < NCH; c++) {
            float h[8];
            layer1_8<OP>(P1 + c * 8 * OP, xe, h);
#pragma unroll
            for (int i = 0; i < 4; i++) h1p[c * 4 + i] = pack2<FMT>(h[2 * i], h[2 * i + 1]);
        }
        H_TS(1);
        if (pendingW) { mbar_wait(doneW, pw); pw ^= 1u; }
        H_TS(2);
#pragma unroll
        for (int c = 0; c < NCH; c++)
            *reinterpret_cast<uint4*>(xchunk(c)) = make_uint4(h1p[c * 4], h1p[c * 4 + 1], h1p[c * 4 + 2], h1p[c * 4 + 3]);
        if (owner) {
            float x8[8];
#pragma unroll
            for (int i = 0; i < 8; i++) x8[i] = i < OP ? xe[i < OP ? i : 0] : 0.0f;
            st8<FMT>(NB + HR * 16 + wt * 16, x8);
        }
        fence_async_smem();
        mbar_arrive(ready);
        H_TS(3);

        // ---- P2a: h2 = tanh(Z + b2) -> X; partial logits; the row's loss; dOut -> NB
        mbar_wait(doneZ, pz); pz ^= 1u;
        tc_fence_after();
        H_TS(4);
        float lg[KMAX];
#pragma unroll
        for (int k = 0; k < KMAX; k++) lg[k] = 0.0f;
#pragma unroll
        for (int c = 0; c < CPT / 16; c++) {
            float v[16];
            tmem_ld16(zaddr + c * 16, v);
            tmem_ld_wait();
#pragma unroll
            for (int i4 = 0; i4 < 4; i4++) {
                const float4 b = *reinterpret_cast<const float4*>(B2 + c * 16 + i4 * 4);
                v[i4 * 4 + 0] = tanh_fast256(v[i4 * 4 + 0] + b.x); v[i4 * 4 + 1] = tanh_fast256(v[i4 * 4 + 1] + b.y);
                v[i4 * 4 + 2] = tanh_fast256(v[i4 * 4 + 2] + b.z); v[i4 * 4 + 3] = tanh_fast256(v[i4 * 4 + 3] + b.w);
            }
#pragma unroll
            for (int k = 0; k < KMAX; k++) {
                if (k < k3) {
                    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
                    for (int i4 = 0; i4 < 4; i4++) {
                        const float4 w = *reinterpret_cast<const float4*>(W3 + k * HH + c * 16 + i4 * 4);
                        s0 = fmaf(v[i4 * 4 + 0], w.x, s0); s1 = fmaf(v[i4 * 4 + 1], w.y, s1);
                        s0 = fmaf(v[i4 * 4 + 2], w.z, s0); s1 = fmaf(v[i4 * 4 + 3], w.w, s1);
                    }
                    lg[k] += s0 + s1;
                }
            }
            st8<FMT>(xchunk(c * 2), v);          // h2 stays in X: G5 reads it, P2b re-reads it
            st8<FMT>(xchunk(c * 2 + 1), v + 8);
        }
        {
            float* pp = part + (q * HR + wt) * L.ldp;
#pragma unroll
            for (int k = 0; k < KMAX; k++)
                if (k < k3) pp[k] = lg[k];
        }
        H_TS(5);
        bar_sync(1, NCOMP);
        H_TS(6);
        if (owner) {
            float xr[KMAX];
#pragma unroll
            for (int k = 0; k < KMAX; k++) {
                xr[k] = 0.0f;
                if (k < k3) {
                    float s = B3[k];
#pragma unroll
                    for (int qq = 0; qq < NQ; qq++) s += part[(qq * HR + wt) * L.ldp + k];
                    xr[k] = s;
                }
            }
            float d8[8];
#pragma unroll
            for (int k = 0; k < 8; k++) d8[k] = 0.0f;
            if (net == 0) {
                if (heads == 1 && !masked) {   // PPO_Discrete: the row stays in registers
                    if (cur.on) actor_row_loss_fast_reg<KMAX>(xr, K, c_act0, c_old_logp, c_adv, u.hp, meanf, rstd, rs);
                    else {
#pragma unroll
                        for (int k = 0; k < KMAX; k++) xr[k] = 0.0f;
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < KMAX; k++)
                        if (k < K) xrow[k] = xr[k];
                    if (cur.on) {
                        const uint64_t mbits = masked ? pack_mask(u.mask + cur.r * K, K) : ~0ull;
                        actor_row_loss(xrow, u.d, masked, mbits, u.actions + cur.r * heads, c_old_logp, c_adv, u.hp, meanf, stdf,
                                       1.0f, rs);
                    } else {
                        for (int k = 0; k < K; k++) xrow[k] = 0.0f;
                    }
#pragma unroll
                    for (int k = 0; k < KMAX; k++) xr[k] = k < K ? xrow[k] : 0.0f;
                }
#pragma unroll
                for (int k = 0; k < KMAX; k++)
                    if (k < K) { d8[k] = xr[k]; gbrow[k * HR] += xr[k]; }
            } else {
                const float gv = cur.on ? critic_row_loss(xr[0], c_ret, c_old_val, u.hp, 1.0f, rs) : 0.0f;
                xr[0] = gv;
                d8[0] = gv;
                gbrow[0] += gv;
            }
            if (!dh_mma) {   // P2b's FMA path reads dL/dlogits of the row from shared memory
#pragma unroll
                for (int k = 0; k < KMAX; k++)
                    if (k < k3) xrow[k] = xr[k];
            }
            st8<FMT>(NB + wt * 16, d8);
        }
        H_TS(7);
        fence_async_smem();
        bar_sync(1, NCOMP);
        tc_fence_before();
        mbar_arrive(ready);
        H_TS(8);

        // ---- P2b: dZ2 = (dOut W3) (1 - h2^2) -> X (over h2, once G5 has read it)
        uint32_t dzp[CPT / 2];
        if (dh_mma) {
            // dOut W3 comes from the tensor core (G6, issued behind G5): this thread only applies the tanh derivative
            mbar_wait(doneZ, pz); pz ^= 1u;
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < CPT / 16; c++) {
                float v[16];
                tmem_ld16(zaddr + c * 16, v);
                tmem_ld_wait();
#pragma unroll
                for (int hc = 0; hc < 2; hc++) {
                    const uint4 hq = *reinterpret_cast<const uint4*>(xchunk(c * 2 + hc));   // h2, written in P2a
                    const uint32_t hw[4] = {hq.x, hq.y, hq.z, hq.w};
#pragma unroll
                    for (int i = 0; i < 4; i++)
                        dzp[c * 8 + hc * 4 + i] = dtanh2<FMT>(v[hc * 8 + 2 * i], v[hc * 8 + 2 * i + 1], hw[i]);
                }
            }
            tc_fence_before();   // Z is read: G2 may overwrite it
        } else {
            float dq[KMAX];
#pragma unroll
            for (int k = 0; k < KMAX; k++) dq[k] = k < k3 ? xrow[k] : 0.0f;
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                float d[8];
#pragma unroll
                for (int i = 0; i < 8; i++) d[i] = 0.0f;
#pragma unroll
                for (int k = 0; k < KMAX; k++) {
                    if (k < k3) {
                        const float4 wa = *reinterpret_cast<const float4*>(W3 + k * HH + c * 8);
                        const float4 wb = *reinterpret_cast<const float4*>(W3 + k * HH + c * 8 + 4);
                        d[0] = fmaf(dq[k], wa.x, d[0]); d[1] = fmaf(dq[k], wa.y, d[1]);
                        d[2] = fmaf(dq[k], wa.z, d[2]); d[3] = fmaf(dq[k], wa.w, d[3]);
                        d[4] = fmaf(dq[k], wb.x, d[4]); d[5] = fmaf(dq[k], wb.y, d[5]);
                        d[6] = fmaf(dq[k], wb.z, d[6]); d[7] = fmaf(dq[k], wb.w, d[7]);
                    }
                }
                const uint4 hq = *reinterpret_cast<const uint4*>(xchunk(c));   // h2, written in P2a
                const uint32_t hw[4] = {hq.x, hq.y, hq.z, hq.w};
#pragma unroll
                for (int i = 0; i < 4; i++) dzp[c * 4 + i] = dtanh2<FMT>(d[2 * i], d[2 * i + 1], hw[i]);
            }
        }
        H_TS(9);
        mbar_wait(doneW, pw); pw ^= 1u;   // G5 has read X (h2) and NB
        H_TS(10);
#pragma unroll
        for (int c = 0; c < NCH; c++)
            *reinterpret_cast<uint4*>(xchunk(c)) = make_uint4(dzp[c * 4], dzp[c * 4 + 1], dzp[c * 4 + 2], dzp[c * 4 + 3]);
        fence_async_smem();
        mbar_arrive(ready);

        // ---- P4: dZ1 = Z (1 - h1^2) -> X (once G2 / G3b have read dZ2 and the dZ2 store has left X)
        H_TS(11);
        mbar_wait(doneZ, pz); pz ^= 1u;
        mbar_wait(doneW, pw); pw ^= 1u;
        tc_fence_after();
        H_TS(12);
#pragma unroll
        for (int c = 0; c < CPT / 16; c++) {
            float v[16];
            tmem_ld16(zaddr + c * 16, v);
            tmem_ld_wait();
            uint32_t o8[8];
#pragma unroll
            for (int i = 0; i < 8; i++) o8[i] = dtanh2<FMT>(v[2 * i], v[2 * i + 1], h1p[c * 8 + i]);
            *reinterpret_cast<uint4*>(xchunk(c * 2)) = make_uint4(o8[0], o8[1], o8[2], o8[3]);
            *reinterpret_cast<uint4*>(xchunk(c * 2 + 1)) = make_uint4(o8[4], o8[5], o8[6], o8[7]);
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(ready);
        H_TS(13);
        pendingW = true;
    }
    if (pendingW) { mbar_wait(doneW, pw); pw ^= 1u; }

    // ---------------- drain: skinny accumulators -> this CTA's slab (x 1/n), diagnostics, output-bias gradients --------
    tc_fence_before();
    bar_sync(1, NCOMP);
    tc_fence_after();
    const int P = num_params(O, HH, K);
    float* slab = u.slabs + (size_t)blockIdx.x * slab_stride(P);
    const float sc = u.inv_n;
    if (q == 0) {   // warps 0..3: thread wt owns hidden units wt and 128 + wt (M = 128 accumulators: row m in lane m)
#pragma unroll
        for (int jb = 0; jb < 2; jb++) {
            const int j = jb * 128 + wt;
            const uint32_t la = tmem_addr(tmem, (uint32_t)w4 * 32u, 0);
            float v[16];
            tmem_ld16(la + h_acc_col(0, jb), v);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k < k3) slab[of.W3 + k * HH + j] = v[k] * sc;
            tmem_ld16(la + h_acc_col(1, jb), v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (i == O) slab[of.b2 + j] = v[8 + i] * sc;
            tmem_ld16(la + h_acc_col(2, jb), v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (i < O) slab[of.W1 + j * O + i] = v[8 + i] * sc;
                if (i == O) slab[of.b1 + j] = v[8 + i] * sc;
            }
        }
    }
    double* red = reinterpret_cast<double*>(sm + L.x);                  // [5][128]   (X is free: G4 has been waited for)
    const float* gb = reinterpret_cast<const float*>(sm + L.gb);        // [k3][128]
    if (owner) {
        red[0 * HR + wt] = (double)rs.pg; red[1 * HR + wt] = (double)rs.v; red[2 * HR + wt] = (double)rs.ent;
        red[3 * HR + wt] = (double)rs.kl; red[4 * HR + wt] = (double)rs.clip;
    }
    bar_sync(1, NCOMP);
    // one warp per quantity (5 diagnostics, k3 output-bias gradients, then the unused diagnostic slots): lane-strided
    // partial sums and a fixed xor tree => deterministic
    for (int item = warp; item < 5 + k3 + 1; item += 4 * NQ) {
        if (item < 5) {
            double acc = 0.0;
            for (int m = lane; m < HR; m += 32) acc += red[item * HR + m];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) u.dpart[(size_t)blockIdx.x * UPD_NDIAG + item] = acc;
        } else if (item < 5 + k3) {
            const int k = item - 5;
            float acc = 0.0f;
            for (int m = lane; m < HR; m += 32) acc += gb[k * HR + m];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) slab[of.b3 + k] = acc * sc;
        } else if (lane < UPD_NDIAG - 5) {
            u.dpart[(size_t)blockIdx.x * UPD_NDIAG + 5 + lane] = 0.0;
        }
    }
    tc_fence_before();
    bar_sync(1, NCOMP);
    if (warp == 0) tmem_dealloc(tmem, 512);
}

// ---- dW2 = dZ2^T h1 over the minibatch rows: TMA-fed split-K GEMM, one net per CTA ------------------------------------------
constexpr int DW_STAGES = 3;
constexpr int DW_SLICE = 64;                          // rows per pipeline stage
constexpr int DW_PANEL = DW_SLICE * 128;              // 8 KB: [64 rows][64 columns] of 16-bit, 128-byte swizzle
constexpr int DW_STAGE_BYTES = 8 * DW_PANEL;          // dZ2 panels 0..3, then h1 panels 0..3
constexpr int DW_NTHR = 192;                          // warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..5 drain

template <int FMT>
__global__ void __launch_bounds__(DW_NTHR, 1) dw2_h256_kernel(const H256Args a, const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const UpdArgs& u = a.u;
    const int O = u.d.O, K = u.d.K;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int net = (int)blockIdx.x & 1;
    const int cta = (int)blockIdx.x >> 1, ncta = (int)gridDim.x >> 1;
    const NetOff of = net == 0 ? actor_off(O, HH, K) : critic_off(O, HH);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + DW_STAGES * DW_STAGE_BYTES);   // full[3], empty[3], done
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    uint64_t *full = bars, *empty = bars + DW_STAGES, *done = bars + 2 * DW_STAGES;
    const int nslices = a.ntiles * (HR / DW_SLICE);
    const int P = num_params(O, HH, K);
    float* slab = u.slabs + (size_t)blockIdx.x * slab_stride(P);

    if (warp == 0) tmem_alloc(tmem_slot, 512);
    if (tid == 0) {
        for (int i = 0; i < 2 * DW_STAGES + 1; i++) mbar_init(bars + i, 1);
        mbar_init_fence();
        tma::prefetch_map(&tmap);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            const int reg_h1 = (int)((int64_t)(net * 2) * a.rows_pad), reg_dz = (int)((int64_t)(net * 2 + 1) * a.rows_pad);
            int it = 0;
            for (int s = cta; s < nslices; s += ncta, it++) {
                const int stg = it % DW_STAGES, round = it / DW_STAGES;
                mbar_wait(empty + stg, (uint32_t)(round & 1) ^ 1u);
                tma::mbar_expect_tx(full + stg, DW_STAGE_BYTES);
                const uint32_t dst = smem_u32(sm + stg * DW_STAGE_BYTES);
                const int row0 = s * DW_SLICE;
#pragma unroll
                for (int p = 0; p < 4; p++) tma::load_2d(&tmap, dst + p * DW_PANEL, full + stg, 64 * p, reg_dz + row0);
#pragma unroll
                for (int p = 0; p < 4; p++) tma::load_2d(&tmap, dst + (4 + p) * DW_PANEL, full + stg, 64 * p, reg_h1 + row0);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: D[jb] (128 hidden units j x 256 inputs i) += dZ2[:, j]^T h1 =====
        constexpr uint32_t id_dw = idesc_f16(FMT, 128, 256, 1, 1);
        int it = 0;
        for (int s = cta; s < nslices; s += ncta, it++) {
            const int stg = it % DW_STAGES, round = it / DW_STAGES;
            mbar_wait(full + stg, (uint32_t)(round & 1));
            tc_fence_after();
            if (lane == 0) {
                const uint32_t base = smem_u32(sm + stg * DW_STAGE_BYTES);
#pragma unroll
                for (int jb = 0; jb < 2; jb++)
#pragma unroll
                    for (int ks = 0; ks < DW_SLICE / 16; ks++)
                        mma_f16(tmem + (uint32_t)jb * 256u, mn_desc(base + (uint32_t)jb * 2u * DW_PANEL, DW_PANEL, ks),
                                mn_desc(base + 4u * DW_PANEL, DW_PANEL, ks), id_dw, (it | ks) ? 1u : 0u);
                mma_commit(empty + stg);
            }
            __syncwarp();
        }
        if (lane == 0) mma_commit(done);
        __syncwarp();
    }
    // every thread: zero the OTHER net's range of this CTA's slab (slab b carries net b & 1 only)
    {
        const int ncrit = critic_size(O, HH);
        const int z0 = net == 0 ? 0 : ncrit, z1 = net == 0 ? ncrit : P;
        for (int i = z0 + tid; i < z1; i += DW_NTHR) slab[i] = 0.0f;
    }
    if (warp >= 2) {
        // ===== drain: 256 x 256 fp32 accumulator -> slab (x 1/n) =====
        mbar_wait(done, 0u);
        tc_fence_after();
        const int w4 = warp & 3;
        const float sc = u.inv_n;
#pragma unroll 1
        for (int jb = 0; jb < 2; jb++) {
            const int j = jb * 128 + w4 * 32 + lane;
            float* dst = slab + of.W2 + (size_t)j * HH;
#pragma unroll 1
            for (int c = 0; c < 16; c++) {
                float v[16];
                tmem_ld16(tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)(jb * 256 + c * 16)), v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; i++) dst[c * 16 + i] = v[i] * sc;
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

// ---- host side -----------------------------------------------------------------------------------------------------------
bool update_h256_supported(const Dims& d) {
    if (d.H != HH || d.O + 1 > 8 || d.K > H_MAX_K || d.K < 1) return false;
    return (size_t)h_layout(d.O, d.K).end + 1024 <= 227 * 1024;
}

int h256_reserve(int64_t mb_rows) {
    const int64_t rows_pad = ((mb_rows + HR - 1) / HR) * HR;
    int dev = 0;
    PPO_CUDA(cudaGetDevice(&dev));
    if (g_scratch.ptr && g_scratch.dev == dev && g_scratch.rows_pad >= rows_pad) return PPO_OK;
    if (g_scratch.ptr) {
        PPO_CUDA(cudaDeviceSynchronize());
        cudaFree(g_scratch.ptr);
        g_scratch = H256Scratch{};
    }
    PPO_REQUIRE(rows_pad * 4 < (1ll << 31), "hidden = 256 update: %lld rows per minibatch is beyond the TMA coordinate range",
                (long long)mb_rows);
    PPO_CUDA(cudaMalloc(&g_scratch.ptr, (size_t)rows_pad * 4 * HH * 2));
    g_scratch.rows_pad = rows_pad;
    g_scratch.dev = dev;
    return PPO_OK;
}

template <int FMT, int OP, int KMAX>
static int launch_h256_t(H256Args& ha, int grid, const CUtensorMap& m128, const CUtensorMap& m64, cudaStream_t st) {
    constexpr int NQ = 4;   // 16 compute warps of 112 registers; 8 warps of 232 (NQ = 2) measured 4 % slower on B200
    const HLayout L = h_layout(ha.u.d.O, ha.u.d.K);
    const size_t smemA = (size_t)L.end + 1024, smemB = (size_t)DW_STAGES * DW_STAGE_BYTES + 128 + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(update_h256_kernel<FMT, OP, NQ, KMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      227 * 1024));
        PPO_CUDA(cudaFuncSetAttribute(dw2_h256_kernel<FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done = true;
    }
    update_h256_kernel<FMT, OP, NQ, KMAX><<<grid, 128 * (NQ + 1), smemA, st>>>(ha, m128);
    PPO_LAUNCH_CHECK();
    dw2_h256_kernel<FMT><<<grid, DW_NTHR, smemB, st>>>(ha, m64);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
template <int FMT, int OP>
static int launch_h256_nq(H256Args& ha, int grid, const CUtensorMap& m128, const CUtensorMap& m64, cudaStream_t st) {
    return ha.u.d.K <= 4 ? launch_h256_t<FMT, OP, 4>(ha, grid, m128, m64, st) : launch_h256_t<FMT, OP, 8>(ha, grid, m128, m64, st);
}

int launch_update_h256(UpdArgs& a, int fmt, int* grid_out, cudaStream_t st) {
    if (!update_h256_supported(a.d)) return PPO_ERR_UNSUPPORTED;
    H256Args ha;
    ha.u = a;
    ha.ntiles = (a.mb_size + HR - 1) / HR;
    const int64_t need = (int64_t)ha.ntiles * HR;
    if (!g_scratch.ptr || g_scratch.rows_pad < need) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        PPO_CUDA(cudaStreamIsCapturing(st, &cs));
        PPO_REQUIRE(cs == cudaStreamCaptureStatusNone,
                    "hidden = 256 update: the activation scratch must be reserved before stream capture");
        if (int rc = h256_reserve(a.mb_size)) return rc;
    }
    ha.rows_pad = g_scratch.rows_pad;
    CUtensorMap m128, m64;
    const int e1 = tma::encode_2d_sw128(&m128, g_scratch.ptr, (uint64_t)g_scratch.rows_pad * 4, HH, HR);
    const int e2 = tma::encode_2d_sw128(&m64, g_scratch.ptr, (uint64_t)g_scratch.rows_pad * 4, HH, DW_SLICE);
    PPO_REQUIRE(e1 == 0 && e2 == 0, "cuTensorMapEncodeTiled failed (%d, %d)", e1, e2);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int per_net = sms / 2;
    if (per_net > ha.ntiles) per_net = ha.ntiles;
    const int grid = 2 * per_net;
    *grid_out = grid;
    const bool wide = a.d.O + 1 > 4;
    if (fmt == FMT_BF16)
        return wide ? launch_h256_nq<FMT_BF16, 8>(ha, grid, m128, m64, st) : launch_h256_nq<FMT_BF16, 4>(ha, grid, m128, m64, st);
    return wide ? launch_h256_nq<FMT_F16, 8>(ha, grid, m128, m64, st) : launch_h256_nq<FMT_F16, 4>(ha, grid, m128, m64, st);
}

}  // namespace ppo

extern "C" int ppo_debug_h256_timeline(long long* d_ts) {
    PPO_CUDA(cudaMemcpyToSymbol(ppo::g_h256_ts, &d_ts, sizeof(d_ts)));
    return PPO_OK;
}
