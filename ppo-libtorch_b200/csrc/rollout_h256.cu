// rollout_h256.cu — the fused rollout (PPO/PPO_Discrete.cpp:524-554) for hidden = 256 with the 256 x 256 layer on the
// tensor core, in the 16-bit arithmetic mode (fp16 / bf16 operands, fp32 accumulation — the reference's own CUDA mode is
// TF32, PPO_Discrete.cpp:56; PPO_MATH_FP32 keeps the fp32 FMA kernels of rollout.cu).
//
// W2 of ONE net fills 128 KB of shared memory (h256.cuh), so the rollout is cut along the one seam the reference's loop
// offers: the critic's values are not consumed until calcAdvantage (:274-331).  Hence
//   rollout_h256_kernel<Env>  T steps of { actor forward -> Categorical sample -> env step -> rollout row } with the
//                             ACTOR's W2 resident; a 128-env tile = the M = 128 rows of the layer-2 MMA; four
//                             compute warpgroups split the 256 hidden units (thread = env x 64-unit panel), the thread
//                             of panel 0 owns the env (state, counters) and runs sampling + the env step;
//   value_h256_kernel         one batched critic forward over the T x N stored observations and the N bootstrap
//                             observations (m_values, next_value) with the CRITIC's W2 resident.
// Layer 1 (K = obs_dim) runs on the CUDA cores in fp32; the arithmetic of the actor forward is instruction for
// instruction the one update_h256_kernel replays, so the first minibatch sees ratio = 1 up to the softmax routine.
#include "categorical.cuh"
#include "h256.cuh"
#include "rollout.cuh"
#include "umma.cuh"

namespace ppo {
using namespace umma;

namespace {

struct RhLayout {   // byte offsets into the 1024-aligned dynamic shared memory
    int w2, x, p1, b2, w3, b3, part, lg, xs, bars, tmem_slot, end;
    int op, ldp;
};
__host__ __device__ inline RhLayout rh_layout(int O, int K) {
    RhLayout L;
    L.op = (O + 1 <= 4) ? 4 : 8;
    L.ldp = K | 1;
    L.w2 = 0;
    L.x = 4 * H_PANEL_W;
    int o = L.x + 4 * H_PANEL_X;
    L.p1 = o; o += HH * L.op * 4;
    L.b2 = o; o += HH * 4;
    L.w3 = o; o += K * HH * 4;
    L.b3 = o; o += 8 * 4;
    L.part = o; o += 4 * HR * L.ldp * 4;   // partial logits [panel][row][ldp]
    L.lg = o; o += HR * L.ldp * 4;         // the row's logits
    L.xs = o; o += HR * 8 * 4;             // the tile's observations [row][8]
    o = (o + 15) & ~15;
    L.bars = o; o += 4 * 8;
    L.tmem_slot = o; o += 16;
    L.end = o;
    return L;
}

// small parameters of one net -> shared memory (fp32)
template <int OP>
__device__ __forceinline__ void stage_small(uint8_t* sm, const RhLayout& L, const float* params, const NetOff& of, int O,
                                            int k3, int tid, int nthreads) {
    float* P1 = reinterpret_cast<float*>(sm + L.p1);
    for (int e = tid; e < HH * OP; e += nthreads) {   // P1[j] = {W1[j][0..O), b1[j], 0..}
        const int j = e / OP, i = e - j * OP;
        float v = 0.0f;
        if (i < O) v = __ldg(params + of.W1 + j * O + i);
        else if (i == O) v = __ldg(params + of.b1 + j);
        P1[e] = v;
    }
    float* B2 = reinterpret_cast<float*>(sm + L.b2);
    for (int e = tid; e < HH; e += nthreads) B2[e] = __ldg(params + of.b2 + e);
    float* W3 = reinterpret_cast<float*>(sm + L.w3);
    for (int e = tid; e < k3 * HH; e += nthreads) W3[e] = __ldg(params + of.W3 + e);
    float* B3 = reinterpret_cast<float*>(sm + L.b3);
    if (tid < 8) B3[tid] = tid < k3 ? __ldg(params + of.b3 + tid) : 0.0f;
}

// P1 of one row: h1 = tanh(P1 xe) for the 64 hidden units of panel q -> X (16-bit)
template <int FMT, int OP>
__device__ __forceinline__ void fwd_p1(const float* P1q, const float (&xe)[OP], uint8_t* Xq, int wt) {
#pragma unroll
    for (int c = 0; c < 8; c++) {
        float h[8];
        layer1_8<OP>(P1q + c * 8 * OP, xe, h);
        st8<FMT>(Xq + act_chunk_off(wt, c), h);
    }
}
// P2 of one row: h2 = tanh(Z + b2) for panel q's 64 units, partial output-layer sums lg[k] over them
template <int KMAX>
__device__ __forceinline__ void fwd_p2(uint32_t zaddr, const float* B2q, const float* W3q, int k3, float (&lg)[KMAX]) {
#pragma unroll
    for (int k = 0; k < KMAX; k++) lg[k] = 0.0f;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        float v[16];
        tmem_ld16(zaddr + c * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int i4 = 0; i4 < 4; i4++) {
            const float4 b = *reinterpret_cast<const float4*>(B2q + c * 16 + i4 * 4);
            v[i4 * 4 + 0] = tanh_fast256(v[i4 * 4 + 0] + b.x); v[i4 * 4 + 1] = tanh_fast256(v[i4 * 4 + 1] + b.y);
            v[i4 * 4 + 2] = tanh_fast256(v[i4 * 4 + 2] + b.z); v[i4 * 4 + 3] = tanh_fast256(v[i4 * 4 + 3] + b.w);
        }
#pragma unroll
        for (int k = 0; k < KMAX; k++) {
            if (k < k3) {
                float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
                for (int i4 = 0; i4 < 4; i4++) {
                    const float4 w = *reinterpret_cast<const float4*>(W3q + k * HH + c * 16 + i4 * 4);
                    s0 = fmaf(v[i4 * 4 + 0], w.x, s0); s1 = fmaf(v[i4 * 4 + 1], w.y, s1);
                    s0 = fmaf(v[i4 * 4 + 2], w.z, s0); s1 = fmaf(v[i4 * 4 + 3], w.w, s1);
                }
                lg[k] += s0 + s1;
            }
        }
    }
}
// the layer-2 GEMM of a tile: Z[128 x 256] = X W2^T, 16 MMAs issued by one lane
template <int FMT>
__device__ __forceinline__ void issue_fwd(uint32_t X, uint32_t W2, uint32_t tmem, uint64_t* doneZ) {
    constexpr uint32_t id_fwd = idesc_f16(FMT, 128, 256, 0, 0);
#pragma unroll
    for (int p = 0; p < 4; p++)
#pragma unroll
        for (int ks = 0; ks < 4; ks++)
            mma_f16(tmem, act_desc_kmajor(X + (uint32_t)p * H_PANEL_X, ks), act_desc_kmajor(W2 + (uint32_t)p * H_PANEL_W, ks),
                    id_fwd, (p | ks) ? 1u : 0u);
    mma_commit(doneZ);
}

template <class Env, int FMT>
__global__ void __launch_bounds__(H_NTHR, 1) rollout_h256_kernel(const RollArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int S = Env::S, O = Env::OBS, OP = (O + 1 <= 4) ? 4 : 8;
    static_assert(S == O && O <= 7, "observation = state, at most 7 floats");
    const int K = a.d.K, heads = a.d.heads;
    const RhLayout L = rh_layout(O, K);
    const NetOff of = actor_off(O, HH, K);
    const int n = a.cfg.n_envs, T = a.cfg.n_steps;
    const bool masked = a.cfg.masked != 0;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);   // ready (512 arrivals), doneZ (1)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const uint64_t ctr0 = a.cfg.counter + (a.cfg.d_counter ? *a.cfg.d_counter : 0ull);
    const int ntiles = (n + HR - 1) / HR;

    if (warp == 0) tmem_alloc(tmem_slot, 256);
    if (tid == 0) {
        mbar_init(bars + 0, H_NCOMP);
        mbar_init(bars + 1, 1);
        mbar_init_fence();
    }
    stage_w2_panels<FMT>(sm + L.w2, a.params + of.W2, tid, H_NTHR);
    stage_small<OP>(sm, L, a.params, of, O, K, tid, H_NTHR);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    uint64_t *ready = bars, *doneZ = bars + 1;

    if (warp >= 16) {   // control warpgroup: warp 16 issues one layer-2 GEMM per (tile, step)
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        if (warp == 16) {
            const uint32_t X = smem_u32(sm + L.x), W2 = smem_u32(sm + L.w2);
            uint32_t ph = 0u;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
                for (int step = 0; step < T; step++) {
                    mbar_wait(ready, ph); ph ^= 1u;
                    tc_fence_after();
                    if (lane == 0) issue_fwd<FMT>(X, W2, tmem, doneZ);
                    __syncwarp();
                }
        }
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");

    const int wt = tid & 127, q = warp >> 2, w4 = warp & 3;
    const bool owner = q == 0;
    uint8_t* Xq = sm + L.x + q * H_PANEL_X;
    const float* P1q = reinterpret_cast<const float*>(sm + L.p1) + (q * 64) * OP;
    const float* B2q = reinterpret_cast<const float*>(sm + L.b2) + q * 64;
    const float* W3q = reinterpret_cast<const float*>(sm + L.w3) + q * 64;
    const float* B3 = reinterpret_cast<const float*>(sm + L.b3);
    float* part = reinterpret_cast<float*>(sm + L.part);
    float* x = reinterpret_cast<float*>(sm + L.lg) + wt * L.ldp;
    float* xs = reinterpret_cast<float*>(sm + L.xs) + wt * 8;
    const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)q * 64u);
    uint32_t pz = 0u;
    double fin_cnt = 0.0, fin_len = 0.0, fin_ret = 0.0;

    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int e = tile * HR + wt;
        const bool on = e < n;
        float s[S];
        int32_t len = 0;
        float ret = 0.0f, nd = 0.0f;
#pragma unroll
        for (int k = 0; k < S; k++) s[k] = 0.0f;
        if (owner && on) {
#pragma unroll
            for (int k = 0; k < S; k++) s[k] = a.state[(size_t)k * n + e];
            len = a.ep_len[e];
            ret = a.ep_ret[e];
            nd = a.next_done[e];
        }
        const uint32_t env_id = a.cfg.env_id_base + (uint32_t)e;
        for (int step = 0; step < T; step++) {
            const size_t row = (size_t)step * n + e;
            if (owner) {
                // m_obs[step] = next_obs; m_dones[step] = next_done   (:529-530)
                if (on) {
                    if constexpr (O == 4) *reinterpret_cast<float4*>(a.obs_buf + row * 4) = make_float4(s[0], s[1], s[2], s[3]);
                    else if constexpr (O == 2) *reinterpret_cast<float2*>(a.obs_buf + row * 2) = make_float2(s[0], s[1]);
                    else
                        for (int k = 0; k < O; k++) a.obs_buf[row * O + k] = s[k];
                    a.dones_buf[row] = nd;
                }
#pragma unroll
                for (int k = 0; k < S; k++) xs[k] = s[k];
            }
            bar_sync(1, H_NCOMP);
            {
                float xe[OP];
#pragma unroll
                for (int i = 0; i < OP; i++) xe[i] = i < O ? xs[i < O ? i : 0] : (i == O ? 1.0f : 0.0f);
                fwd_p1<FMT, OP>(P1q, xe, Xq, wt);
            }
            tc_fence_before();   // this thread's tensor-memory reads of the previous step are complete
            fence_async_smem();
            mbar_arrive(ready);
            mbar_wait(doneZ, pz); pz ^= 1u;
            tc_fence_after();
            {
                float lg[H_MAX_K];
                fwd_p2<H_MAX_K>(zaddr, B2q, W3q, K, lg);
                float* pp = part + (q * HR + wt) * L.ldp;
#pragma unroll
                for (int k = 0; k < H_MAX_K; k++)
                    if (k < K) pp[k] = lg[k];
            }
            bar_sync(1, H_NCOMP);
            if (owner && on) {
                for (int k = 0; k < K; k++) {
                    float v = B3[k];
#pragma unroll
                    for (int qq = 0; qq < 4; qq++) v += part[(qq * HR + wt) * L.ldp + k];
                    x[k] = v;
                }
                int32_t acts[PPO_MAX_HEADS];
                const uint64_t ctr = ctr0 + (uint64_t)step;
                const uint64_t mbits = masked ? (uint64_t)Env::action_mask(s) : ~0ull;
                const RowOut ro = row_eval(x, heads, a.d.nvec, mbits, masked, nullptr, nullptr, a.cfg.seed, env_id, ctr,
                                           a.cfg.entropy_mode, acts);
                for (int h = 0; h < heads; h++) a.actions_buf[row * heads + h] = acts[h];
                a.logprobs_buf[row] = ro.logp;
                if (a.mask_buf)
                    for (int j = 0; j < K; j++) a.mask_buf[row * K + j] = (uint8_t)((mbits >> j) & 1ull);
                // stepEnvs (:413-483): step, truncate, auto-reset, episode statistics
                StepOut so = Env::step_multi(s, acts);
                len += 1;
                ret = __fadd_rn(ret, so.reward);
                const bool fin = so.terminated || (len == a.cfg.max_episode_steps);
                const int32_t f_len = fin ? len : 0;
                const float f_ret = fin ? ret : 0.0f;
                if (fin) {
                    fin_cnt += 1.0; fin_len += (double)len; fin_ret += (double)ret;
                    float u[4];
                    reset_uniforms(a.cfg.seed, env_id, ctr, u);
                    Env::reset(s, u);
                    len = 0; ret = 0.0f;
                }
                a.rewards_buf[row] = so.reward;
                if (a.cfg.fin_len) { a.cfg.fin_len[row] = f_len; a.cfg.fin_ret[row] = f_ret; }
                nd = fin ? 1.0f : 0.0f;
            }
        }
        if (owner && on) {
#pragma unroll
            for (int k = 0; k < S; k++) a.state[(size_t)k * n + e] = s[k];
            a.ep_len[e] = len;
            a.ep_ret[e] = ret;
            a.next_done[e] = nd;
        }
    }
    tc_fence_before();
    bar_sync(1, H_NCOMP);
    if (a.ep_stats && owner) {   // finished-episode statistics: warp sums, one atomic per warp and quantity
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            fin_cnt += __shfl_xor_sync(0xffffffffu, fin_cnt, o);
            fin_len += __shfl_xor_sync(0xffffffffu, fin_len, o);
            fin_ret += __shfl_xor_sync(0xffffffffu, fin_ret, o);
        }
        if (lane == 0 && fin_cnt != 0.0) {
            atomicAdd(a.ep_stats + 0, fin_cnt);
            atomicAdd(a.ep_stats + 1, fin_len);
            atomicAdd(a.ep_stats + 2, fin_ret);
        }
    }
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem, 256);
    }
}

// critic forward over rows_a observation rows [rows_a][O] (-> out_a) followed by rows_b rows read from a struct-of-arrays
// state [O][rows_b] (-> out_b): m_values of the whole rollout and the bootstrap value in one launch
struct ValArgs {
    const float* params;
    Dims d;
    const float* obs_a;
    float* out_a;
    int64_t rows_a;
    const float* soa_b;
    float* out_b;
    int rows_b;
};
template <int FMT, int OP>
__global__ void __launch_bounds__(H_NTHR, 1) value_h256_kernel(const ValArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int O = a.d.O;
    const RhLayout L = rh_layout(O, 1);
    const NetOff of = critic_off(O, HH);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const int64_t rows = a.rows_a + a.rows_b;
    const int ntiles = (int)((rows + HR - 1) / HR);

    if (warp == 0) tmem_alloc(tmem_slot, 256);
    if (tid == 0) {
        mbar_init(bars + 0, H_NCOMP);
        mbar_init(bars + 1, 1);
        mbar_init_fence();
    }
    stage_w2_panels<FMT>(sm + L.w2, a.params + of.W2, tid, H_NTHR);
    stage_small<OP>(sm, L, a.params, of, O, 1, tid, H_NTHR);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    uint64_t *ready = bars, *doneZ = bars + 1;

    if (warp >= 16) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        if (warp == 16) {
            const uint32_t X = smem_u32(sm + L.x), W2 = smem_u32(sm + L.w2);
            uint32_t ph = 0u;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                mbar_wait(ready, ph); ph ^= 1u;
                tc_fence_after();
                if (lane == 0) issue_fwd<FMT>(X, W2, tmem, doneZ);
                __syncwarp();
            }
        }
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");

    const int wt = tid & 127, q = warp >> 2, w4 = warp & 3;
    uint8_t* Xq = sm + L.x + q * H_PANEL_X;
    const float* P1q = reinterpret_cast<const float*>(sm + L.p1) + (q * 64) * OP;
    const float* B2q = reinterpret_cast<const float*>(sm + L.b2) + q * 64;
    const float* W3q = reinterpret_cast<const float*>(sm + L.w3) + q * 64;
    const float b3 = reinterpret_cast<const float*>(sm + L.b3)[0];
    float* part = reinterpret_cast<float*>(sm + L.part);
    const uint32_t zaddr = tmem_addr(tmem, (uint32_t)w4 * 32u, (uint32_t)q * 64u);
    uint32_t pz = 0u;

    auto load_xe = [&](int tile, float (&xe)[OP]) {
        const int64_t r = (int64_t)tile * HR + wt;
#pragma unroll
        for (int i = 0; i < OP; i++) xe[i] = 0.0f;
        if (tile < ntiles && r < rows) {
            if (r < a.rows_a) {
#pragma unroll
                for (int i = 0; i < OP; i++)
                    if (i < O) xe[i] = __ldg(a.obs_a + r * O + i);
            } else {
                const int64_t eb = r - a.rows_a;
#pragma unroll
                for (int i = 0; i < OP; i++)
                    if (i < O) xe[i] = __ldg(a.soa_b + (size_t)i * a.rows_b + eb);
            }
        }
#pragma unroll
        for (int i = 0; i < OP; i++)
            if (i == O) xe[i] = 1.0f;
    };
    float xn[OP];
    load_xe(blockIdx.x, xn);
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        float xe[OP];
#pragma unroll
        for (int i = 0; i < OP; i++) xe[i] = xn[i];
        load_xe(tile + gridDim.x, xn);       // the next tile's observations fly while this one computes
        fwd_p1<FMT, OP>(P1q, xe, Xq, wt);    // X is free: the previous tile's GEMM was waited for below
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(ready);
        mbar_wait(doneZ, pz); pz ^= 1u;
        tc_fence_after();
        float lg[1];
        fwd_p2<1>(zaddr, B2q, W3q, 1, lg);
        part[q * HR + wt] = lg[0];
        bar_sync(1, H_NCOMP);
        if (q == 0) {
            const int64_t r = (int64_t)tile * HR + wt;
            if (r < rows) {
                const float v = b3 + part[wt] + part[HR + wt] + part[2 * HR + wt] + part[3 * HR + wt];
                if (r < a.rows_a) a.out_a[r] = v;
                else a.out_b[r - a.rows_a] = v;
            }
        }
        bar_sync(2, H_NCOMP);   // the partial sums are consumed before the next tile overwrites them
    }
    tc_fence_before();
    bar_sync(1, H_NCOMP);
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem, 256);
    }
}

}  // namespace

bool rollout_h256_supported(const Dims& d) {
    return d.H == HH && d.O <= 7 && d.K >= 1 && d.K <= H_MAX_K && (size_t)rh_layout(d.O, d.K).end + 1024 <= 227 * 1024;
}

int launch_value_h256(const float* params, const Dims& d, const float* obs_a, float* out_a, int64_t rows_a,
                      const float* soa_b, float* out_b, int rows_b, int fmt, cudaStream_t st) {
    ValArgs v{params, d, obs_a, out_a, rows_a, soa_b, out_b, rows_b};
    const RhLayout L = rh_layout(d.O, 1);
    const size_t smem = (size_t)L.end + 1024;
    const int64_t rows = rows_a + rows_b;
    if (rows <= 0) return PPO_OK;
    const int ntiles = (int)((rows + HR - 1) / HR);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < sms ? ntiles : sms;
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(value_h256_kernel<FMT_F16, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        PPO_CUDA(cudaFuncSetAttribute(value_h256_kernel<FMT_F16, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        PPO_CUDA(cudaFuncSetAttribute(value_h256_kernel<FMT_BF16, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        PPO_CUDA(cudaFuncSetAttribute(value_h256_kernel<FMT_BF16, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done = true;
    }
    const bool wide = d.O + 1 > 4;
    if (fmt == FMT_BF16) {
        if (wide) value_h256_kernel<FMT_BF16, 8><<<grid, H_NTHR, smem, st>>>(v);
        else value_h256_kernel<FMT_BF16, 4><<<grid, H_NTHR, smem, st>>>(v);
    } else {
        if (wide) value_h256_kernel<FMT_F16, 8><<<grid, H_NTHR, smem, st>>>(v);
        else value_h256_kernel<FMT_F16, 4><<<grid, H_NTHR, smem, st>>>(v);
    }
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

template <class Env>
int launch_rollout_h256(const RollArgs& a, int fmt, cudaStream_t st) {
    const RhLayout L = rh_layout(Env::OBS, a.d.K);
    const size_t smem = (size_t)L.end + 1024;
    PPO_REQUIRE(smem <= 227 * 1024, "rollout_h256: needs %zu B of shared memory", smem);
    const int ntiles = (a.cfg.n_envs + HR - 1) / HR;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < sms ? ntiles : sms;
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(rollout_h256_kernel<Env, FMT_BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        PPO_CUDA(cudaFuncSetAttribute(rollout_h256_kernel<Env, FMT_F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done = true;
    }
    if (a.cfg.n_steps > 0) {
        if (fmt == FMT_BF16) rollout_h256_kernel<Env, FMT_BF16><<<grid, H_NTHR, smem, st>>>(a);
        else rollout_h256_kernel<Env, FMT_F16><<<grid, H_NTHR, smem, st>>>(a);
        PPO_LAUNCH_CHECK();
    }
    // m_values[t] = critic(m_obs[t]) and next_value = critic(next_obs) (calcAdvantage :280) in one batched pass
    return launch_value_h256(a.params, a.d, a.obs_buf, a.values_buf, (int64_t)a.cfg.n_steps * a.cfg.n_envs, a.state,
                             a.next_value, a.cfg.n_envs, fmt, st);
}
template int launch_rollout_h256<CartPoleEnv>(const RollArgs&, int, cudaStream_t);
template int launch_rollout_h256<MountainCarEnv>(const RollArgs&, int, cudaStream_t);
template int launch_rollout_h256<MultiWalkEnv>(const RollArgs&, int, cudaStream_t);

}  // namespace ppo
