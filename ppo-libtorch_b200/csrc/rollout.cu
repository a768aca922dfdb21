// rollout.cu — fused rollout: T x (actor-critic forward + Categorical sample + env step + rollout
// store) + bootstrap value in ONE launch, followed by the GAE scan (gae.cu).
// Replaces the rollout phase of PPO_Discrete::train (PPO/PPO_Discrete.cpp:524-554): per step
// ~25 LibTorch op dispatches, a D2H of the actions, a ThreadPool barrier and 3*N tiny H2D copies.
//
// The policy is frozen during a rollout and envs are independent, so a CTA owns a tile of TM envs
// for all T steps: env state, episode counters and the done flag stay in registers; the tile's
// observations feed the shared-memory MLP directly; nothing but the [T,N,...] rollout rows touches HBM.
#include "rollout.cuh"

namespace ppo {

template <class Env, int TM, int H>
__global__ void __launch_bounds__(MLP_THREADS) rollout_kernel(const RollArgs a) {
    extern __shared__ __align__(16) float sm[];
    constexpr int S = Env::S, O = Env::OBS;
    static_assert(S == O, "observation = state for the built-in envs");
    const int K = a.d.K, heads = a.d.heads;
    const TileLayout L = tile_layout(TM, H, O, K);
    const NetOff co = critic_off(O, H), ao = actor_off(O, H, K);
    const int n = a.cfg.n_envs, T = a.cfg.n_steps;
    const bool masked = a.cfg.masked != 0;
    const int t = threadIdx.x;
    double fin_cnt = 0.0, fin_len = 0.0, fin_ret = 0.0;
    const uint64_t ctr0 = a.cfg.counter + (a.cfg.d_counter ? *a.cfg.d_counter : 0ull);

    const int ntiles = (n + TM - 1) / TM;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int e0 = tile * TM;
        const int rows = min(TM, n - e0);
        const int e = e0 + t;
        const bool on = t < rows;
        float s[S];
        int32_t len = 0;
        float ret = 0.0f, nd = 0.0f;
        if (on) {
#pragma unroll
            for (int k = 0; k < S; k++) s[k] = a.state[(size_t)k * n + e];
            len = a.ep_len[e];
            ret = a.ep_ret[e];
            nd = a.next_done[e];
        } else {
#pragma unroll
            for (int k = 0; k < S; k++) s[k] = 0.0f;
        }
        __syncthreads();
        for (int step = 0; step <= T; step++) {
            // m_obs[step] = next_obs; m_dones[step] = next_done   (:529-530)
            if (t < TM) {
#pragma unroll
                for (int k = 0; k < O; k++) sm[L.xs + t * O + k] = s[k];
                if (on && step < T) {
                    const size_t row = (size_t)step * n + e;
                    if constexpr (O == 4)
                        *reinterpret_cast<float4*>(a.obs_buf + row * 4) = make_float4(s[0], s[1], s[2], s[3]);
                    else if constexpr (O == 2)
                        *reinterpret_cast<float2*>(a.obs_buf + row * 2) = make_float2(s[0], s[1]);
                    else
                        for (int k = 0; k < O; k++) a.obs_buf[row * O + k] = s[k];
                    a.dones_buf[row] = nd;
                }
            }
            __syncthreads();
            if (step == T) {  // bootstrap: next_value = critic(next_obs)  (calcAdvantage :280)
                net_forward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1);
                if (on) {
                    a.next_value[e] = sm[L.vs + t];
                    a.next_done[e] = nd;
                }
                break;
            }
            net_forward<TM, H>(sm, L, O, a.params, ao, K, sm + L.ls, L.ldo);  // computeActionLogic (:534)
            net_forward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1);
            if (on) {
                const size_t row = (size_t)step * n + e;
                const uint32_t env_id = a.cfg.env_id_base + (uint32_t)e;
                const uint64_t ctr = ctr0 + (uint64_t)step;
                const uint64_t mbits = masked ? (uint64_t)Env::action_mask(s) : ~0ull;
                int32_t acts[PPO_MAX_HEADS];
                const RowOut o = row_eval(sm + L.ls + t * L.ldo, heads, a.d.nvec, mbits, masked, nullptr, nullptr,
                                          a.cfg.seed, env_id, ctr, a.cfg.entropy_mode, acts);
                for (int h = 0; h < heads; h++) a.actions_buf[row * heads + h] = acts[h];
                a.logprobs_buf[row] = o.logp;
                a.values_buf[row] = sm[L.vs + t];
                if (a.mask_buf)
                    for (int j = 0; j < K; j++) a.mask_buf[row * K + j] = (uint8_t)((mbits >> j) & 1ull);
                // stepEnvs (:413-483): step, truncate, auto-reset, episode statistics
                float u[4] = {0.f, 0.f, 0.f, 0.f};
                float reward, fr = 0.0f;
                int32_t fl = 0;
                StepOut so = Env::step_multi(s, acts);
                len += 1;
                ret = __fadd_rn(ret, so.reward);
                const bool done = so.terminated || (len == a.cfg.max_episode_steps);
                reward = so.reward;
                if (done) {
                    fl = len; fr = ret;
                    fin_cnt += 1.0; fin_len += (double)fl; fin_ret += (double)fr;
                    reset_uniforms(a.cfg.seed, env_id, ctr, u);
                    Env::reset(s, u);
                    len = 0; ret = 0.0f;
                }
                a.rewards_buf[row] = reward;
                if (a.cfg.fin_len) { a.cfg.fin_len[row] = done ? fl : 0; a.cfg.fin_ret[row] = done ? fr : 0.0f; }
                nd = done ? 1.0f : 0.0f;
            }
            // the next iteration's write to xs/ls/vs happens after the barriers inside net_forward
        }
        if (on) {
#pragma unroll
            for (int k = 0; k < S; k++) a.state[(size_t)k * n + e] = s[k];
            a.ep_len[e] = len;
            a.ep_ret[e] = ret;
        }
        __syncthreads();
    }
    // finished-episode statistics of this CTA -> global accumulators (one atomic triple per CTA)
    if (a.ep_stats) {
        __syncthreads();
        double* red = reinterpret_cast<double*>(sm);
        if (t < TM) { red[t] = fin_cnt; red[TM + t] = fin_len; red[2 * TM + t] = fin_ret; }
        __syncthreads();
        if (t < 3) {
            double acc = 0.0;
            for (int m = 0; m < TM; m++) acc += red[t * TM + m];
            if (acc != 0.0) atomicAdd(a.ep_stats + t, acc);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// H = 64 specialisation: the weights never leave the SM.  Each thread keeps one full row of W2 (64 floats) of one
// net in registers for the whole rollout: warp w = (net, half of the 64 hidden units, half of the tile's envs),
// lane = hidden unit, so the 64 x 64 layer of both nets is 64 FMAs per (env, unit) fed by broadcast shared-memory
// reads of h1, with no weight traffic at all after the prologue.  W1 / b1 / b2 sit in registers too, W3 / b3 in
// shared memory.  Four barriers per step; HBM only sees the [T,N,...] rollout rows.
__device__ long long* g_roll_ts = nullptr;   // ppo_debug_rollout_timeline: SM clock of CTA 0 / thread 0 at phase edges
#define R64_TS(i)                                                           \
    do {                                                                    \
        if (ts && step == 2) ts[(i)] = clock64();                           \
    } while (0)
constexpr int R64_LDH = 68;

// Two fp32 FMAs per instruction (sm_100 FFMA2): the scalar FFMA issues at half rate per SMSP, so the 64 x 64 layer is
// FMA-pipe-bound unless its products are paired.  acc.{x,y} += a.{x,y} * b.{x,y}
__device__ inline void ffma2(unsigned long long& acc, unsigned long long a, unsigned long long b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
__device__ inline unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ inline float sum2(unsigned long long v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}
// tanh(x) = 1 - 2 / (exp(2x) + 1) on the special-function unit (ex2 + rcp): absolute error ~1e-7, no branches.
// (tanhf's two-path implementation diverges inside a warp and costs ~4x the instructions.)
__device__ inline float tanh_fast(float x) {
    const float t = exp2f(2.8853900817779268f * x);   // exp(2x); inf / 0 at the extremes give +-1
    return 1.0f - __fdividef(2.0f, t + 1.0f);
}   // activation row stride (floats): 16 B aligned, rows shift by 4 banks

struct R64Layout {
    int h1, h2, w3, b3, xs, ls, vs, end, ldo;
};
__host__ __device__ inline R64Layout r64_layout(int TM, int O, int K) {
    R64Layout L;
    L.ldo = K | 1;
    L.h1 = 0;
    L.h2 = L.h1 + 2 * TM * R64_LDH;
    L.w3 = L.h2 + 2 * TM * R64_LDH;
    L.b3 = L.w3 + (K + 1) * 64;
    L.xs = L.b3 + ((K + 1 + 3) & ~3);
    L.ls = L.xs + ((TM * O + 3) & ~3);
    L.vs = L.ls + ((TM * L.ldo + 3) & ~3);
    L.end = L.vs + TM;
    return L;
}

template <class Env, int TM>
__global__ void __launch_bounds__(256, 2) rollout64_kernel(const RollArgs a) {
    extern __shared__ __align__(16) float sm[];
    constexpr int S = Env::S, O = Env::OBS, H = 64, MH = TM / 2;
    static_assert(S == O && O <= 4 && TM % 8 == 0, "tile shape");
    const int K = a.d.K, heads = a.d.heads;
    const R64Layout L = r64_layout(TM, O, K);
    const NetOff co = critic_off(O, H), ao = actor_off(O, H, K);
    const int n = a.cfg.n_envs, T = a.cfg.n_steps;
    const bool masked = a.cfg.masked != 0;
    const int t = threadIdx.x;
    double fin_cnt = 0.0, fin_len = 0.0, fin_ret = 0.0;
    const uint64_t ctr0 = a.cfg.counter + (a.cfg.d_counter ? *a.cfg.d_counter : 0ull);
    long long* ts = (blockIdx.x == 0 && threadIdx.x == 0) ? g_roll_ts : nullptr;

    // ---- prologue: weights into registers / shared memory (net 0 = actor, 1 = critic) ----
    // layer 2 role: warp = (net, unit half, env half); lane = (group of 4 units jg, quarter of the 64 inputs kq).
    // A thread keeps the 4 x 16 block W2[jbase..+4][kq*16..+16] in registers (as 32 packed pairs): per env it reads only
    // its 16 inputs (4 LDS.128 instead of 16 - shared-memory return bandwidth was the measured limit), forms 4 partial
    // sums and a 3-shuffle transpose-reduce leaves lane kq with the complete sum of unit jbase + kq.
    const int warp = t >> 5, lane = t & 31;
    const int net2 = warp >> 2, mbase2 = (warp & 1) * MH;
    const int kq = lane & 3, jbase = ((warp >> 1) & 1) * 32 + (lane >> 2) * 4;
    const NetOff& o2 = net2 == 0 ? ao : co;
    unsigned long long w2[32];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float* w = a.params + o2.W2 + (jbase + r) * H + kq * 16 + 2 * i;
            w2[r * 8 + i] = pack2(__ldg(w), __ldg(w + 1));
        }
    const float b2 = __ldg(a.params + o2.b2 + jbase + kq);
    // layer 1 role: thread = (unit j1, net, env half)
    const int j1 = t & 63, net1 = (t >> 6) & 1, mbase1 = (t >> 7) * MH;
    const NetOff& o1 = net1 == 0 ? ao : co;
    float w1[O];
#pragma unroll
    for (int i = 0; i < O; i++) w1[i] = __ldg(a.params + o1.W1 + j1 * O + i);
    const float b1 = __ldg(a.params + o1.b1 + j1);
    for (int e = t; e < (K + 1) * 64; e += 256) {
        const int q = e >> 6, k = e & 63;
        sm[L.w3 + e] = q < K ? __ldg(a.params + ao.W3 + q * H + k) : __ldg(a.params + co.W3 + k);
    }
    for (int q = t; q <= K; q += 256) sm[L.b3 + q] = q < K ? __ldg(a.params + ao.b3 + q) : __ldg(a.params + co.b3);

    const int ntiles = (n + TM - 1) / TM;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int e0 = tile * TM;
        const int rows = min(TM, n - e0);
        const int e = e0 + t;
        const bool on = t < rows;
        float s[S];
        int32_t len = 0;
        float ret = 0.0f, nd = 0.0f;
#pragma unroll
        for (int k = 0; k < S; k++) s[k] = on ? a.state[(size_t)k * n + e] : 0.0f;
        if (on) {
            len = a.ep_len[e];
            ret = a.ep_ret[e];
            nd = a.next_done[e];
        }
        __syncthreads();   // W3 / b3 staged; previous tile's readers done
        for (int step = 0; step <= T; step++) {
            // m_obs[step] = next_obs; m_dones[step] = next_done   (:529-530)
            R64_TS(0);
            if (t < TM) {
#pragma unroll
                for (int k = 0; k < O; k++) sm[L.xs + t * O + k] = s[k];
                if (on && step < T) {
                    const size_t row = (size_t)step * n + e;
                    if constexpr (O == 4)
                        *reinterpret_cast<float4*>(a.obs_buf + row * 4) = make_float4(s[0], s[1], s[2], s[3]);
                    else if constexpr (O == 2)
                        *reinterpret_cast<float2*>(a.obs_buf + row * 2) = make_float2(s[0], s[1]);
                    else
                        for (int k = 0; k < O; k++) a.obs_buf[row * O + k] = s[k];
                    a.dones_buf[row] = nd;
                }
            }
            __syncthreads();
            R64_TS(1);
            {   // layer 1: h1 = tanh(W1 x + b1), both nets.  All loads, then all math, then all stores: the compiler must
                // assume shared-memory stores alias later loads, which would serialise the per-env dependency chains
                float* h1 = sm + L.h1 + (net1 * TM + mbase1) * R64_LDH + j1;
                float acc[MH];
#pragma unroll
                for (int i = 0; i < MH; i++) {
                    const float* x = sm + L.xs + (mbase1 + i) * O;
                    acc[i] = b1;
#pragma unroll
                    for (int o = 0; o < O; o++) acc[i] = fmaf(x[o], w1[o], acc[i]);
                }
#pragma unroll
                for (int i = 0; i < MH; i++) acc[i] = tanh_fast(acc[i]);
#pragma unroll
                for (int i = 0; i < MH; i++) h1[i * R64_LDH] = acc[i];
            }
            __syncthreads();
            R64_TS(2);
            {   // layer 2: h2 = tanh(W2 h1 + b2); two envs in flight per iteration (loads, math, shuffles, stores batched)
                const float* hin = sm + L.h1 + (net2 * TM + mbase2) * R64_LDH + kq * 16;
                float* hout = sm + L.h2 + (net2 * TM + mbase2) * R64_LDH + jbase + kq;
                const bool b0 = kq & 1, b1 = kq & 2;
#pragma unroll 1
                for (int m = 0; m < MH; m += 2) {
                    unsigned long long hp[2][8];
#pragma unroll
                    for (int e2 = 0; e2 < 2; e2++)
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const ulonglong2 h = *reinterpret_cast<const ulonglong2*>(hin + (m + e2) * R64_LDH + 4 * i);
                            hp[e2][2 * i] = h.x;
                            hp[e2][2 * i + 1] = h.y;
                        }
                    unsigned long long acc[2][4];
#pragma unroll
                    for (int e2 = 0; e2 < 2; e2++)
#pragma unroll
                        for (int r = 0; r < 4; r++) acc[e2][r] = 0ull;
#pragma unroll
                    for (int i = 0; i < 8; i++)
#pragma unroll
                        for (int r = 0; r < 4; r++)
#pragma unroll
                            for (int e2 = 0; e2 < 2; e2++) ffma2(acc[e2][r], hp[e2][i], w2[r * 8 + i]);
                    float keep[2];
#pragma unroll
                    for (int e2 = 0; e2 < 2; e2++) {
                        const float s0 = sum2(acc[e2][0]), s1 = sum2(acc[e2][1]), s2 = sum2(acc[e2][2]),
                                    s3 = sum2(acc[e2][3]);
                        float ka = b0 ? s1 : s0, kb = b0 ? s3 : s2;
                        ka += __shfl_xor_sync(0xffffffffu, b0 ? s0 : s1, 1);
                        kb += __shfl_xor_sync(0xffffffffu, b0 ? s2 : s3, 1);
                        keep[e2] = (b1 ? kb : ka) + __shfl_xor_sync(0xffffffffu, b1 ? ka : kb, 2);
                    }
#pragma unroll
                    for (int e2 = 0; e2 < 2; e2++) keep[e2] = tanh_fast(keep[e2] + b2);
#pragma unroll
                    for (int e2 = 0; e2 < 2; e2++) hout[(m + e2) * R64_LDH] = keep[e2];
                }
            }
            __syncthreads();
            R64_TS(3);
            {   // layer 3: K logits + 1 value per env; 8 lanes share one output
                const int part = t & 7, total = TM * (K + 1);
                for (int base = 0; base < total; base += 32) {
                    const int idx = base + (t >> 3);
                    const bool live = idx < total;
                    const int q = live ? idx / TM : 0, m = live ? idx % TM : 0;
                    const float* h = sm + L.h2 + ((q < K ? 0 : 1) * TM + m) * R64_LDH + part * 4;
                    const float* w = sm + L.w3 + q * 64 + part * 4;
                    const float4 ha = ld4(h), hb = ld4(h + 32), wa = ld4(w), wb = ld4(w + 32);
                    float acc = ha.x * wa.x;
                    acc = fmaf(ha.y, wa.y, acc); acc = fmaf(ha.z, wa.z, acc); acc = fmaf(ha.w, wa.w, acc);
                    acc = fmaf(hb.x, wb.x, acc); acc = fmaf(hb.y, wb.y, acc);
                    acc = fmaf(hb.z, wb.z, acc); acc = fmaf(hb.w, wb.w, acc);
                    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
                    if (live && part == 0) {
                        acc += sm[L.b3 + q];
                        if (q < K) sm[L.ls + m * L.ldo + q] = acc;
                        else sm[L.vs + m] = acc;
                    }
                }
            }
            __syncthreads();
            R64_TS(4);
            if (step == T) {  // bootstrap: next_value = critic(next_obs)  (calcAdvantage :280)
                if (on) {
                    a.next_value[e] = sm[L.vs + t];
                    a.next_done[e] = nd;
                }
                break;
            }
            if (on) {
                const size_t row = (size_t)step * n + e;
                const uint32_t env_id = a.cfg.env_id_base + (uint32_t)e;
                const uint64_t ctr = ctr0 + (uint64_t)step;
                const uint64_t mbits = masked ? (uint64_t)Env::action_mask(s) : ~0ull;
                int32_t acts[PPO_MAX_HEADS];
                const RowOut o = row_eval(sm + L.ls + t * L.ldo, heads, a.d.nvec, mbits, masked, nullptr, nullptr,
                                          a.cfg.seed, env_id, ctr, a.cfg.entropy_mode, acts);
                for (int h = 0; h < heads; h++) a.actions_buf[row * heads + h] = acts[h];
                a.logprobs_buf[row] = o.logp;
                a.values_buf[row] = sm[L.vs + t];
                if (a.mask_buf)
                    for (int j = 0; j < K; j++) a.mask_buf[row * K + j] = (uint8_t)((mbits >> j) & 1ull);
                // stepEnvs (:413-483): step, truncate, auto-reset, episode statistics
                StepOut so = Env::step_multi(s, acts);
                len += 1;
                ret = __fadd_rn(ret, so.reward);
                const bool done = so.terminated || (len == a.cfg.max_episode_steps);
                const int32_t f_len = done ? len : 0;
                const float f_ret = done ? ret : 0.0f;
                if (done) {
                    fin_cnt += 1.0; fin_len += (double)len; fin_ret += (double)ret;
                    float u[4];
                    reset_uniforms(a.cfg.seed, env_id, ctr, u);
                    Env::reset(s, u);
                    len = 0; ret = 0.0f;
                }
                a.rewards_buf[row] = so.reward;
                if (a.cfg.fin_len) { a.cfg.fin_len[row] = f_len; a.cfg.fin_ret[row] = f_ret; }
                nd = done ? 1.0f : 0.0f;
            }
            R64_TS(5);
        }
        if (on) {
#pragma unroll
            for (int k = 0; k < S; k++) a.state[(size_t)k * n + e] = s[k];
            a.ep_len[e] = len;
            a.ep_ret[e] = ret;
        }
        __syncthreads();
    }
    if (a.ep_stats) {
        __syncthreads();
        double* red = reinterpret_cast<double*>(sm);
        if (t < TM) { red[t] = fin_cnt; red[TM + t] = fin_len; red[2 * TM + t] = fin_ret; }
        __syncthreads();
        if (t < 3) {
            double acc = 0.0;
            for (int m = 0; m < TM; m++) acc += red[t * TM + m];
            if (acc != 0.0) atomicAdd(a.ep_stats + t, acc);
        }
    }
}

template <class Env, int TM, int H>
static int launch_rollout(const RollArgs& a, cudaStream_t st) {
    const TileLayout L = tile_layout(TM, H, Env::OBS, a.d.K);
    const size_t smem = (size_t)L.end * sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(rollout_kernel<Env, TM, H>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      200 * 1024));
        attr_done = true;
    }
    PPO_REQUIRE(smem <= 200 * 1024, "rollout: tile needs %zu B of shared memory", smem);
    const int ntiles = (a.cfg.n_envs + TM - 1) / TM;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < 4 * sms ? ntiles : 4 * sms;
    rollout_kernel<Env, TM, H><<<grid, MLP_THREADS, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

template <class Env, int TM>
static int launch_rollout64(const RollArgs& a, cudaStream_t st) {
    const R64Layout L = r64_layout(TM, Env::OBS, a.d.K);
    const size_t smem = (size_t)L.end * sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(rollout64_kernel<Env, TM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      100 * 1024));
        attr_done = true;
    }
    PPO_REQUIRE(smem <= 100 * 1024, "rollout: tile needs %zu B of shared memory", smem);
    const int ntiles = (a.cfg.n_envs + TM - 1) / TM;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < 2 * sms ? ntiles : 2 * sms;   // two resident CTAs per SM
    rollout64_kernel<Env, TM><<<grid, 256, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

static int g_rollout_impl = 1;   // 0 = generic kernel for every shape (ppo_debug_rollout_impl)
static int g_rollout_h256_min = 1024;  // 128-env MMA tiles: below ~8 tiles the 32-env FMA tiles spread over more SMs
static int g_rollout_tc_min = 8192;   // below ~8k envs a step is latency-bound and the FMA tiles (16 envs per CTA) win

template <class Env>
static int dispatch_rollout(const RollArgs& a, cudaStream_t st) {
    PPO_REQUIRE(a.d.O == Env::OBS, "rollout: obs_dim=%d but the env observes %d floats", a.d.O, Env::OBS);
    PPO_REQUIRE(a.d.heads >= Env::HEADS, "rollout: the env consumes %d sub-actions per step but the policy has %d heads",
                Env::HEADS, a.d.heads);
    // 128-env tensor-core tiles once there are enough envs to fill them (ppo_debug_rollout_impl: 2 = never, 3 = always)
    if (rollout_tc_supported(a.d) && (g_rollout_impl == 3 || (g_rollout_impl == 1 && a.cfg.n_envs >= g_rollout_tc_min)))
        return launch_rollout_tc<Env>(a, st);
    if (a.d.H == 64 && g_rollout_impl) {
        // small env counts: 16-env tiles so that two CTAs per SM overlap each other's serial env / sampling phase
        const int sms = sm_count() > 0 ? sm_count() : 148;
        if (a.cfg.n_envs <= 32 * sms) return launch_rollout64<Env, 16>(a, st);
        return launch_rollout64<Env, 32>(a, st);
    }
    // hidden = 256 in the 16-bit modes: layer 2 on the tensor core (rollout_h256.cu); PPO_MATH_FP32 keeps the FMA tiles
    if (rollout_h256_supported(a.d) && math_mode() != PPO_MATH_FP32 &&
        (g_rollout_impl == 3 || (g_rollout_impl == 1 && a.cfg.n_envs >= g_rollout_h256_min)))
        return launch_rollout_h256<Env>(a, math_mode() == PPO_MATH_BF16 ? 1 : 0, st);
    switch (a.d.H) {
        case 64: return launch_rollout<Env, 32, 64>(a, st);
        case 128: return launch_rollout<Env, 32, 128>(a, st);
        case 256: return launch_rollout<Env, 32, 256>(a, st);
    }
    set_error("hidden=%d unsupported", a.d.H);
    return PPO_ERR_UNSUPPORTED;
}

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_debug_rollout_timeline(long long* d_ts) {
    PPO_CUDA(cudaMemcpyToSymbol(g_roll_ts, &d_ts, sizeof(d_ts)));
    return PPO_OK;
}

extern "C" int ppo_debug_rollout_impl(int impl) {
    g_rollout_impl = impl;
    return PPO_OK;
}

extern "C" int ppo_rollout(const float* params, const ppo_dims* dims, const ppo_rollout_cfg* cfg, float* state_soa,
                           int32_t* ep_len, float* ep_ret, float* next_done, float* next_value, float* obs_buf,
                           int32_t* actions_buf, float* logprobs_buf, float* rewards_buf, float* dones_buf,
                           float* values_buf, float* adv_buf, float* ret_buf, uint8_t* mask_buf, double* ep_stats,
                           ppo_stream_t stream) {
    RollArgs a{};
    if (int rc = make_dims(dims, &a.d)) return rc;
    PPO_REQUIRE(cfg && params && state_soa && ep_len && ep_ret && next_done && next_value && obs_buf && actions_buf &&
                    logprobs_buf && rewards_buf && dones_buf && values_buf,
                "ppo_rollout: null buffer");
    PPO_REQUIRE(cfg->n_envs >= 0 && cfg->n_steps >= 0, "ppo_rollout: negative shape");
    if (cfg->n_envs == 0) return PPO_OK;
    a.params = params; a.cfg = *cfg; a.state = state_soa; a.ep_len = ep_len; a.ep_ret = ep_ret;
    a.next_done = next_done; a.next_value = next_value; a.obs_buf = obs_buf; a.actions_buf = actions_buf;
    a.logprobs_buf = logprobs_buf; a.rewards_buf = rewards_buf; a.dones_buf = dones_buf; a.values_buf = values_buf;
    a.mask_buf = mask_buf; a.ep_stats = ep_stats;
    cudaStream_t st = as_stream(stream);
    int rc;
    rc = env_dispatch(cfg->env_kind, [&](auto tag) { return dispatch_rollout<typename decltype(tag)::type>(a, st); }, -12345);
    if (rc == -12345) { set_error("ppo_rollout: unknown env kind %d", cfg->env_kind); return PPO_ERR_ARG; }
    if (rc) return rc;
    if (adv_buf && ret_buf && cfg->n_steps > 0)
        return ppo_gae(cfg->n_steps, cfg->n_envs, rewards_buf, values_buf, dones_buf, next_value, next_done,
                       cfg->gamma, cfg->gae_lambda, cfg->use_gae, adv_buf, ret_buf, stream);
    return PPO_OK;
}
