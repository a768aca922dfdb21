// gae.cu — K4: GAE / discounted returns as a reverse-time scan over [T, N] (time-major, N contiguous).
// Replaces PPO_Discrete::calcAdvantage's T sequential iterations of ~10 tiny LibTorch ops
// (PPO/PPO_Discrete.cpp:289-305 GAE, :309-329 plain returns).
//
// HBM-bound: 20 B per (t, env) element (read r, V, done; write adv, ret).  One thread owns VEC
// adjacent env columns; the carry (lastgaelam, V_{t+1}, done_{t+1}) stays in registers, so each
// element is read once and written once.  Loads are issued UNROLL time steps ahead of the
// dependent recurrence so that UNROLL*3 independent 16 B loads are in flight per thread.
#include "common.cuh"

namespace ppo {

template <int V> struct F;
template <> struct F<1> { using t = float; };
template <> struct F<2> { using t = float2; };
template <> struct F<4> { using t = float4; };

template <int V> __device__ inline void ld_stream(const float* p, float (&r)[V]) {
    if constexpr (V == 4) {
        float4 v;
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
        r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = v.w;
    } else if constexpr (V == 2) {
        float2 v;
        asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
        r[0] = v.x; r[1] = v.y;
    } else {
        r[0] = __ldg(p);
    }
}
template <int V> __device__ inline void st_stream(float* p, const float (&r)[V]) {
    if constexpr (V == 4) {
        asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(r[0]), "f"(r[1]),
                     "f"(r[2]), "f"(r[3]) : "memory");
    } else if constexpr (V == 2) {
        asm volatile("st.global.L1::no_allocate.v2.f32 [%0], {%1,%2};" ::"l"(p), "f"(r[0]), "f"(r[1]) : "memory");
    } else {
        *p = r[0];
    }
}

template <int V, int UNROLL>
__global__ void __launch_bounds__(128) gae_kernel(int T, int n, const float* __restrict__ rewards,
                                                  const float* __restrict__ values,
                                                  const float* __restrict__ dones,
                                                  const float* __restrict__ next_value,
                                                  const float* __restrict__ next_done, float gamma, float lam,
                                                  int use_gae, float* __restrict__ adv_out,
                                                  float* __restrict__ ret_out) {
    const int64_t c0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V;
    if (c0 >= n) return;
    const float gl = __fmul_rn(gamma, lam);
    float nv[V], nnt[V], last[V];
    {
        float nd[V];
        ld_stream<V>(next_value + c0, nv);
        ld_stream<V>(next_done + c0, nd);
#pragma unroll
        for (int v = 0; v < V; v++) { nnt[v] = __fsub_rn(1.0f, nd[v]); last[v] = 0.0f; }
    }
    // non-GAE branch (:309-329): R_{T-1} bootstraps from next_value, afterwards from R_{t+1}
    bool first = true;
    for (int t1 = T; t1 > 0; t1 -= UNROLL) {
        const int cnt = t1 < UNROLL ? t1 : UNROLL;
        float r[UNROLL][V], val[UNROLL][V], d[UNROLL][V];
#pragma unroll
        for (int k = 0; k < UNROLL; k++) {
            if (k < cnt) {
                const size_t o = (size_t)(t1 - 1 - k) * n + c0;
                ld_stream<V>(rewards + o, r[k]);
                ld_stream<V>(values + o, val[k]);
                ld_stream<V>(dones + o, d[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < UNROLL; k++) {
            if (k < cnt) {
                const size_t o = (size_t)(t1 - 1 - k) * n + c0;
                float a[V], rt[V];
#pragma unroll
                for (int v = 0; v < V; v++) {
                    if (use_gae) {
                        // delta = r + gamma*nv*nnt - V ; last = delta + gamma*lambda*nnt*last
                        const float delta = __fsub_rn(
                            __fadd_rn(r[k][v], __fmul_rn(__fmul_rn(gamma, nv[v]), nnt[v])), val[k][v]);
                        last[v] = __fadd_rn(delta, __fmul_rn(__fmul_rn(gl, nnt[v]), last[v]));
                        a[v] = last[v];
                        rt[v] = __fadd_rn(last[v], val[k][v]);
                    } else {
                        const float nr = first ? nv[v] : last[v];
                        last[v] = __fadd_rn(r[k][v], __fmul_rn(__fmul_rn(gamma, nnt[v]), nr));
                        rt[v] = last[v];
                        a[v] = __fsub_rn(last[v], val[k][v]);
                    }
                    nv[v] = val[k][v];
                    nnt[v] = __fsub_rn(1.0f, d[k][v]);
                }
                first = false;
                st_stream<V>(adv_out + o, a);
                st_stream<V>(ret_out + o, rt);
            }
        }
    }
}

template <int V>
static int launch_gae(int T, int n, const float* rewards, const float* values, const float* dones,
                      const float* next_value, const float* next_done, float gamma, float lam, int use_gae,
                      float* adv, float* ret, cudaStream_t st) {
    const int threads = 128;
    const int64_t items = ((int64_t)n + V - 1) / V;
    const int blocks = (int)((items + threads - 1) / threads);
    gae_kernel<V, 8><<<blocks, threads, 0, st>>>(T, n, rewards, values, dones, next_value, next_done, gamma, lam,
                                                use_gae, adv, ret);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_gae(int T, int n, const float* rewards, const float* values, const float* dones,
                       const float* next_value, const float* next_done, float gamma, float gae_lambda, int use_gae,
                       float* adv_out, float* ret_out, ppo_stream_t stream) {
    PPO_REQUIRE(T >= 0 && n >= 0, "ppo_gae: negative shape T=%d n=%d", T, n);
    if (T == 0 || n == 0) return PPO_OK;
    PPO_REQUIRE(rewards && values && dones && next_value && next_done && adv_out && ret_out,
                "ppo_gae: null buffer");
    auto al = [](const void* p, int b) { return (reinterpret_cast<uintptr_t>(p) & (uintptr_t)(b - 1)) == 0; };
    auto all_al = [&](int b) {
        return al(rewards, b) && al(values, b) && al(dones, b) && al(next_value, b) && al(next_done, b) &&
               al(adv_out, b) && al(ret_out, b);
    };
    // Vector width: widest that keeps every row 16/8 B aligned, but do not starve the 148 SMs of
    // threads when N is small (config B: 4096 columns -> scalar columns, 32 CTAs of 128 threads).
    const int sms = sm_count() > 0 ? sm_count() : 148;
    if (n % 4 == 0 && all_al(16) && (int64_t)n / 4 >= (int64_t)sms * 128 * 4)
        return launch_gae<4>(T, n, rewards, values, dones, next_value, next_done, gamma, gae_lambda, use_gae,
                             adv_out, ret_out, as_stream(stream));
    if (n % 2 == 0 && all_al(8) && (int64_t)n / 2 >= (int64_t)sms * 128 * 2)
        return launch_gae<2>(T, n, rewards, values, dones, next_value, next_done, gamma, gae_lambda, use_gae,
                             adv_out, ret_out, as_stream(stream));
    return launch_gae<1>(T, n, rewards, values, dones, next_value, next_done, gamma, gae_lambda, use_gae, adv_out,
                         ret_out, as_stream(stream));
}
