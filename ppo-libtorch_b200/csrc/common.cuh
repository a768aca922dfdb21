// common.cuh — shared helpers for the sm_100a PPO kernels: error plumbing for the C ABI,
// Philox4x32-10, the flat parameter-slab layout (reference order, PPO/Agent.cpp:41-66).
#pragma once
#include <utility>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/ppo_b200.h"
#include "../../include/ppo_b200_debug.h"

namespace ppo {

// ---------------------------------------------------------------- error plumbing
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define PPO_CUDA(call)                                                        \
    do {                                                                      \
        cudaError_t _e = (call);                                              \
        if (_e != cudaSuccess) return ::ppo::cuda_fail(_e, #call, __FILE__, __LINE__); \
    } while (0)
#define PPO_LAUNCH_CHECK() PPO_CUDA(cudaGetLastError())
#define PPO_REQUIRE(cond, ...)                  \
    do {                                        \
        if (!(cond)) {                          \
            ::ppo::set_error(__VA_ARGS__);      \
            return PPO_ERR_ARG;                 \
        }                                       \
    } while (0)

int math_mode();  // PPO_MATH_* selected with ppo_set_math_mode
int sm_count();  // cached multiprocessor count of the current device (0 if no device)

// ---------------------------------------------------------------- dims helpers
struct Dims {
    int O, H, heads, K;  // K = total logits
    int nvec[PPO_MAX_HEADS];
};
int make_dims(const ppo_dims* d, Dims* out);  // validates; PPO_OK or PPO_ERR_ARG

// Offsets (in floats) into the flat slab; critic first (PPO/Agent.cpp:41-47), then actor (:52-66).
struct NetOff {
    int W1, b1, W2, b2, W3, b3;
};
__host__ __device__ inline NetOff critic_off(int O, int H) {
    NetOff n;
    n.W1 = 0;
    n.b1 = n.W1 + H * O;
    n.W2 = n.b1 + H;
    n.b2 = n.W2 + H * H;
    n.W3 = n.b2 + H;
    n.b3 = n.W3 + H;
    return n;
}
__host__ __device__ inline int critic_size(int O, int H) { return H * O + H + H * H + H + H + 1; }
__host__ __device__ inline NetOff actor_off(int O, int H, int K) {
    const int base = critic_size(O, H);
    NetOff n;
    n.W1 = base;
    n.b1 = n.W1 + H * O;
    n.W2 = n.b1 + H;
    n.b2 = n.W2 + H * H;
    n.W3 = n.b2 + H;
    n.b3 = n.W3 + K * H;
    return n;
}
__host__ __device__ inline int num_params(int O, int H, int K) {
    return critic_size(O, H) + H * O + H + H * H + H + K * H + K;
}

// ---------------------------------------------------------------- Philox4x32-10
// Counter-based RNG (Salmon et al. 2011).  key = 64-bit seed, counter = (c0,c1,c2,c3).
struct Philox {
    static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    static constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    __host__ __device__ static inline void round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#ifdef __CUDA_ARCH__
        const uint32_t hi0 = __umulhi(M0, c[0]), hi1 = __umulhi(M1, c[2]);
#else
        const uint32_t hi0 = (uint32_t)(((uint64_t)M0 * c[0]) >> 32);
        const uint32_t hi1 = (uint32_t)(((uint64_t)M1 * c[2]) >> 32);
#endif
        const uint32_t lo0 = M0 * c[0], lo1 = M1 * c[2];
        const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    }
    // 4 x 32 random bits for (seed, a, b, c, d)
    __host__ __device__ static inline void gen(uint64_t seed, uint32_t a, uint32_t b, uint32_t c_, uint32_t d,
                                               uint32_t (&out)[4]) {
        uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
        out[0] = a; out[1] = b; out[2] = c_; out[3] = d;
#pragma unroll
        for (int r = 0; r < 10; r++) {
            round(out, k0, k1);
            k0 += W0; k1 += W1;
        }
    }
    // canonical uniform in [0,1): top 24 bits
    __host__ __device__ static inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
};
// stream ids so that resets, sampling and permutations never share a counter
enum { RNG_STREAM_RESET = 1, RNG_STREAM_SAMPLE = 2 };

// Permutation of [0, batch) (replaces torch::randperm, PPO/PPO_Discrete.cpp:569): a 4-round Feistel network on the
// smallest even-width bit domain that covers `batch`, with cycle walking back into range - a bijection for every batch
// size, evaluable per index, so kernels can gather through it without materialising the index buffer.
struct PermFn {
    int inline_on;          // 1: evaluate perm_index() in the consuming kernel (no perm buffer)
    int half_bits;
    int64_t batch;
    uint64_t seed, epoch;
    const uint64_t* d_epoch;   // optional device scalar added to `epoch` (CUDA-graph replay)
};
// Domain of the bijection = 2^bits with bits = ceil(log2(batch)) (>= 2): an UNBALANCED Feistel network (left half
// bits / 2, right half the rest; the halves swap size every round and are back after four), so a power-of-two batch is
// permuted without any cycle walking and any other batch rejects < 50 % of the draws.  (A balanced network needs an even
// bit count: batch = 2^19 then walked a 2^20 domain, and a warp iterates until its slowest lane is inside - ~6 network
// evaluations per index instead of 1.)  The field keeps its old name.
__host__ inline int perm_half_bits(int64_t batch) {
    int bits = 2;
    while (bits < 32 && (1ll << bits) < batch) bits++;
    return bits;
}
__device__ inline void perm_keys(uint64_t seed, uint64_t epoch, uint32_t (&k)[4]) {
    Philox::gen(seed, 0x7065726Du /* 'perm' */, 3u, (uint32_t)epoch, (uint32_t)(epoch >> 32), k);
}
__device__ inline uint32_t feistel(uint32_t x, int bits, const uint32_t (&k)[4]) {
    int lb = bits >> 1, rb = bits - lb;   // l: the high lb bits, r: the low rb bits
    uint32_t l = x >> rb, r = x & ((1u << rb) - 1u);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        uint32_t f = (r ^ k[i]) * 0x9E3779B1u;
        f ^= f >> 15; f *= 0x85EBCA77u; f ^= f >> 13;
        const uint32_t nl = r, nr = l ^ (f & ((1u << lb) - 1u));
        l = nl; r = nr;
        const int t = lb; lb = rb; rb = t;
    }
    return (l << rb) | r;
}
// inverse of feistel(): the rounds backwards (the halves' widths are back to (bits / 2, rest) after four rounds)
__device__ inline uint32_t feistel_inv(uint32_t y, int bits, const uint32_t (&k)[4]) {
    int lb = bits >> 1, rb = bits - lb;
    uint32_t l = y >> rb, r = y & ((1u << rb) - 1u);
#pragma unroll
    for (int i = 3; i >= 0; i--) {
        // this round produced (l, r) = (r_prev, l_prev ^ f(r_prev)) from widths (lb_prev, rb_prev) = (rb, lb)
        const int plb = rb, prb = lb;
        const uint32_t pr = l;
        uint32_t f = (pr ^ k[i]) * 0x9E3779B1u;
        f ^= f >> 15; f *= 0x85EBCA77u; f ^= f >> 13;
        const uint32_t pl = r ^ (f & ((1u << plb) - 1u));
        l = pl; r = pr;
        lb = plb; rb = prb;
    }
    return (l << rb) | r;
}
// position of row j in the permuted order: perm_index(perm_position(j)) == j (cycle walking inverts the same way)
__device__ inline int64_t perm_position(int64_t j, int64_t batch, int half_bits, const uint32_t (&k)[4]) {
    uint32_t x = (uint32_t)j;
    do { x = feistel_inv(x, half_bits, k); } while ((int64_t)x >= batch);
    return (int64_t)x;
}
__device__ inline int64_t perm_index(int64_t i, int64_t batch, int half_bits, const uint32_t (&k)[4]) {
    uint32_t x = (uint32_t)i;
    do { x = feistel(x, half_bits, k); } while ((int64_t)x >= batch);
    return (int64_t)x;
}

static inline cudaStream_t as_stream(ppo_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- programmatic dependent launch (the per-minibatch kernel chain update -> reduce/AdamW -> update -> ...) --------------
// A kernel launched with launch_pdl may be scheduled while its predecessor on the stream is still running; it must execute
// pdl_wait() before touching anything the predecessor writes (the wait returns once the predecessor grid has completed
// and its memory is visible).  pdl_trigger() in the predecessor lets the successor's launch proceed early.  Both are no-ops
// for a kernel launched the ordinary way.  Captured into CUDA graphs as programmatic edges.
extern int g_pdl;   // ppo_debug_pdl
extern long long g_spin_cycles;   // ppo_set_spin_timeout_ms
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <class... KArgs, class... Args>
static inline cudaError_t launch_pdl(bool allow, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = allow ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace ppo
