// update.cuh — argument block and workspace layout shared by the two K5 implementations (update.cu: fp32 FMA;
// update_tc.cu: tcgen05 tensor cores) and the cross-CTA reduction that follows either.
#pragma once
#include "common.cuh"

namespace ppo {

constexpr int UPD_MAX_GRID = 296;  // 2 CTAs x 148 SMs
constexpr int UPD_NDIAG = 8;       // pg, v, ent, kl, clip (+3 spare) partial sums per CTA

__host__ __device__ inline int64_t slab_stride(int P) { return ((int64_t)P + 31) & ~(int64_t)31; }

struct UpdArgs {
    const float* params;
    Dims d;
    const float* obs;
    const uint8_t* mask;
    const int32_t* actions;
    const float* old_logprob;
    const float* adv;
    const float* ret;
    const float* old_val;
    const int32_t* perm;
    int64_t mb_start;
    int mb_size;
    float inv_n;
    const float* adv_stats;  // {mean, std} or nullptr
    ppo_hyper hp;
    float* slabs;    // [grid][slab_stride(P)]
    double* dpart;   // [grid][UPD_NDIAG]
};

// update_tc.cu: returns PPO_ERR_UNSUPPORTED (without touching the error string) when the shape is outside the
// tensor-core kernel's envelope (hidden != 64, obs_dim > 7, sumK > 32), PPO_OK after launching, else an error.
int launch_update_tc(const UpdArgs& a, int fmt, int* grid_out, cudaStream_t st);

}  // namespace ppo
