// update.cuh — argument block and workspace layout shared by the two K5 implementations (update.cu: fp32 FMA;
// update_tc.cu: tcgen05 tensor cores) and the cross-CTA reduction that follows either.
#pragma once
#include "common.cuh"

namespace ppo {

constexpr int UPD_MAX_GRID = 296;  // 2 CTAs x 148 SMs
constexpr int UPD_NDIAG = 8;       // pg, v, ent, kl, clip (+3 spare) partial sums per CTA

__host__ __device__ inline int64_t slab_stride(int P) { return ((int64_t)P + 31) & ~(int64_t)31; }

// Optional tail of the slab reduction: clip_grad_norm_ + AdamW (PPO/PPO_Discrete.cpp:640-641) applied by the LAST reduce
// CTA to finish (single-GPU path: no allreduce sits between the gradient and the optimiser, so reduce + norm + step are
// one launch instead of two).  Workspace tail: ticket (zeroed once, left zeroed) + one double per reduce CTA.
struct FuseAdam {
    float* params;
    float* exp_avg;
    float* exp_avg_sq;
    const float* d_lr;
    float beta1, beta2, eps, weight_decay;
    int64_t step_host;
    const int64_t* d_step;
    float max_grad_norm;
    float* norm_out;
    float* diag_hist;   // 8 floats of this optimiser step
    unsigned* err;      // optional device flag set (and the step skipped) when the intra-GPU ticket spin times out
    long long spin_cycles;   // 0 = default
};
// Slab reduction + clip + AdamW executed by the update kernel itself behind a grid-wide barrier (update_tc.cu:
// fused_reduce_apply) instead of a second launch.  on != 0 only when every CTA of the launch is resident and there are at
// least as many CTAs as 64-column blocks.  ticket: 16 uints (slots 2..4 used here), sq_part: one double per block.
struct TailArgs {
    int on, P;
    int64_t stride;
    float ent_coef, vf_coef;
    float* grads_out;
    float* diag_out;
    FuseAdam f;
    unsigned* ticket;
    double* sq_part;
};

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
    const float4* rec;   // optional packed rows (ppo_pack_rows): 64 B per row = one aligned DRAM burst instead of six
                         // scattered 4-byte gathers; single-head, unmasked, obs_dim <= 4 only
    PermFn pf;       // perm == nullptr && pf.inline_on: rows are perm_index(mb_start + i) evaluated in the kernel
    int64_t mb_start;
    int mb_size;
    float inv_n;
    const float* adv_stats;  // {mean, std} or nullptr
    ppo_hyper hp;
    float* slabs;    // [grid][slab_stride(P)]
    double* dpart;   // [grid][UPD_NDIAG]
    TailArgs tail;
    int pdl;         // launch the tensor-core kernel with programmatic stream serialization (common.cuh: launch_pdl)
};

// Data-parallel variant of the fused tail: the allreduce happens INSIDE the reduction kernel over peer memory (NVLink
// P2P stores into the other ranks' exchange buffers, mapped with CUDA IPC), so gradient exchange + clip + AdamW is one
// launch per rank instead of reduce_kernel -> ncclAllReduce -> clip_adamw_kernel.  PUSH model, LL words: exchange buffer of
// a rank (one cudaMalloc) = two slots (optimiser step parity) of uint64 [world][peer_px(P)].  CTA b of rank r STORES its 64
// reduced columns as {step number, fp32 bits} words into row r of EVERY rank's buffer (its own included) and then polls the
// words of its columns in its OWN buffer until every row carries this step's tag, summing them in rank order (=>
// bit-identical totals on every rank).  One one-way NVLink latency per optimiser step, no flags, no system fence.
// Slot reuse two steps later is safe: a peer publishes step s+1 only after its step-s reads.
constexpr int PEER_MAX = 8, PEER_NB = 256;
__host__ __device__ inline int64_t peer_px(int P) { return (((int64_t)P + 63) & ~(int64_t)63) + 64; }   // + diagnostics
__host__ __device__ inline int64_t peer_slot_bytes(int P, int world) { return (int64_t)world * peer_px(P) * 8; }
struct PeerArgs {
    int world, rank;
    char* xchg[PEER_MAX];   // every rank's exchange buffer as mapped in THIS process (xchg[rank] = the local one)
    unsigned* err;          // local device flag: set when a peer did not arrive within the spin limit (sticky: once set,
                            // every later optimiser step is skipped so that ranks cannot step on partial gradients)
    long long spin_cycles;  // spin limit in SM cycles (ppo_set_spin_timeout_ms)
};
constexpr int RED_COLS = 64, RED_PARTS = 4;
__host__ __device__ inline int64_t upd_tail_bytes(int P) { return 64 + (int64_t)((P + RED_COLS - 1) / RED_COLS) * 8; }
// ppo_update_minibatch with the optimiser step fused behind it (trainer.cu)
int update_minibatch_fused(const float* params, const ppo_dims* dims, const float* obs, const uint8_t* mask,
                           const int32_t* actions, const float* old_logprob, const float* adv, const float* ret,
                           const float* old_val, const int32_t* perm, int64_t mb_start, int mb_size, float inv_n,
                           const float* adv_stats, const ppo_hyper* h_hyper, void* workspace, float* grads_out,
                           float* diag_out, const FuseAdam* fuse, cudaStream_t st, const PeerArgs* peer = nullptr,
                           const PermFn* pf = nullptr, const float* rec = nullptr);

// update_tc.cu: returns PPO_ERR_UNSUPPORTED (without touching the error string) when the shape is outside the
// tensor-core kernel's envelope (hidden != 64, obs_dim > 7, sumK > 32), PPO_OK after launching, else an error.
// fuse != nullptr: the kernel may run the reduction + optimiser tail itself; *tail_done says whether it did
int launch_update_tc(UpdArgs& a, int fmt, int* grid_out, cudaStream_t st, const FuseAdam* fuse = nullptr,
                     float* grads_out = nullptr, float* diag_out = nullptr, bool* tail_done = nullptr);

bool update_tc_supported(const Dims& d);     // update_tc.cu: hidden = 64 envelope
bool update_h256_supported(const Dims& d);   // update_h256.cu: hidden = 256 envelope
// update_h256.cu: forward / loss / backward kernel + dW2 GEMM kernel for hidden = 256; PPO_ERR_UNSUPPORTED outside the envelope
int launch_update_h256(UpdArgs& a, int fmt, int* grid_out, cudaStream_t st);
// grows the 16-bit activation scratch of the hidden = 256 path to `mb_rows` minibatch rows (call outside stream capture)
int h256_reserve(int64_t mb_rows);

}  // namespace ppo
