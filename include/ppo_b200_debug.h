/* ppo_b200_debug.h — diagnostics of libppo_b200.so: tensor-core building-block probes, A/B switches between kernel
 * variants and in-kernel timelines.  NOT part of the drop-in boundary (include/ppo_b200.h): nothing here has a reference
 * counterpart; tests/, tools/ and bench.py use these to measure and to pin kernel variants against each other. */
#ifndef PPO_B200_DEBUG_H
#define PPO_B200_DEBUG_H
#include "ppo_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Known-answer probe of the tcgen05 (16-bit operands, fp32 TMEM accumulator) building blocks the fused kernels
 * use (csrc/umma.cuh).  One 128-thread CTA, one GEMM with operands rounded to fp16 (fmt 0) or bf16 (fmt 1):
 * mode 0: D[128,64] = A[128,64] B[64,64]^T; 1: D = A B[64,64]; 2: D[64,64] = A^T B[128,64]; 3: D[64,16] = A^T B[128,16].
 * D receives the raw tensor-memory image [128 lanes][64 columns] (M = 64 results sit in lanes
 * (m % 16) + 32 (m / 16)).  No reference counterpart. */
int ppo_debug_umma_probe(int mode, int fmt, const float* A, const float* B, float* D, ppo_stream_t stream);
/* Same GEMM issued `reps` times back to back by one thread; d_cycles[0] (device int64) = SM clocks from the first issue
 * to the completion of the last (how fast the tensor pipe retires each operand layout). */
/* Rollout kernel choice: 1 (default) = the dispatcher's pick (hidden = 64: register-resident FMA tiles below 8 192 envs,
 * tcgen05 tiles above; hidden = 256 in the 16-bit modes: tcgen05 tiles from 1 024 envs), 0 = the generic tiled FMA kernel
 * for every shape, 2 = never the tensor-core tiles, 3 = always the tensor-core tiles where the shape allows them. */
int ppo_debug_rollout_impl(int impl);
/* 1 (default): trainers whose shape allows it (single head, unmasked, obs_dim <= 4) repack the rollout rows with
 * ppo_pack_rows once per iteration and the update kernel gathers records; 0: gather from the six row arrays. */
int ppo_debug_pack_rows(int on);
/* 1: on one GPU, when the update kernel's grid covers every 64-column block of the parameter vector, it runs the slab
 * reduction + clip_grad_norm_ + AdamW itself behind a grid-wide barrier (one launch per optimiser step); 0 (default): the
 * separate reduce_apply_kernel launch, which measured 1-2 % faster. */
int ppo_debug_fuse_tail(int on);
/* Bit 0 (default on): on ONE GPU the update kernel and the reduction / optimiser kernel are launched with programmatic
 * stream serialization (each calls griddepcontrol.wait before reading its predecessor's output), so a launch overlaps the
 * tail of the kernel in front of it.  Bit 1 (default off): the same next to the peer-memory gradient exchange (N > 1).
 * 0: ordinary stream order. */
int ppo_debug_pdl(int on);
/* d_ts = device buffer of 8 int64 (NULL switches it off): CTA 0 / thread 0 of the hidden = 64 rollout kernel stamps the
 * SM clock at the six phase edges of rollout step 2 (tools/rollout_bench.py prints the deltas). */
int ppo_debug_rollout_timeline(long long* d_ts);
/* In-kernel timeline of the tensor-core update kernel: d_ts = device buffer of 256 int64 (NULL switches it off);
 * later launches stamp the SM clock of CTA 0, thread `who` at fixed points (tools/tc_timeline.py prints them). */
int ppo_debug_tc_timeline(long long* d_ts, int who);
int ppo_debug_umma_time(int mode, int fmt, const float* A, const float* B, float* D, int reps, long long* d_cycles,
                        ppo_stream_t stream);
/* tcgen05.mma throughput / concurrency probe (tools/umma_conc.py): `nwarps` warps each issue reps x 8 MMAs of one shape
 * (0 M128N64 K/K, 1 M64N64 MN/MN, 2 M64N8 MN/narrow, 3 M128N8 MN/narrow, 4 M128N64 K/MN) from their own tiles into their
 * own tensor-memory columns; d_cycles[0] = SM cycles from the first issue to the last completion. */
int ppo_debug_umma_conc(int shape, int nwarps, int reps, long long* d_cycles, ppo_stream_t stream);
/* 10 * (tile streams per CTA) + (K as a template parameter, 0 = generic head path) of the update_tc_kernel variant that serves
 * a shape, -1 outside the tensor-core envelope (host logic only, no device needed). */
int ppo_debug_update_tc_variant(const ppo_dims* dims, int masked);

/* 1 (default): ppo_categorical runs the warp-cooperative kernel (one (row, head) per lane) whenever every head has at most
 * 8 logits; 0: the row-per-thread kernel (IEEE expf / logf / division per logit) for every shape. */
int ppo_debug_categorical_impl(int impl);
/* 1 (default): ppo_adv_stats_all_epochs reads the advantages in one coalesced pass per epoch through the inverse
 * permutation once there are at least 2^25 rows (beyond what stays L2-resident); 0: the permuted-gather kernel for every
 * size; 2: the coalesced kernel for every size (at most 8 minibatches per epoch). */
int ppo_debug_adv_impl(int impl);
/* In-kernel timeline of update_h256_kernel: d_ts = device buffer of 64 int64 (NULL switches it off); later launches stamp the
 * SM clock of CTA 0 / thread 0 at 14 phase edges of its first 4 tiles, 16 slots per tile (tools/h256_timeline.py). */
int ppo_debug_h256_timeline(long long* d_ts);


#ifdef __cplusplus
}
#endif
#endif /* PPO_B200_DEBUG_H */
