// trainer.cu — host side of one PPO iteration: the body of PPO_Discrete::train /
// PPO_MultiDiscrete::train (PPO/PPO_Discrete.cpp:511-673 minus printing and checkpointing) as a
// stream of kernel launches over one device arena, with no host round trip inside an iteration.
//
//   iterate():  rollout (1 launch) -> GAE (1) -> advantage statistics of all E x MB minibatches (1; the minibatch
//   permutation is a keyed bijection evaluated inside the kernels, never materialised) -> row packing (1) ->
//   per minibatch { fused fwd+loss+bwd (1; 6 at hidden = 256), slab reduction + [NVLink exchange] + clip + AdamW (1) }
//   -> explained variance (1), episode tail (1) -> one D2H of the diagnostics.
//   The split-phase API (rollout / begin_epoch / minibatch_grads / apply / finish) and NCCL data parallelism keep the
//   per-epoch permutation + statistics launches and the separate reduce / allreduce / clip + AdamW launches.
//
// With use_cuda_graph the whole iteration is captured once and replayed; the values that change
// between iterations (learning rate, optimiser step, RNG counters) live in device scalars.
#include <cstddef>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "common.cuh"
#include "update.cuh"

extern "C" int ppo_adv_stats_all_epochs(const float* adv, int64_t batch, int mb_size, int n_mb, int n_epochs, uint64_t seed,
                                        uint64_t epoch0, const uint64_t* d_epoch, float* stats_out, double* partial_out,
                                        void* workspace, ppo_stream_t stream);

namespace ppo {
int clip_adamw_launch(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_params, float lr,
                      const float* d_lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                      const int64_t* d_step, float max_grad_norm, float grad_scale, float* norm_out,
                      const float* diag_src, float* diag_dst, cudaStream_t st);
}
struct ppo_comm;
extern "C" int ppo_comm_allreduce_f32(ppo_comm* c, float* buf, int64_t count, ppo_stream_t stream);
extern "C" int ppo_comm_allreduce_f64(ppo_comm* c, double* buf, int64_t count, ppo_stream_t stream);
extern "C" int ppo_comm_world_size(const ppo_comm* c);

using namespace ppo;

namespace {
struct Scalars {           // device scalars a captured graph reads
    uint64_t rollout_counter;  // Philox counter base of this rollout
    uint64_t epoch_counter;    // permutation epoch base of this update
    int64_t step_base;         // optimiser steps taken before this iteration
    float lr;
    float pad;
};
struct Readback {          // one D2H per iteration
    float ev, grad_norm;
    double ep_stats[3];
    unsigned peer_err, pad;
    int32_t tail_count, pad2;            // newest finished episodes of the iteration, oldest first (ppo_episode_tail)
    int32_t tail_len[PPO_EPISODE_TAIL];
    float tail_ret[PPO_EPISODE_TAIL];
};
}  // namespace

struct ppo_trainer {
    ppo_trainer_cfg cfg;
    Dims d;
    int P = 0;
    int64_t B = 0;
    int mb_size = 0, n_mb = 0, n_opt = 0;  // n_opt = optimiser steps per iteration
    char* arena = nullptr;
    size_t arena_bytes = 0;
    std::map<std::string, std::pair<void*, int64_t>> buf;
    // device
    float *params, *grads, *exp_avg, *exp_avg_sq;
    float *state, *ep_ret, *next_done, *next_value;
    int32_t* ep_len;
    float *obs, *logprobs, *rewards, *dones, *values, *adv, *ret;
    int32_t *actions, *perm;
    uint8_t* masks;
    float *adv_stats, *diag_hist, *ev_out, *norm_out;
    double *adv_partial, *ep_stats;
    void *upd_ws, *ev_ws, *adv_ws;
    int32_t* fin_len = nullptr;   // episode log [T,N] + its tail {count, len[100], ret[100]}
    float* fin_ret = nullptr;
    int32_t* tail = nullptr;
    bool use_rec = false;
    float* rec = nullptr;   // packed rollout rows [B][16] (ppo_pack_rows) or nullptr when the shape has no packed form
    Scalars* d_scal;
    // host
    float* h_diag = nullptr;    // pinned [n_opt, 8]
    Readback* h_rb = nullptr;   // pinned
    int64_t step = 0;
    uint64_t rollout_counter = 0, epoch_counter = 0;
    ppo_comm* comm = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    cudaStream_t graph_stream = nullptr;
    int opt_in_iter = 0;  // optimiser steps issued in the current iteration (split-phase API)
    // peer-memory gradient exchange (update.cuh: PeerArgs): own exchange buffer + the peers' IPC mappings
    char* xchg = nullptr;
    unsigned* err_flag = nullptr;   // device word: an in-kernel wait (grid ticket, peer exchange) ran out; sticky
    ppo::PeerArgs peer{};
    bool peer_ready = false;
    // optional per-phase device timing (cudaEvent pairs on the launch stream; eager mode only)
    bool profile = false;
    std::vector<cudaEvent_t> ev_pool;
    std::vector<int> ev_phase;  // phase of pair i (events 2i, 2i+1)
    size_t ev_used = 0;
};

enum { PH_ROLLOUT = 0, PH_GAE, PH_EPOCH, PH_UPDATE, PH_APPLY, PH_TAIL, PH_COUNT };

namespace {
struct PhaseTimer {  // records an event pair around a phase when profiling is on
    ppo_trainer* t;
    cudaStream_t st;
    bool on;
    PhaseTimer(ppo_trainer* t_, int phase, cudaStream_t st_) : t(t_), st(st_), on(t_->profile) {
        if (!on) return;
        if (t->ev_used + 2 > t->ev_pool.size()) {
            cudaEvent_t a, b;
            cudaEventCreate(&a);
            cudaEventCreate(&b);
            t->ev_pool.push_back(a);
            t->ev_pool.push_back(b);
        }
        t->ev_phase.resize(t->ev_used / 2 + 1);
        t->ev_phase[t->ev_used / 2] = phase;
        cudaEventRecord(t->ev_pool[t->ev_used], st);
    }
    ~PhaseTimer() {
        if (!on) return;
        cudaEventRecord(t->ev_pool[t->ev_used + 1], st);
        t->ev_used += 2;
    }
};
}  // namespace

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" void ppo_trainer_destroy(ppo_trainer* t) {
    if (!t) return;
    if (t->graph_exec) cudaGraphExecDestroy(t->graph_exec);
    for (cudaEvent_t e : t->ev_pool) cudaEventDestroy(e);
    for (int r = 0; r < ppo::PEER_MAX; r++)
        if (t->peer_ready && r != t->peer.rank && t->peer.xchg[r]) cudaIpcCloseMemHandle(t->peer.xchg[r]);
    if (t->xchg) cudaFree(t->xchg);
    if (t->arena) cudaFree(t->arena);
    if (t->h_diag) cudaFreeHost(t->h_diag);
    if (t->h_rb) cudaFreeHost(t->h_rb);
    delete t;
}

extern "C" ppo_trainer* ppo_trainer_create(const ppo_trainer_cfg* c) {
    if (!c) { set_error("ppo_trainer_create: null cfg"); return nullptr; }
    if (ppo_device_count() <= 0) {
        set_error("ppo_trainer_create: no CUDA device (there is no CPU fallback)");
        return nullptr;
    }
    ppo_trainer* t = new ppo_trainer();
    t->cfg = *c;
    if (make_dims(&c->dims, &t->d)) { delete t; return nullptr; }
    // PPO_ENV_HOST: the environments live on the host (f4 bridge): no device env state is stepped, the caller fills the
    // rollout rows and runs ppo_trainer_update
    const int S = c->env_kind == PPO_ENV_HOST ? t->d.O : ppo_env_state_dim(c->env_kind);
    if (S < 0 || S != t->d.O) {
        set_error("ppo_trainer_create: env kind %d observes %d floats but obs_dim=%d", c->env_kind, S, t->d.O);
        delete t;
        return nullptr;
    }
    if (c->num_envs < 1 || c->num_steps < 1 || c->num_minibatches < 1 || c->update_epochs < 0) {
        set_error("ppo_trainer_create: bad sizes");
        delete t;
        return nullptr;
    }
    t->P = num_params(t->d.O, t->d.H, t->d.K);
    t->B = (int64_t)c->num_envs * c->num_steps;
    t->mb_size = (int)(t->B / c->num_minibatches);  // PPO_Discrete.cpp:246-247 (integer division)
    if (t->mb_size < 1) { set_error("ppo_trainer_create: minibatch size 0"); delete t; return nullptr; }
    t->n_mb = (int)((t->B + t->mb_size - 1) / t->mb_size);  // :573 loop count (last slice may be short)
    t->n_opt = t->n_mb * c->update_epochs;
    const bool masked = c->masked != 0;

    struct Item { const char* name; void** ptr; size_t bytes; };
    const int N = c->num_envs, T = c->num_steps, O = t->d.O, K = t->d.K, heads = t->d.heads;
    const size_t P = (size_t)t->P;
    std::vector<Item> items = {
        {"params", (void**)&t->params, P * 4},
        {"grads", (void**)&t->grads, (P + 8) * 4},  // + 8 diagnostics in the tail (one allreduce)
        {"exp_avg", (void**)&t->exp_avg, P * 4},
        {"exp_avg_sq", (void**)&t->exp_avg_sq, P * 4},
        {"state", (void**)&t->state, (size_t)S * N * 4},
        {"ep_len", (void**)&t->ep_len, (size_t)N * 4},
        {"ep_ret", (void**)&t->ep_ret, (size_t)N * 4},
        {"next_done", (void**)&t->next_done, (size_t)N * 4},
        {"next_value", (void**)&t->next_value, (size_t)N * 4},
        {"obs", (void**)&t->obs, (size_t)T * N * O * 4},
        {"actions", (void**)&t->actions, (size_t)T * N * heads * 4},
        {"logprobs", (void**)&t->logprobs, (size_t)T * N * 4},
        {"rewards", (void**)&t->rewards, (size_t)T * N * 4},
        {"dones", (void**)&t->dones, (size_t)T * N * 4},
        {"values", (void**)&t->values, (size_t)T * N * 4},
        {"adv", (void**)&t->adv, (size_t)T * N * 4},
        {"ret", (void**)&t->ret, (size_t)T * N * 4},
        {"masks", (void**)&t->masks, masked ? (size_t)T * N * K : 0},
        {"perm", (void**)&t->perm, (size_t)t->B * 4},
        {"adv_stats", (void**)&t->adv_stats, (size_t)t->n_mb * (c->update_epochs > 0 ? c->update_epochs : 1) * 2 * 4},
        {"adv_partial", (void**)&t->adv_partial, (size_t)t->n_mb * (c->update_epochs > 0 ? c->update_epochs : 1) * 3 * 8},
        {"diag", (void**)&t->diag_hist, (size_t)(t->n_opt > 0 ? t->n_opt : 1) * 8 * 4},
        {"ev_out", (void**)&t->ev_out, 4},
        {"norm_out", (void**)&t->norm_out, 4},
        {"ep_stats", (void**)&t->ep_stats, 3 * 8},
        {"scalars", (void**)&t->d_scal, sizeof(Scalars)},
        {"upd_ws", (void**)&t->upd_ws, (size_t)ppo_update_workspace_bytes(&c->dims)},
        {"ev_ws", (void**)&t->ev_ws, (size_t)ppo_explained_variance_workspace_bytes()},
        {"adv_ws", (void**)&t->adv_ws,
         (size_t)ppo_adv_stats_workspace_bytes(t->n_mb * (c->update_epochs > 0 ? c->update_epochs : 1))},
        {"fin_len", (void**)&t->fin_len, (size_t)T * N * 4},
        {"fin_ret", (void**)&t->fin_ret, (size_t)T * N * 4},
        {"tail", (void**)&t->tail, (size_t)(2 + 2 * PPO_EPISODE_TAIL) * 4},
        {"err_flag", (void**)&t->err_flag, 16},
        {"rec", (void**)&t->rec, (heads == 1 && !masked && O <= 4 && (t->d.H == 64 || t->d.H == 256)) ? (size_t)t->B * 64 : 0},
    };
    size_t total = 0;
    for (auto& it : items) total += al256(it.bytes);
    if (cudaMalloc(&t->arena, total) != cudaSuccess) {
        set_error("ppo_trainer_create: cudaMalloc(%zu B) failed: %s", total, cudaGetErrorString(cudaGetLastError()));
        t->arena = nullptr;
        ppo_trainer_destroy(t);
        return nullptr;
    }
    t->arena_bytes = total;
    cudaMemset(t->arena, 0, total);
    size_t off = 0;
    for (auto& it : items) {
        *it.ptr = it.bytes ? (void*)(t->arena + off) : nullptr;
        t->buf[it.name] = {*it.ptr, (int64_t)it.bytes};
        off += al256(it.bytes);
    }
    if (cudaMallocHost(&t->h_diag, (size_t)(t->n_opt > 0 ? t->n_opt : 1) * 8 * 4) != cudaSuccess ||
        cudaMallocHost(&t->h_rb, sizeof(Readback)) != cudaSuccess) {
        set_error("ppo_trainer_create: cudaMallocHost failed");
        ppo_trainer_destroy(t);
        return nullptr;
    }
    // hidden = 256 tensor-core update path: its 16-bit activation scratch (h1, dZ2 of one minibatch) is allocated here,
    // outside any stream capture
    if (c->update_epochs > 0 && ppo::update_h256_supported(t->d) && ppo::h256_reserve(t->mb_size) != PPO_OK) {
        ppo_trainer_destroy(t);
        return nullptr;
    }
    cudaDeviceSynchronize();
    return t;
}

// ---- peer-memory gradient exchange: CUDA IPC plumbing ---------------------------------------------------------------
static size_t xchg_bytes(const ppo_trainer* t) { return 2 * (size_t)ppo::peer_slot_bytes(t->P, t->cfg.world_size) + 256; }

extern "C" int ppo_trainer_peer_handle(ppo_trainer* t, char* h_handle64) {
    PPO_REQUIRE(t && h_handle64, "ppo_trainer_peer_handle: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!t->xchg) {
        PPO_CUDA(cudaMalloc(&t->xchg, xchg_bytes(t)));
        PPO_CUDA(cudaMemset(t->xchg, 0, xchg_bytes(t)));
        PPO_CUDA(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    PPO_CUDA(cudaIpcGetMemHandle(&h, t->xchg));
    memcpy(h_handle64, &h, 64);
    return PPO_OK;
}

extern "C" int ppo_trainer_set_peers(ppo_trainer* t, const char* h_handles, int world) {
    PPO_REQUIRE(t && h_handles, "ppo_trainer_set_peers: null argument");
    PPO_REQUIRE(world == t->cfg.world_size && world >= 2 && world <= ppo::PEER_MAX,
                "ppo_trainer_set_peers: world=%d (cfg.world_size=%d, at most %d)", world, t->cfg.world_size, ppo::PEER_MAX);
    PPO_REQUIRE(t->xchg, "ppo_trainer_set_peers: call ppo_trainer_peer_handle first");
    PPO_REQUIRE(!t->peer_ready, "ppo_trainer_set_peers: peers already set");
    const int blocks = (t->P + ppo::RED_COLS - 1) / ppo::RED_COLS;
    if (blocks > ppo::PEER_NB || blocks > sm_count()) {
        set_error("peer-memory exchange supports up to %d parameters on this device; NCCL path stays in use",
                  (sm_count() < ppo::PEER_NB ? sm_count() : ppo::PEER_NB) * ppo::RED_COLS);
        return PPO_ERR_UNSUPPORTED;
    }
    t->peer.world = world;
    t->peer.rank = t->cfg.rank;
    for (int r = 0; r < world; r++) {
        if (r == t->cfg.rank) { t->peer.xchg[r] = t->xchg; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, h_handles + (size_t)r * 64, 64);
        void* p = nullptr;
        PPO_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        t->peer.xchg[r] = (char*)p;
    }
    t->peer.err = t->err_flag;
    t->peer_ready = true;
    if (t->graph_exec) { cudaGraphExecDestroy(t->graph_exec); t->graph_exec = nullptr; }
    return PPO_OK;
}

extern "C" int ppo_trainer_set_comm(ppo_trainer* t, ppo_comm* comm) {
    PPO_REQUIRE(t, "null trainer");
    t->comm = comm;
    if (comm) PPO_REQUIRE(ppo_comm_world_size(comm) == t->cfg.world_size, "comm world size != cfg.world_size");
    return PPO_OK;
}

extern "C" void* ppo_trainer_buffer(ppo_trainer* t, const char* name, int64_t* bytes) {
    if (!t || !name) return nullptr;
    auto it = t->buf.find(name);
    if (it == t->buf.end()) { set_error("ppo_trainer_buffer: unknown buffer '%s'", name); return nullptr; }
    if (bytes) *bytes = it->second.second;
    return it->second.first;
}

extern "C" int ppo_trainer_set_params(ppo_trainer* t, const float* h) {
    PPO_REQUIRE(t && h, "null argument");
    PPO_CUDA(cudaMemcpy(t->params, h, (size_t)t->P * 4, cudaMemcpyHostToDevice));
    return PPO_OK;
}
extern "C" int ppo_trainer_get_params(ppo_trainer* t, float* h) {
    PPO_REQUIRE(t && h, "null argument");
    PPO_CUDA(cudaMemcpy(h, t->params, (size_t)t->P * 4, cudaMemcpyDeviceToHost));
    return PPO_OK;
}
extern "C" int ppo_trainer_get_opt_state(ppo_trainer* t, float* m, float* v, int64_t* step) {
    PPO_REQUIRE(t, "null trainer");
    if (m) PPO_CUDA(cudaMemcpy(m, t->exp_avg, (size_t)t->P * 4, cudaMemcpyDeviceToHost));
    if (v) PPO_CUDA(cudaMemcpy(v, t->exp_avg_sq, (size_t)t->P * 4, cudaMemcpyDeviceToHost));
    if (step) *step = t->step;
    return PPO_OK;
}
extern "C" int ppo_trainer_set_opt_state(ppo_trainer* t, const float* m, const float* v, int64_t step) {
    PPO_REQUIRE(t && step >= 0, "bad argument");
    if (m) PPO_CUDA(cudaMemcpy(t->exp_avg, m, (size_t)t->P * 4, cudaMemcpyHostToDevice));
    if (v) PPO_CUDA(cudaMemcpy(t->exp_avg_sq, v, (size_t)t->P * 4, cudaMemcpyHostToDevice));
    if (t->xchg && step != t->step) {
        // flags in the exchange buffer are step numbers: after a jump of the step counter stale ones could satisfy a wait
        // (or block one).  Every rank calls this collectively and synchronises (host barrier) before the next iterate.
        PPO_CUDA(cudaDeviceSynchronize());
        PPO_CUDA(cudaMemset(t->xchg, 0, xchg_bytes(t)));
        PPO_CUDA(cudaDeviceSynchronize());
    }
    t->step = step;
    return PPO_OK;
}

// initEnvs (PPO_Discrete.cpp:365-402): reset every env; next_done = 0 (:502)
extern "C" int ppo_trainer_reset_envs(ppo_trainer* t) {
    PPO_REQUIRE(t, "null trainer");
    if (t->cfg.env_kind == PPO_ENV_HOST) return PPO_OK;   // nothing on the device to reset
    const int N = t->cfg.num_envs;
    // env RNG key = global env index; ppo_env_reset keys on the local index, so offset the seed stream
    // through the counter: use the dedicated init counter ~0 and a rank-salted seed.
    int rc = ppo_env_reset(t->cfg.env_kind, N, t->state, t->ep_len, t->ep_ret, nullptr,
                           t->cfg.seed + 0x9E3779B97F4A7C15ull * (uint64_t)(t->cfg.rank + 1), ~0ull, nullptr, nullptr);
    if (rc) return rc;
    PPO_CUDA(cudaMemsetAsync(t->next_done, 0, (size_t)N * 4, nullptr));
    PPO_CUDA(cudaDeviceSynchronize());
    return PPO_OK;
}

// ---- phases ---------------------------------------------------------------------------------------
// by-value kernel arguments are snapshotted at launch, so back-to-back asynchronous iterations
// cannot race on a staging buffer
__global__ void set_scalars_kernel(Scalars* dst, Scalars v) { *dst = v; }
static int push_scalars(ppo_trainer* t, float lr, cudaStream_t st) {
    Scalars v;
    v.rollout_counter = t->rollout_counter;
    v.epoch_counter = t->epoch_counter;
    v.step_base = t->step;
    v.lr = lr;
    v.pad = 0.0f;
    set_scalars_kernel<<<1, 1, 0, st>>>(t->d_scal, v);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

static int g_pack_rows = 1;   // ppo_debug_pack_rows
extern "C" int ppo_debug_pack_rows(int on) {
    g_pack_rows = on;
    return PPO_OK;
}

static int phase_rollout(ppo_trainer* t, cudaStream_t st, bool external = false) {
    const ppo_trainer_cfg& c = t->cfg;
    if (external) {   // the rollout rows, next_value / next_done (and ep_stats / the episode log) were filled by the caller
        PhaseTimer pt(t, PH_GAE, st);
        return ppo_gae(c.num_steps, c.num_envs, t->rewards, t->values, t->dones, t->next_value, t->next_done, c.gamma,
                       c.gae_lambda, c.use_gae, t->adv, t->ret, st);
    }
    PPO_REQUIRE(c.env_kind != PPO_ENV_HOST, "host environments: fill the rollout rows and call ppo_trainer_update");
    PPO_CUDA(cudaMemsetAsync(t->ep_stats, 0, 3 * 8, st));
    ppo_rollout_cfg rc{};
    rc.env_kind = c.env_kind; rc.n_envs = c.num_envs; rc.n_steps = c.num_steps;
    rc.max_episode_steps = c.max_episode_steps; rc.gamma = c.gamma; rc.gae_lambda = c.gae_lambda;
    rc.use_gae = c.use_gae; rc.entropy_mode = c.hyper.entropy_mode; rc.masked = c.masked;
    rc.env_id_base = (uint32_t)c.rank * (uint32_t)c.num_envs; rc.seed = c.seed;
    rc.counter = 0; rc.d_counter = &t->d_scal->rollout_counter;
    rc.fin_len = t->fin_len; rc.fin_ret = t->fin_ret;
    {
        PhaseTimer pt(t, PH_ROLLOUT, st);
        int r = ppo_rollout(t->params, &c.dims, &rc, t->state, t->ep_len, t->ep_ret, t->next_done, t->next_value,
                            t->obs, t->actions, t->logprobs, t->rewards, t->dones, t->values, nullptr, nullptr,
                            t->masks, t->ep_stats, st);
        if (r) return r;
    }
    PhaseTimer pt(t, PH_GAE, st);
    return ppo_gae(c.num_steps, c.num_envs, t->rewards, t->values, t->dones, t->next_value, t->next_done, c.gamma,
                   c.gae_lambda, c.use_gae, t->adv, t->ret, st);
}

static uint64_t perm_seed(const ppo_trainer_cfg& c) {   // rank-salted: every rank permutes its own shard
    return c.seed ^ (0xA5A5A5A5ull * (uint64_t)(c.rank + 1));
}

static int phase_begin_epoch(ppo_trainer* t, int epoch, cudaStream_t st) {
    const ppo_trainer_cfg& c = t->cfg;
    PhaseTimer pt(t, PH_EPOCH, st);
    // rank-salted seed: every rank permutes its own shard independently
    int rc = ppo_make_permutation(t->B, perm_seed(c), (uint64_t)epoch, &t->d_scal->epoch_counter, t->perm, st);
    if (rc) return rc;
    if (!c.hyper.norm_adv) return PPO_OK;
    if (c.world_size > 1 && t->comm) {
        rc = ppo_adv_stats(t->adv, t->perm, t->B, t->mb_size, t->n_mb, nullptr, t->adv_partial, t->adv_ws, st);
        if (rc) return rc;
        rc = ppo_comm_allreduce_f64(t->comm, t->adv_partial, (int64_t)t->n_mb * 3, st);
        if (rc) return rc;
        return ppo_adv_stats_finalize(t->adv_partial, t->n_mb, t->adv_stats, st);
    }
    return ppo_adv_stats(t->adv, t->perm, t->B, t->mb_size, t->n_mb, t->adv_stats, nullptr, t->adv_ws, st);
}

static int phase_minibatch_grads(ppo_trainer* t, int mb, cudaStream_t st) {
    const ppo_trainer_cfg& c = t->cfg;
    const int64_t start = (int64_t)mb * t->mb_size;
    int64_t cnt = t->B - start;
    if (cnt > t->mb_size) cnt = t->mb_size;
    PPO_REQUIRE(cnt >= 1, "minibatch %d is empty", mb);
    const float inv_n = 1.0f / ((float)cnt * (float)c.world_size);
    PhaseTimer pt(t, PH_UPDATE, st);
    return ppo::update_minibatch_fused(t->params, &c.dims, t->obs, t->masks, t->actions, t->logprobs, t->adv, t->ret,
                                       t->values, t->perm, start, (int)cnt, inv_n, t->adv_stats + 2 * mb, &c.hyper,
                                       t->upd_ws, t->grads, t->grads + t->P, nullptr, st, nullptr, nullptr,
                                       t->use_rec ? t->rec : nullptr);
}

// single-GPU iterations: gradient + reduction + clip + AdamW back to back, the optimiser fused into the reduction launch
// whole-iteration path: the E x MB minibatch statistics in one launch (+ one allreduce under data parallelism)
static int phase_all_epoch_stats(ppo_trainer* t, cudaStream_t st) {
    const ppo_trainer_cfg& c = t->cfg;
    if (!c.hyper.norm_adv || c.update_epochs == 0) return PPO_OK;
    const int n_all = t->n_mb * c.update_epochs;
    const bool dp = c.world_size > 1 && t->comm;
    int rc = ppo_adv_stats_all_epochs(t->adv, t->B, t->mb_size, t->n_mb, c.update_epochs, perm_seed(c), 0,
                                      &t->d_scal->epoch_counter, dp ? nullptr : t->adv_stats, dp ? t->adv_partial : nullptr,
                                      t->adv_ws, st);
    if (rc || !dp) return rc;
    if ((rc = ppo_comm_allreduce_f64(t->comm, t->adv_partial, (int64_t)n_all * 3, st))) return rc;
    return ppo_adv_stats_finalize(t->adv_partial, n_all, t->adv_stats, st);
}

static int phase_minibatch_step(ppo_trainer* t, int epoch, int mb, cudaStream_t st) {
    const ppo_trainer_cfg& c = t->cfg;
    const int64_t start = (int64_t)mb * t->mb_size;
    int64_t cnt = t->B - start;
    if (cnt > t->mb_size) cnt = t->mb_size;
    PPO_REQUIRE(cnt >= 1, "minibatch %d is empty", mb);
    const int k = t->opt_in_iter++;
    PPO_REQUIRE(k < t->n_opt, "more optimiser steps than update_epochs * num_minibatches in one iteration");
    ppo::FuseAdam f{};
    f.params = t->params; f.exp_avg = t->exp_avg; f.exp_avg_sq = t->exp_avg_sq; f.d_lr = &t->d_scal->lr;
    f.beta1 = c.beta1; f.beta2 = c.beta2; f.eps = c.adam_eps; f.weight_decay = c.weight_decay;
    f.step_host = (int64_t)k + 1; f.d_step = &t->d_scal->step_base; f.max_grad_norm = c.max_grad_norm;
    f.norm_out = t->norm_out; f.diag_hist = t->diag_hist + (size_t)k * 8;
    f.err = t->err_flag; f.spin_cycles = ppo::g_spin_cycles;
    t->peer.spin_cycles = ppo::g_spin_cycles;
    const bool peer = c.world_size > 1 && t->peer_ready;   // enqueue_iteration only takes this path with peers mapped
    ppo::PermFn pf{};
    pf.inline_on = 1; pf.half_bits = ppo::perm_half_bits(t->B); pf.batch = t->B; pf.seed = perm_seed(c);
    pf.epoch = (uint64_t)epoch; pf.d_epoch = &t->d_scal->epoch_counter;
    return ppo::update_minibatch_fused(t->params, &c.dims, t->obs, t->masks, t->actions, t->logprobs, t->adv, t->ret,
                                       t->values, nullptr, start, (int)cnt, 1.0f / ((float)cnt * (float)c.world_size),
                                       t->adv_stats + 2 * ((size_t)epoch * t->n_mb + mb), &c.hyper, t->upd_ws, t->grads,
                                       t->grads + t->P, &f, st, peer ? &t->peer : nullptr, &pf,
                                       t->use_rec ? t->rec : nullptr);
}

static int phase_apply(ppo_trainer* t, cudaStream_t st) {
    const ppo_trainer_cfg& c = t->cfg;
    PhaseTimer pt(t, PH_APPLY, st);
    if (c.world_size > 1 && t->comm) {
        int rc = ppo_comm_allreduce_f32(t->comm, t->grads, (int64_t)t->P + 8, st);
        if (rc) return rc;
    }
    const int k = t->opt_in_iter++;
    PPO_REQUIRE(k < t->n_opt, "more optimiser steps than update_epochs * num_minibatches in one iteration");
    return clip_adamw_launch(t->params, t->grads, t->exp_avg, t->exp_avg_sq, t->P, 0.0f, &t->d_scal->lr, c.beta1,
                             c.beta2, c.adam_eps, c.weight_decay, (int64_t)k + 1, &t->d_scal->step_base,
                             c.max_grad_norm, 1.0f, t->norm_out, t->grads + t->P, t->diag_hist + (size_t)k * 8, st);
}

static int phase_tail(ppo_trainer* t, cudaStream_t st) {
    PhaseTimer pt(t, PH_TAIL, st);
    int rc = ppo_explained_variance(t->B, t->ret, t->values, t->ev_ws, t->ev_out, st);
    if (rc) return rc;
    if (t->n_opt > 0)
        PPO_CUDA(cudaMemcpyAsync(t->h_diag, t->diag_hist, (size_t)t->n_opt * 32, cudaMemcpyDeviceToHost, st));
    PPO_CUDA(cudaMemcpyAsync(&t->h_rb->ev, t->ev_out, 4, cudaMemcpyDeviceToHost, st));
    PPO_CUDA(cudaMemcpyAsync(&t->h_rb->grad_norm, t->norm_out, 4, cudaMemcpyDeviceToHost, st));
    PPO_CUDA(cudaMemcpyAsync(t->h_rb->ep_stats, t->ep_stats, 24, cudaMemcpyDeviceToHost, st));
    // the episodes stepEnvs would have left in its CircularBuffer(100) (PPO/PPO_Discrete.cpp:474-480)
    if ((rc = ppo_episode_tail(t->fin_len, t->fin_ret, t->B, PPO_EPISODE_TAIL, t->tail, t->tail + 2,
                               reinterpret_cast<float*>(t->tail + 2 + PPO_EPISODE_TAIL), st)))
        return rc;
    static_assert(offsetof(Readback, tail_ret) - offsetof(Readback, tail_count) == (2 + PPO_EPISODE_TAIL) * 4,
                  "Readback tail block mirrors the device tail buffer");
    PPO_CUDA(cudaMemcpyAsync(&t->h_rb->tail_count, t->tail, (size_t)(2 + 2 * PPO_EPISODE_TAIL) * 4,
                             cudaMemcpyDeviceToHost, st));
    PPO_CUDA(cudaMemcpyAsync(&t->h_rb->peer_err, t->err_flag, 4, cudaMemcpyDeviceToHost, st));
    return PPO_OK;
}

static int enqueue_iteration(ppo_trainer* t, cudaStream_t st, bool external_rollout = false) {
    int rc;
    t->opt_in_iter = 0;
    if ((rc = phase_rollout(t, st, external_rollout))) return rc;
    const bool dp = t->cfg.world_size > 1 && t->comm;
    // world_size > 1 needs a transport: the NCCL communicator (advantage moments, and gradients without peer memory)
    PPO_REQUIRE(t->cfg.world_size == 1 || t->comm,
                "world_size = %d but no communicator: call ppo_trainer_set_comm (and optionally ppo_trainer_set_peers) first",
                t->cfg.world_size);
    const bool fused = (!dp || t->peer_ready) && !t->profile;
    if (fused) {   // no per-epoch launches: statistics of every epoch up front, permutation evaluated inside the update kernel
        if ((rc = phase_all_epoch_stats(t, st))) return rc;
        const bool packed = t->rec && g_pack_rows && t->cfg.update_epochs > 0;
        if (packed) {
            PhaseTimer pt(t, PH_EPOCH, st);
            if ((rc = ppo_pack_rows(t->B, t->d.O, t->obs, t->actions, t->logprobs, t->adv, t->ret, t->values, t->rec, st)))
                return rc;
        }
        t->use_rec = packed;
        for (int e = 0; e < t->cfg.update_epochs; e++)
            for (int mb = 0; mb < t->n_mb; mb++)
                if ((rc = phase_minibatch_step(t, e, mb, st))) return rc;
        return phase_tail(t, st);
    }
    // split path inside iterate (NCCL data parallelism, profiling): same packed gather, rows indexed through `perm`
    t->use_rec = t->rec && g_pack_rows && t->cfg.update_epochs > 0;
    if (t->use_rec) {
        PhaseTimer pt(t, PH_EPOCH, st);
        if ((rc = ppo_pack_rows(t->B, t->d.O, t->obs, t->actions, t->logprobs, t->adv, t->ret, t->values, t->rec, st)))
            return rc;
    }
    for (int e = 0; e < t->cfg.update_epochs; e++) {
        if ((rc = phase_begin_epoch(t, e, st))) return rc;
        for (int mb = 0; mb < t->n_mb; mb++) {
            if ((rc = phase_minibatch_grads(t, mb, st))) return rc;
            if ((rc = phase_apply(t, st))) return rc;
        }
    }
    return phase_tail(t, st);
}

static void advance_counters(ppo_trainer* t) {
    t->rollout_counter += (uint64_t)t->cfg.num_steps;
    t->epoch_counter += (uint64_t)t->cfg.update_epochs;
    t->step += t->n_opt;
}

static void fill_result(ppo_trainer* t, float lr, ppo_iter_result* r) {
    memset(r, 0, sizeof(*r));
    if (t->n_opt > 0) {
        const float* last = t->h_diag + (size_t)(t->n_opt - 1) * 8;
        r->pg_loss = last[0]; r->v_loss = last[1]; r->entropy = last[2]; r->loss = last[3]; r->approx_kl = last[4];
        double cf = 0.0;  // clip_fraction = mean over all minibatches of the update (Utils.cpp:5-14)
        for (int k = 0; k < t->n_opt; k++) cf += t->h_diag[(size_t)k * 8 + 5];
        r->clipfrac = (float)(cf / t->n_opt);
    }
    r->explained_variance = t->h_rb->ev;
    r->grad_norm = t->h_rb->grad_norm;
    r->episodes = t->h_rb->ep_stats[0];
    r->ep_len_sum = t->h_rb->ep_stats[1];
    r->ep_ret_sum = t->h_rb->ep_stats[2];
    r->lr = lr;
}

// Everything of an iteration BEHIND the rollout, on rollout rows the caller filled (obs, actions, logprobs, rewards, dones,
// values [T, N ..], next_value, next_done [N], optionally fin_len / fin_ret / ep_stats): calcAdvantage + the E x MB optimiser
// steps + diagnostics (PPO_Discrete.cpp:556-673).  For host-side environments (f4) and callers with their own rollout.
extern "C" int ppo_trainer_update(ppo_trainer* t, float h_lr, ppo_iter_result* h_result, ppo_stream_t stream) {
    PPO_REQUIRE(t, "null trainer");
    cudaStream_t st = as_stream(stream);
    int rc;
    if ((rc = push_scalars(t, h_lr, st))) return rc;
    if ((rc = enqueue_iteration(t, st, true))) return rc;
    advance_counters(t);
    if (h_result) {
        PPO_CUDA(cudaStreamSynchronize(st));
        fill_result(t, h_lr, h_result);
        if (t->h_rb->peer_err) {
            set_error("an in-kernel wait ran out: the optimiser steps from there on were skipped");
            return PPO_ERR_COMM;
        }
    }
    return PPO_OK;
}

extern "C" int ppo_trainer_iterate(ppo_trainer* t, float h_lr, ppo_iter_result* h_result, ppo_stream_t stream) {
    PPO_REQUIRE(t, "null trainer");
    cudaStream_t st = as_stream(stream);
    int rc;
    if ((rc = push_scalars(t, h_lr, st))) return rc;
    if (t->cfg.use_cuda_graph && !t->profile) {
        if (t->graph_exec && t->graph_stream != st) {  // a graph is tied to nothing, but keep it simple
            cudaGraphExecDestroy(t->graph_exec);
            t->graph_exec = nullptr;
        }
        if (!t->graph_exec) {
            cudaGraph_t g = nullptr;
            PPO_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            rc = enqueue_iteration(t, st);
            cudaError_t e = cudaStreamEndCapture(st, &g);
            if (rc) { if (g) cudaGraphDestroy(g); return rc; }
            PPO_CUDA(e);
            PPO_CUDA(cudaGraphInstantiate(&t->graph_exec, g, 0));
            cudaGraphDestroy(g);
            t->graph_stream = st;
        }
        PPO_CUDA(cudaGraphLaunch(t->graph_exec, st));
    } else {
        if ((rc = enqueue_iteration(t, st))) return rc;
    }
    advance_counters(t);
    if (h_result) {
        PPO_CUDA(cudaStreamSynchronize(st));
        fill_result(t, h_lr, h_result);
        if (t->h_rb->peer_err) {
            set_error("an in-kernel wait ran out (%s): the optimiser steps from there on were skipped",
                      t->peer_ready ? "a rank did not publish its gradient block within the spin limit" : "grid ticket");
            return PPO_ERR_COMM;
        }
    }
    return PPO_OK;
}

// ---- split-phase API (host loops that interleave their own work between phases) ---------------------
extern "C" int ppo_trainer_last_episodes(ppo_trainer* t, int max_n, int32_t* h_len, float* h_ret) {
    if (!t || !h_len || !h_ret || max_n < 0) return 0;
    int n = t->h_rb->tail_count;
    if (n > PPO_EPISODE_TAIL) n = PPO_EPISODE_TAIL;
    const int skip = n > max_n ? n - max_n : 0;   // keep the newest max_n
    for (int i = skip; i < n; i++) { h_len[i - skip] = t->h_rb->tail_len[i]; h_ret[i - skip] = t->h_rb->tail_ret[i]; }
    return n - skip;
}

extern "C" int ppo_trainer_rollout(ppo_trainer* t, float h_lr, ppo_stream_t stream) {
    PPO_REQUIRE(t, "null trainer");
    int rc = push_scalars(t, h_lr, as_stream(stream));
    if (rc) return rc;
    t->opt_in_iter = 0;
    t->use_rec = false;   // the phase-by-phase API reads the row arrays (callers may edit them between phases)
    return phase_rollout(t, as_stream(stream));
}
extern "C" int ppo_trainer_begin_epoch(ppo_trainer* t, int epoch, ppo_stream_t stream) {
    PPO_REQUIRE(t && epoch >= 0, "bad argument");
    return phase_begin_epoch(t, epoch, as_stream(stream));
}
extern "C" int ppo_trainer_minibatch_grads(ppo_trainer* t, int mb, ppo_stream_t stream) {
    PPO_REQUIRE(t && mb >= 0 && mb < t->n_mb, "minibatch index out of range");
    return phase_minibatch_grads(t, mb, as_stream(stream));
}
extern "C" int ppo_trainer_apply(ppo_trainer* t, ppo_stream_t stream) {
    PPO_REQUIRE(t, "null trainer");
    return phase_apply(t, as_stream(stream));
}
extern "C" int ppo_trainer_finish(ppo_trainer* t, float h_lr, ppo_iter_result* h_result, ppo_stream_t stream) {
    PPO_REQUIRE(t, "null trainer");
    cudaStream_t st = as_stream(stream);
    int rc = phase_tail(t, st);
    if (rc) return rc;
    // counters advance by what was actually issued
    t->rollout_counter += (uint64_t)t->cfg.num_steps;
    t->epoch_counter += (uint64_t)t->cfg.update_epochs;
    t->step += t->opt_in_iter;
    if (h_result) {
        PPO_CUDA(cudaStreamSynchronize(st));
        fill_result(t, h_lr, h_result);
    }
    return PPO_OK;
}

// ---- per-phase device timing -------------------------------------------------------------------------
// ppo_trainer_profile(t, 1) makes every later phase record a cudaEvent pair on its launch stream (graph
// replay is bypassed while profiling).  ppo_trainer_phase_ms synchronises, returns the accumulated
// milliseconds and launch counts per phase {rollout, gae, epoch setup, update, apply, tail} and resets.
extern "C" int ppo_trainer_profile(ppo_trainer* t, int enable) {
    PPO_REQUIRE(t, "null trainer");
    t->profile = enable != 0;
    t->ev_used = 0;
    return PPO_OK;
}
extern "C" int ppo_trainer_phase_ms(ppo_trainer* t, float* h_ms6, int32_t* h_count6) {
    PPO_REQUIRE(t && h_ms6, "null argument");
    PPO_CUDA(cudaDeviceSynchronize());
    for (int i = 0; i < PH_COUNT; i++) { h_ms6[i] = 0.0f; if (h_count6) h_count6[i] = 0; }
    for (size_t i = 0; i + 1 < t->ev_used; i += 2) {
        float ms = 0.0f;
        PPO_CUDA(cudaEventElapsedTime(&ms, t->ev_pool[i], t->ev_pool[i + 1]));
        h_ms6[t->ev_phase[i / 2]] += ms;
        if (h_count6) h_count6[t->ev_phase[i / 2]] += 1;
    }
    t->ev_used = 0;
    return PPO_OK;
}
