// oracle/ref_shim.cpp — TEST INFRASTRUCTURE, not product code.
//
// A thin C ABI over the UNMODIFIED reference sources (compiled where they lie under
// /root/reference by oracle/Makefile; nothing is copied into this repo).  It lets the
// pytest harness drive the reference's own classes on identical inputs so that
// (1) oracle/ppo_oracle.c (the portable CPU restatement) can be pinned against the real
// reference, and (2) tests/golden/*.npz fixtures can be generated (tests/golden/make_golden.py).
//
// Reference symbols exercised:
//   CartPole::{reset,step}            Environments/CartPole.cpp:34-45,47-94
//   MountainCar::step                 Environments/MountainCar.cpp:29-57
//   Agent::{getValue,getActionAndValueDiscrete,getActionAndValueMasked}  PPO/Agent.cpp:107-170
//   Categorical / CategoricalMasked   Distributions/*.cpp
//   PPO_Discrete::{calcAdvantage,stepEnvs,initEnvs,getApproxKLAndClippedObj}  PPO/PPO_Discrete.cpp:274-483
//   torch::nn::utils::clip_grad_norm_ + torch::optim::AdamW   (LibTorch, PPO_Discrete.cpp:76-78,637-641)
// The minibatch loss block (PPO_Discrete.cpp:585-631) is inline in train() and therefore
// cannot be called; ref_minibatch_step() restates it with the same ATen ops and lets the
// real LibTorch autograd / clip_grad_norm_ / AdamW produce gradients and the parameter step.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu/reference legs may load this.

#include <ftw.h>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <filesystem>
#include <string>
#include <vector>

#include "PPO/PPO_Discrete.h"

namespace {
thread_local std::string g_err;
int fail(const std::exception& e) { g_err = e.what(); return -1; }

torch::Tensor f32(const float* p, std::vector<int64_t> shape) {
    return torch::from_blob(const_cast<float*>(p), shape, torch::kFloat32).clone();
}

struct RefTrainer {
    std::shared_ptr<torch::Device> device;
    std::shared_ptr<Agent> agent;
    std::shared_ptr<torch::optim::AdamW> opt;
};

// Scratch cwd helpers in plain POSIX (std::filesystem::remove_all from a dlopen()ed library
// crashed under python in this image, so the shim itself stays off std::filesystem).
int rm_cb(const char* p, const struct stat*, int, struct FTW*) { return remove(p); }
struct ScratchCwd {
    char old[4096];
    char tmp[64];
    ScratchCwd() {
        if (!getcwd(old, sizeof old)) old[0] = 0;
        std::snprintf(tmp, sizeof tmp, "/tmp/ppo_ref_XXXXXX");
        if (!mkdtemp(tmp)) throw std::runtime_error("mkdtemp failed");
    }
    void enter() { if (chdir(tmp) != 0) throw std::runtime_error("chdir failed"); }
    ~ScratchCwd() {
        if (old[0] && chdir(old) != 0) {}
        nftw(tmp, rm_cb, 16, FTW_DEPTH | FTW_PHYS);
    }
};

void copy_out(const torch::Tensor& t, float* dst) {
    torch::Tensor c = t.detach().to(torch::kFloat32).contiguous();
    std::memcpy(dst, c.data_ptr<float>(), sizeof(float) * c.numel());
}
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

// ---- environments -----------------------------------------------------------------
// n_resets consecutive reset() calls on one CartPole(seed): out[n_resets*4].
int ref_cartpole_reset_seq(int64_t seed, int n_resets, float* out) {
    try {
        CartPole env(seed);
        for (int r = 0; r < n_resets; r++) {
            std::vector<float> s = env.reset();
            std::memcpy(out + 4 * r, s.data(), 4 * sizeof(float));
        }
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// One step() per row from an injected state.
int ref_cartpole_step(const float* state_in, const int64_t* actions, int n,
                      float* state_out, float* reward, int* terminated) {
    try {
        CartPole env(0);
        for (int i = 0; i < n; i++) {
            env.state.assign(state_in + 4 * i, state_in + 4 * i + 4);
            env.terminated = false;
            auto [obs, r, term, info] = env.step(actions[i]);
            std::memcpy(state_out + 4 * i, obs.data(), 4 * sizeof(float));
            reward[i] = r;
            terminated[i] = term ? 1 : 0;
        }
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

int ref_mountaincar_step(const float* state_in, const int64_t* actions, int n,
                         float* state_out, float* reward, int* terminated) {
    try {
        MountainCar env;
        for (int i = 0; i < n; i++) {
            env.state = {state_in[2 * i], state_in[2 * i + 1]};
            auto [obs, r, term, info] = env.step(actions[i]);
            state_out[2 * i] = obs[0];
            state_out[2 * i + 1] = obs[1];
            reward[i] = r;
            terminated[i] = term ? 1 : 0;
        }
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// ---- distributions ----------------------------------------------------------------
// mask == nullptr -> Categorical, else CategoricalMasked (mask u8 [n,K]).
int ref_categorical(const float* logits, const unsigned char* mask, int n, int K,
                    const int64_t* actions, float* norm_logits, float* probs,
                    float* logprob, float* entropy) {
    try {
        auto dev = std::make_shared<torch::Device>(torch::kCPU);
        torch::Tensor lg = f32(logits, {n, K});
        torch::Tensor act = torch::from_blob(const_cast<int64_t*>(actions), {n}, torch::kInt64).clone();
        if (mask == nullptr) {
            Categorical d(lg, dev);
            copy_out(d.m_logits, norm_logits);
            copy_out(d.m_probs, probs);
            copy_out(d.log_prob(act), logprob);
            copy_out(d.entropy(), entropy);
        } else {
            torch::Tensor mk = torch::from_blob(const_cast<unsigned char*>(mask), {n, K}, torch::kUInt8).clone();
            CategoricalMasked d(lg, mk, dev);
            copy_out(d.m_logits, norm_logits);
            copy_out(d.m_probs, probs);
            copy_out(d.log_prob(act), logprob);
            copy_out(d.entropy(), entropy);
        }
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// ---- agent / trainer handle ---------------------------------------------------------
void* ref_trainer_create(int64_t obs_size, int64_t action_size, float lr, int64_t seed) {
    try {
        torch::manual_seed(seed);
        auto* t = new RefTrainer();
        t->device = std::make_shared<torch::Device>(torch::kCPU);
        t->agent = std::make_shared<Agent>(obs_size, action_size, t->device);
        // PPO/PPO_Discrete.cpp:76-78
        t->opt = std::make_shared<torch::optim::AdamW>(t->agent->parameters(),
                                                      torch::optim::AdamWOptions(lr).eps(1e-5f));
        return t;
    } catch (const std::exception& e) { fail(e); return nullptr; }
}

void ref_trainer_destroy(void* h) { delete static_cast<RefTrainer*>(h); }

// The reference's checkpoint calls on the reference's own Agent class and AdamW:
// torch::save(m_agent, ...) / torch::save(*m_optimizer, ...) (PPO/PPO_Discrete.cpp:670-672) and
// torch::load(m_agent, ...) / torch::load(*m_optimizer, ...) (:813, :832).  (The reference BINARY cannot be used for
// this on Linux: loadPolicyFromCheckpoint compares last_write_time against a default-constructed file_time_type, which
// is the year-2174 epoch of libstdc++'s file_clock, so it never finds a newest file and throws on the empty name.)
int ref_trainer_save(void* h, const char* agent_file, const char* optimizer_file) {
    auto* t = static_cast<RefTrainer*>(h);
    try {
        if (agent_file) torch::save(t->agent, agent_file);
        if (optimizer_file) torch::save(*t->opt, optimizer_file);
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}
int ref_trainer_load(void* h, const char* agent_file, const char* optimizer_file) {
    auto* t = static_cast<RefTrainer*>(h);
    try {
        if (agent_file) torch::load(t->agent, agent_file);
        if (optimizer_file) torch::load(*t->opt, optimizer_file);
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}
// AdamW state of parameter `index` after a load / step: step count and the two moment tensors (flat)
int ref_trainer_opt_state(void* h, int index, int64_t* step, float* exp_avg, float* exp_avg_sq) {
    auto* t = static_cast<RefTrainer*>(h);
    try {
        auto& p = t->opt->param_groups()[0].params()[index];
        auto it = t->opt->state().find(p.unsafeGetTensorImpl());
        if (it == t->opt->state().end()) { *step = -1; return 0; }
        auto& st = static_cast<torch::optim::AdamWParamState&>(*it->second);
        *step = st.step();
        if (exp_avg) copy_out(st.exp_avg(), exp_avg);
        if (exp_avg_sq) copy_out(st.exp_avg_sq(), exp_avg_sq);
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// Multi-head nvec for getActionAndValueMasked (PPO/Agent.cpp:21 keeps a single head; the
// member is public, the forward at :137-170 is generic over it).  Sum must equal action_size.
int ref_trainer_set_nvec(void* h, const int64_t* nvec, int heads) {
    auto* t = static_cast<RefTrainer*>(h);
    t->agent->m_actionSpace.assign(nvec, nvec + heads);
    return 0;
}

int64_t ref_trainer_num_params(void* h) {
    auto* t = static_cast<RefTrainer*>(h);
    int64_t n = 0;
    for (auto& p : t->agent->parameters()) n += p.numel();
    return n;
}

// Flat order == Agent::parameters() order == checkpoint order:
// critic {W1,b1,W2,b2,W3,b3}, actor {W1,b1,W2,b2,W3,b3}; weights [out,in] row-major.
int ref_trainer_get_params(void* h, float* flat) {
    auto* t = static_cast<RefTrainer*>(h);
    int64_t off = 0;
    for (auto& p : t->agent->parameters()) { copy_out(p, flat + off); off += p.numel(); }
    return 0;
}

int ref_trainer_set_params(void* h, const float* flat) {
    auto* t = static_cast<RefTrainer*>(h);
    torch::NoGradGuard ng;
    int64_t off = 0;
    for (auto& p : t->agent->parameters()) {
        p.copy_(torch::from_blob(const_cast<float*>(flat + off), p.sizes(), torch::kFloat32));
        off += p.numel();
    }
    return 0;
}

int ref_trainer_get_grads(void* h, float* flat) {
    auto* t = static_cast<RefTrainer*>(h);
    int64_t off = 0;
    for (auto& p : t->agent->parameters()) {
        if (p.grad().defined()) copy_out(p.grad(), flat + off);
        else std::memset(flat + off, 0, sizeof(float) * p.numel());
        off += p.numel();
    }
    return 0;
}

// Raw network outputs: logits [n,K] (m_Actor) and value [n] (getValue).
int ref_agent_forward(void* h, const float* obs, int n, int O, float* logits, float* value) {
    try {
        auto* t = static_cast<RefTrainer*>(h);
        torch::NoGradGuard ng;
        torch::Tensor x = f32(obs, {n, O});
        copy_out(t->agent->m_Actor->forward(x), logits);
        copy_out(t->agent->getValue(x), value);
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// getActionAndValueDiscrete / getActionAndValueMasked with GIVEN actions (update-time path).
// mask == nullptr -> Discrete (actions [n]); else Masked (actions [n,heads], mask u8 [n,sumK]).
int ref_agent_evaluate(void* h, const float* obs, int n, int O, const unsigned char* mask,
                       const int64_t* actions, int heads, float* logprob, float* entropy, float* value) {
    try {
        auto* t = static_cast<RefTrainer*>(h);
        torch::NoGradGuard ng;
        torch::Tensor x = f32(obs, {n, O});
        AgentOutput out;
        if (mask == nullptr) {
            torch::Tensor act = torch::from_blob(const_cast<int64_t*>(actions), {n}, torch::kInt64).clone();
            out = t->agent->getActionAndValueDiscrete(x, act);
        } else {
            int64_t sumK = t->agent->m_actionSpaceSum;
            torch::Tensor mk = torch::from_blob(const_cast<unsigned char*>(mask), {n, sumK}, torch::kUInt8).clone();
            // the reference indexes action[i] per head => action is [heads, n] inside (Agent.cpp:157-160)
            torch::Tensor act = torch::from_blob(const_cast<int64_t*>(actions), {n, heads}, torch::kInt64).clone().t().contiguous();
            out = t->agent->getActionAndValueMasked(x, mk, act);
        }
        copy_out(out.logprob, logprob);
        copy_out(out.entropy, entropy);
        copy_out(out.value.reshape({-1}), value);
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// One optimiser step on an explicit minibatch, restating PPO/PPO_Discrete.cpp:579-641 with the
// same ATen ops; autograd, clip_grad_norm_ and AdamW are the real LibTorch ones.
// diag[8] = {pg_loss, v_loss, entropy_loss, loss, approx_kl, clipfrac, grad_norm(pre-clip), lr}
// grads_out (optional) receives the PRE-clip gradients in flat parameter order.
int ref_minibatch_step(void* h, const float* obs, int n, int O, const unsigned char* mask, int heads,
                       const int64_t* actions, const float* old_logprob, const float* advantages,
                       const float* returns, const float* old_values,
                       float lr, float clip_coef, int norm_adv, int clip_vloss,
                       float ent_coef, float vf_coef, float max_grad_norm,
                       float* diag, float* grads_out) {
    try {
        auto* t = static_cast<RefTrainer*>(h);
        static_cast<torch::optim::AdamWOptions&>(t->opt->param_groups()[0].options()).lr(lr);
        torch::Tensor x = f32(obs, {n, O});
        torch::Tensor b_logp = f32(old_logprob, {n});
        torch::Tensor mb_adv = f32(advantages, {n});
        torch::Tensor b_ret = f32(returns, {n});
        torch::Tensor b_val = f32(old_values, {n});
        AgentOutput out;
        if (mask == nullptr) {
            torch::Tensor act = torch::from_blob(const_cast<int64_t*>(actions), {n}, torch::kInt64).clone();
            out = t->agent->getActionAndValueDiscrete(x, act);
        } else {
            int64_t sumK = t->agent->m_actionSpaceSum;
            torch::Tensor mk = torch::from_blob(const_cast<unsigned char*>(mask), {n, sumK}, torch::kUInt8).clone();
            torch::Tensor act = torch::from_blob(const_cast<int64_t*>(actions), {n, heads}, torch::kInt64).clone().t().contiguous();
            out = t->agent->getActionAndValueMasked(x, mk, act);
        }
        torch::Tensor logratio = out.logprob - b_logp;                       // :585
        torch::Tensor ratio = logratio.exp();                                // :586
        float clipfrac;
        torch::Tensor approx_kl;
        {
            torch::NoGradGuard ng;                                           // :343-356
            clipfrac = static_cast<float>(torch::gt((ratio - 1.0).abs(), clip_coef).count_nonzero().item<int64_t>()) /
                       static_cast<float>(n);
            approx_kl = ((ratio - 1) - logratio).mean();
        }
        if (norm_adv) mb_adv = (mb_adv - mb_adv.mean()) / (mb_adv.std() + 1e-8f);   // :591-594
        torch::Tensor pg1 = -mb_adv * ratio;                                 // :597-599
        torch::Tensor pg2 = -mb_adv * torch::clamp(ratio, 1 - clip_coef, 1 + clip_coef);
        torch::Tensor pg_loss = torch::max(pg1, pg2).mean();
        torch::Tensor newvalue = out.value.view(-1);                         // :602
        torch::Tensor v_loss;
        if (clip_vloss) {                                                    // :603-619
            torch::Tensor vu = (newvalue - b_ret) * (newvalue - b_ret);
            torch::Tensor vc = b_val + torch::clamp(newvalue - b_val, -clip_coef, clip_coef);
            torch::Tensor vcl = (vc - b_ret) * (vc - b_ret);
            v_loss = 0.5f * torch::max(vu, vcl).mean();
        } else {                                                             // :621-625
            v_loss = 0.5f * ((newvalue - b_ret) * (newvalue - b_ret)).mean();
        }
        torch::Tensor entropy_loss = out.entropy.mean();                     // :628
        torch::Tensor loss = pg_loss - ent_coef * entropy_loss + v_loss * vf_coef;  // :631
        t->opt->zero_grad();                                                 // :637-641
        loss.backward();
        if (grads_out) ref_trainer_get_grads(h, grads_out);
        double gn = torch::nn::utils::clip_grad_norm_(t->agent->parameters(), max_grad_norm);
        t->opt->step();
        diag[0] = pg_loss.item<float>();
        diag[1] = v_loss.item<float>();
        diag[2] = entropy_loss.item<float>();
        diag[3] = loss.item<float>();
        diag[4] = approx_kl.item<float>();
        diag[5] = clipfrac;
        diag[6] = static_cast<float>(gn);
        diag[7] = lr;
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

// ---- GAE through PPO_Discrete::calcAdvantage ------------------------------------------
// Builds a PPO_Discrete in a scratch cwd (its ctor reads ./PPOConfig.toml), fills the public
// rollout buffers, swaps the critic for an identity Linear(1,1) so that
// getValue(next_obs=[N,1]) == injected next_value, then calls the reference method.
int ref_gae(int T, int N, const float* rewards, const float* values, const float* dones,
            const float* next_value, const float* next_done, float gamma, float lam, int use_gae,
            float* returns_out, float* adv_out) {
    try {
        ScratchCwd scratch;
        {
            std::ofstream f(std::string(scratch.tmp) + "/PPOConfig.toml");
            f.precision(9);
            f << "[environment]\nobs_size = 4\naction_size = 2\n[general]\nuse_cuda = false\n[ppo]\n"
              << "num_envs = " << N << "\nnum_steps = " << T << "\nnum_minibatches = 1\n"
              << "gamma = " << gamma << "\ngae_lambda = " << lam << "\nuse_gae = " << (use_gae ? "true" : "false") << "\n";
        }
        scratch.enter();
        int rc = 0;
        {
            PPO_Discrete algo;
            algo.m_gamma = gamma;          // avoid the toml text round-trip
            algo.m_gae_lambda = lam;
            algo.m_rewards = f32(rewards, {T, N});
            algo.m_values = f32(values, {T, N});
            algo.m_dones = f32(dones, {T, N});
            torch::nn::Linear ident(1, 1);
            {
                torch::NoGradGuard ng;
                ident->weight.fill_(1.0f);
                ident->bias.fill_(0.0f);
            }
            algo.m_agent->m_Critic = torch::nn::Sequential(ident);
            torch::Tensor nv = f32(next_value, {N, 1});
            torch::Tensor nd = f32(next_done, {N});
            auto ra = algo.calcAdvantage(nv, nd);
            copy_out(ra[0], returns_out);
            copy_out(ra[1], adv_out);
        }
        return rc;
    } catch (const std::exception& e) { return fail(e); }
}

// ---- vectorised env stepping through PPO_Discrete::initEnvs / stepEnvs -------------------
// Runs `steps` calls of stepEnvs on N CartPoles (all seeded with `seed`, reference quirk Q4),
// with the given action stream [steps,N].  Outputs obs [steps+1,N,4] (row 0 = initEnvs),
// reward [steps,N], done [steps,N].  Exercises truncation + auto-reset (PPO_Discrete.cpp:443-458).
int ref_cartpole_vec_rollout(int N, int steps, int64_t seed, int max_episode_steps,
                             const int64_t* actions, float* obs_out, float* reward_out, int* done_out) {
    try {
        ScratchCwd scratch;
        {
            std::ofstream f(std::string(scratch.tmp) + "/PPOConfig.toml");
            f << "[environment]\nobs_size = 4\naction_size = 2\nmax_episode_steps = " << max_episode_steps
              << "\n[general]\nuse_cuda = false\nseed = " << seed << "\n[ppo]\nnum_envs = " << N
              << "\nnum_steps = " << steps << "\nnum_minibatches = 1\n";
        }
        scratch.enter();
        {
            PPO_Discrete algo;
            algo.m_threadPool->start();
            torch::Tensor obs = algo.initEnvs();
            copy_out(obs, obs_out);
            for (int s = 0; s < steps; s++) {
                torch::Tensor act = torch::from_blob(const_cast<int64_t*>(actions + (size_t)s * N), {N}, torch::kInt64).clone();
                auto [o, r, d] = algo.stepEnvs(act);
                copy_out(o, obs_out + (size_t)(s + 1) * N * 4);
                copy_out(r.view(-1), reward_out + (size_t)s * N);
                torch::Tensor di = d.view(-1).to(torch::kInt32).contiguous();
                std::memcpy(done_out + (size_t)s * N, di.data_ptr<int>(), sizeof(int) * N);
            }
            algo.m_threadPool->stop();
        }
        return 0;
    } catch (const std::exception& e) { return fail(e); }
}

}  // extern "C"
