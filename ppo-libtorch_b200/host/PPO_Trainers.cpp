// PPO_Trainers.cpp — see PPO_Trainers.h.  Torch-free: LibTorch is only behind Checkpoint.h.
#include "PPO_Trainers.h"

#include <filesystem>
#include <iomanip>
#include <iostream>
#include <stdexcept>

#include <cuda_runtime_api.h>

#include "Checkpoint.h"

namespace fs = std::filesystem;
using namespace ppo_host;

static void check(int rc, const char* what) {
    if (rc != PPO_OK) throw std::runtime_error(std::string(what) + ": " + ppo_last_error());
}

PPOTrainerBase::PPOTrainerBase(bool multidiscrete, const std::string& config_path)
    : m_multidiscrete(multidiscrete), m_config_path(config_path) {
    getArgs();
}

void PPOTrainerBase::getArgs() {
    Config c;
    if (m_multidiscrete) c.max_episode_steps = 200;   // PPO_MultiDiscrete.cpp:31
    if (load_config(m_config_path, c) == 2 && m_multidiscrete) c.max_episode_steps = 200;
    m_config = c;
    m_obs_size = c.obs_size; m_action_size = c.action_size; m_learning_rate = c.learning_rate; m_seed = c.seed;
    m_total_timesteps = c.total_timesteps; m_use_cuda = c.use_cuda; m_torch_deterministic = c.torch_deterministic;
    m_num_envs = c.num_envs; m_num_steps = c.num_steps; m_anneal_lr = c.anneal_lr; m_use_gae = c.use_gae;
    m_gamma = c.gamma; m_gae_lambda = c.gae_lambda; m_num_minibatches = c.num_minibatches;
    m_update_epochs = c.update_epochs; m_norm_adv = c.norm_adv; m_clip_coef = c.clip_coef; m_clip_vloss = c.clip_vloss;
    m_ent_coef = c.ent_coef; m_vf_coef = c.vf_coef; m_max_grad_norm = c.max_grad_norm;
    m_checkpoint_updates = c.checkpoint_updates; m_max_episode_steps = c.max_episode_steps;
    m_batch_size = c.batch_size; m_minibatch_size = c.minibatch_size;
}

void PPOTrainerBase::build() {
    if (!m_use_cuda) throw std::runtime_error("use_cuda = false: the B200 path has no CPU fallback (run the reference's CPU path)");
    if (ppo_device_count() <= 0) throw std::runtime_error("no CUDA device visible: the B200 path has no CPU fallback");
    std::string env = m_config.env.empty() ? (m_multidiscrete ? "mountaincar" : "cartpole") : m_config.env;
    const int kind = deviceEnvKind(env);
    if (kind != PPO_ENV_HOST && ppo_env_state_dim(kind) != m_obs_size)   // initEnvs' size check (PPO_Discrete.cpp:370-375)
        throw std::runtime_error("obs_size " + std::to_string(m_obs_size) + " does not match the environment's observation (" +
                                 std::to_string(ppo_env_state_dim(kind)) + " floats)");
    // action space: one head of action_size logits (Agent.cpp:21) unless PPO_MultiDiscrete is given the
    // MultiDiscrete nvec the reference keeps commented at Agent.cpp:16 through `[b200] nvec = [..]`
    std::vector<int64_t> nvec = {m_action_size};
    if (m_multidiscrete && !m_config.nvec.empty()) nvec = m_config.nvec;
    if (kind != PPO_ENV_HOST && (int)nvec.size() < ppo_env_num_heads(kind))
        throw std::runtime_error("environment \"" + env + "\" consumes " + std::to_string(ppo_env_num_heads(kind)) +
                                 " sub-actions per step: give PPO_MultiDiscrete its action space with [b200] nvec = [..]");
    if (nvec.size() > PPO_MAX_HEADS) throw std::runtime_error("[b200] nvec: at most " + std::to_string(PPO_MAX_HEADS) + " heads");
    int64_t sumK = 0;
    for (int64_t k : nvec) {
        if (k < 1) throw std::runtime_error("[b200] nvec: every head needs at least one logit");
        sumK += k;
    }
    m_shell = checkpoint_shell_create(m_obs_size, m_config.hidden, sumK, m_learning_rate, m_seed);
    ppo_trainer_cfg c{};
    c.dims.obs_dim = (int32_t)m_obs_size; c.dims.hidden = (int32_t)m_config.hidden; c.dims.heads = (int32_t)nvec.size();
    for (size_t h = 0; h < nvec.size(); h++) c.dims.nvec[h] = (int32_t)nvec[h];
    m_env_kind = kind;
    c.env_kind = kind; c.num_envs = (int32_t)m_num_envs; c.num_steps = (int32_t)m_num_steps;
    c.num_minibatches = (int32_t)m_num_minibatches; c.update_epochs = (int32_t)m_update_epochs;
    c.max_episode_steps = (int32_t)m_max_episode_steps; c.use_gae = m_use_gae; c.gamma = m_gamma; c.gae_lambda = m_gae_lambda;
    c.hyper.clip_coef = m_clip_coef; c.hyper.norm_adv = m_norm_adv; c.hyper.clip_vloss = m_clip_vloss;
    c.hyper.ent_coef = m_ent_coef; c.hyper.vf_coef = m_vf_coef;
    c.hyper.entropy_mode = m_config.exact_entropy ? PPO_ENTROPY_EXACT : PPO_ENTROPY_REFERENCE;
    c.max_grad_norm = m_max_grad_norm; c.beta1 = 0.9f; c.beta2 = 0.999f; c.adam_eps = 1e-5f; c.weight_decay = 1e-2f;
    c.seed = (uint64_t)m_seed; c.world_size = 1; c.rank = 0; c.use_cuda_graph = m_config.use_cuda_graph;
    c.masked = m_multidiscrete;
    m_b200 = ppo_trainer_create(&c);
    if (!m_b200) throw std::runtime_error(std::string("ppo_trainer_create: ") + ppo_last_error());
    std::vector<float> flat((size_t)checkpoint_num_params(m_shell));
    if ((int64_t)flat.size() != ppo_num_params(&c.dims)) throw std::runtime_error("agent shell / device slab size mismatch");
    checkpoint_get_params(m_shell, flat.data());
    check(ppo_trainer_set_params(m_b200, flat.data()), "ppo_trainer_set_params");
    loadPolicyFromCheckpoint();
    check(ppo_trainer_reset_envs(m_b200), "ppo_trainer_reset_envs");
    m_episode_stats = std::make_unique<CircularBuffer>(100);   // PPO_Discrete.cpp:98
}

PPOTrainerBase::~PPOTrainerBase() {
    if (m_b200) ppo_trainer_destroy(m_b200);
    if (m_shell) checkpoint_shell_destroy(m_shell);
}

std::vector<float> PPOTrainerBase::getParameters() const {
    std::vector<float> flat((size_t)checkpoint_num_params(m_shell));
    check(ppo_trainer_get_params(m_b200, flat.data()), "ppo_trainer_get_params");
    return flat;
}

static std::string newest_file(const fs::path& dir) {   // newest by last_write_time (:801-806)
    std::string best;
    fs::file_time_type t{};
    for (const auto& e : fs::directory_iterator(dir))
        if (best.empty() || fs::last_write_time(e) > t) { best = e.path().string(); t = fs::last_write_time(e); }
    return best;
}

void PPOTrainerBase::loadPolicyFromCheckpoint() {
    const fs::path model_dir = "./ModelCheckpoints/", optim_dir = "./OptimizerCheckpoints/";
    if (!fs::exists(model_dir) || !fs::exists(optim_dir)) {
        std::cout << "No previous model checkpoint found at " << model_dir << ", initializing new agent!" << std::endl;
        return;
    }
    const size_t n = (size_t)checkpoint_num_params(m_shell);
    const std::string agent_file = newest_file(model_dir);
    if (agent_file.empty()) {
        std::cout << "No previous model checkpoint found at " << model_dir << ", initializing new agent!" << std::endl;
    } else {
        std::cout << "Loading model " << agent_file << "..." << std::endl;
        m_global_step = step_from_filename(agent_file, "PPO_Agent_");   // 64-bit (the reference's stoi overflows at 2^31)
        std::cout << "Continuing training from step " << m_global_step << std::endl;
        checkpoint_load_agent(m_shell, agent_file);
        std::vector<float> flat(n);
        checkpoint_get_params(m_shell, flat.data());
        check(ppo_trainer_set_params(m_b200, flat.data()), "ppo_trainer_set_params");
    }
    const std::string optim_file = newest_file(optim_dir);
    if (optim_file.empty()) {
        std::cout << "No previous optimizer checkpoint found at " << model_dir << ", initializing new optimizer!" << std::endl;
    } else {
        std::cout << "Loading optimizer " << optim_file << "..." << std::endl;
        checkpoint_load_optimizer(m_shell, optim_file);
        std::vector<float> m(n), v(n);
        int64_t step = 0;
        double lr = 0.0;
        if (checkpoint_get_opt_state(m_shell, m.data(), v.data(), &step, &lr)) {
            check(ppo_trainer_set_opt_state(m_b200, m.data(), v.data(), step), "ppo_trainer_set_opt_state");
            m_opt_steps = step;
            // without annealing the reference never rewrites the optimizer's lr, so a resumed run keeps the archive's
            if (!m_anneal_lr && lr > 0.0) m_learning_rate = (float)lr;
        }
    }
}

void PPOTrainerBase::saveCheckpoint(const std::string& agent_file, const std::string& optimizer_file) {
    const size_t n = (size_t)checkpoint_num_params(m_shell);
    std::vector<float> p(n), m(n), v(n);
    int64_t step = 0;
    check(ppo_trainer_get_params(m_b200, p.data()), "ppo_trainer_get_params");
    check(ppo_trainer_get_opt_state(m_b200, m.data(), v.data(), &step), "ppo_trainer_get_opt_state");
    checkpoint_set_params(m_shell, p.data());
    checkpoint_set_opt_state(m_shell, m.data(), v.data(), step, m_last_lr >= 0.0f ? m_last_lr : m_learning_rate);
    checkpoint_save_agent(m_shell, agent_file);
    checkpoint_save_optimizer(m_shell, optimizer_file);
}

// the environment registry (csrc/env.cuh: PPO_ENV_LIST) resolves the name; the reference fixes the env by the element type
// of m_envs (PPO/PPO_Discrete.h:106)
int PPOTrainerBase::deviceEnvKind(const std::string& env) const {
    const int kind = ppo_env_kind_from_name(env.c_str());
    if (kind < 0) {
        std::string known;
        for (int k = 0; ppo_env_name(k); k++) known += std::string(k ? ", " : "") + "\"" + ppo_env_name(k) + "\"";
        throw std::runtime_error("[b200] env = \"" + env + "\": not a registered environment (" + known + ")");
    }
    return kind;
}

void* PPOTrainerBase::ensureStream() {
    if (!m_stream) {   // graph capture cannot run on the legacy default stream
        cudaStream_t st = nullptr;
        if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) throw std::runtime_error("cudaStreamCreate failed");
        m_stream = st;
    }
    return m_stream;
}

ppo_iter_result PPOTrainerBase::iterate(float lr) {
    ppo_iter_result r{};
    check(ppo_trainer_iterate(m_b200, lr, &r, ensureStream()), "ppo_trainer_iterate");
    m_last_lr = lr;   // what the reference's optimizer holds when it is saved (lr_now, PPO_Discrete.cpp:514-518)
    return r;
}

template <typename T> void PPOTrainerBase::printElement(T t, int width) {   // PPO_Discrete.h:111-114
    std::cout << std::left << std::setw(width) << std::setfill(' ') << t << std::right << "|\n";
}

// PPO_Discrete.cpp:700-774, same rows, widths and stream formatting state sequence
void PPOTrainerBase::printPPOResults(int64_t update, int64_t global_step, std::chrono::milliseconds fps,
                                     std::chrono::milliseconds time_elapsed, const ppo_iter_result& r) {
    const double fps_s = fps.count() / 1000.0;
    const int64_t fps_v = fps_s > 0.0 ? (int64_t)((double)m_batch_size / fps_s) : 0;
    if (update == 1) {
        std::cout << "---------------------------------\n";
        if (!m_episode_stats->empty()) {
            std::cout << "| rollout/           |          |\n";
            std::cout << std::setprecision(1) << std::defaultfloat << "|    ep_len_mean     | ";
            printElement(m_episode_stats->avgLength(), 9);
            std::cout << std::setprecision(5) << "|    ep_rew_mean     | ";
            printElement(m_episode_stats->avgReward(), 9);
        }
        std::cout << "| time/              |          |\n";
        std::cout << "|    fps             | ";
        printElement(fps_v, 9);
        std::cout << "|    iterations      | ";
        printElement(update, 9);
        std::cout << "|    time_elapsed    | ";
        printElement((int64_t)(time_elapsed.count() / 1000.0), 9);
        std::cout << "|    total_timesteps | ";
        printElement(global_step, 9);
        std::cout << "---------------------------------" << std::endl << std::endl;
        return;
    }
    std::cout << "------------------------------------------\n";
    if (!m_episode_stats->empty()) {
        std::cout << "| rollout/                |              |\n";
        std::cout << std::setprecision(2) << std::fixed << "|    ep_len_mean          | ";
        printElement(m_episode_stats->avgLength(), 13);
        std::cout << std::setprecision(8) << "|    ep_rew_mean          | ";
        printElement(m_episode_stats->avgReward(), 13);
    }
    std::cout << "| time/                   |              |\n";
    std::cout << "|    fps                  | ";
    printElement(fps_v, 13);
    std::cout << "|    iterations           | ";
    printElement(update, 13);
    std::cout << "|    time_elapsed         | ";
    printElement((int64_t)(time_elapsed.count() / 1000.0), 13);
    std::cout << "|    total_timesteps      | ";
    printElement(global_step, 13);
    std::cout << "| train/                  |              |\n";
    std::cout << std::setprecision(9) << "|    approx_kl            | ";
    printElement(r.approx_kl, 13);
    std::cout << "|    clip_fraction        | ";
    printElement(vector_mean(m_clipfracs), 13);
    std::cout << "|    clip_range           | ";
    printElement(m_clip_coef, 13);
    std::cout << "|    entropy_loss         | " << std::fixed;
    printElement(r.entropy, 13);
    std::cout << "|    explained_variance   | ";
    printElement(r.explained_variance, 13);
    std::cout << "|    learning_rate        | " << std::defaultfloat << std::setprecision(6);
    printElement(r.lr, 13);
    std::cout << "|    loss                 | " << std::fixed << std::setprecision(9);
    printElement(r.loss, 13);
    std::cout << "|    n_updates            | ";
    printElement(update * m_update_epochs, 13);
    std::cout << "|    policy_gradient_loss | ";
    printElement(r.pg_loss, 13);
    std::cout << "|    value_loss           | ";
    printElement(r.v_loss, 13);
    std::cout << "------------------------------------------" << std::endl << std::endl;
}

// PPO_Discrete::train (PPO_Discrete.cpp:485-690)
void PPOTrainerBase::train() {
    const auto start_time = std::chrono::steady_clock::now();
    auto update_time = start_time;
    uint64_t global_step = m_global_step;
    const int64_t num_updates = m_batch_size > 0 ? ((int64_t)m_total_timesteps - (int64_t)global_step) / m_batch_size : 0;   // :496
    for (int64_t update = 1; update <= num_updates; ++update) {
        float lr = m_learning_rate;
        if (m_anneal_lr) {                              // :514-518
            const double frac = 1.0 - ((double)update - 1.0) / (double)num_updates;
            lr = (float)(frac * (double)m_learning_rate);
        }
        const ppo_iter_result r = iterate(lr);
        global_step += (uint64_t)m_batch_size;
        m_clipfracs.assign(1, r.clipfrac);              // already the mean over the update's minibatches (Utils.cpp:5-14)
        // Episode statistics (PPO_Discrete.cpp:474-480): the rollout kernel logs every finished episode at its (step, env)
        // slot, ppo_episode_tail compacts the newest 100 on the device, and they are pushed oldest first - the ring ends
        // up holding exactly what the reference's per-step pushes would have left in it.
        if (!pushesEpisodeStats()) {
            int32_t lens[PPO_EPISODE_TAIL];
            float rets[PPO_EPISODE_TAIL];
            const int n = ppo_trainer_last_episodes(m_b200, PPO_EPISODE_TAIL, lens, rets);
            for (int i = 0; i < n; i++) m_episode_stats->add(rets[i], (int64_t)lens[i]);
        }
        const auto now = std::chrono::steady_clock::now();
        const auto fps = std::chrono::duration_cast<std::chrono::milliseconds>(now - update_time);
        const auto elapsed = std::chrono::duration_cast<std::chrono::milliseconds>(now - start_time);
        printPPOResults(update, (int64_t)global_step, fps, elapsed, r);
        update_time = std::chrono::steady_clock::now();
        if (m_checkpoint_updates > 0 && update % m_checkpoint_updates == 0) {   // :662-673
            fs::create_directories("./ModelCheckpoints/");
            fs::create_directories("./OptimizerCheckpoints/");
            const std::string a = "./ModelCheckpoints/PPO_Agent_" + std::to_string(global_step) + "_steps.pt";
            const std::string o = "./OptimizerCheckpoints/PPO_Optimizer_" + std::to_string(global_step) + "_steps.pt";
            std::cout << "Saving model checkpoint to " << a << "..." << std::endl;
            std::cout << "Saving optimizer checkpoint to " << o << "..." << std::endl;
            saveCheckpoint(a, o);
        }
    }
    m_global_step = global_step;
    fs::create_directories("./Models/");                // :678-685
    const std::string a = "./Models/PPO_Agent_" + std::to_string(m_total_timesteps) + "_steps.pt";
    const std::string o = "./Models/PPO_Optimizer_" + std::to_string(m_total_timesteps) + "_steps.pt";
    std::cout << "Saving model " << a << "..." << std::endl;
    std::cout << "Saving optimizer " << o << "..." << std::endl;
    saveCheckpoint(a, o);
}
