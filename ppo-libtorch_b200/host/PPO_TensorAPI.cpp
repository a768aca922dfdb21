// PPO_TensorAPI.cpp — the tensor-level methods of PPO_Discrete / PPO_MultiDiscrete (PPO_Discrete.h) over the C ABI.
#include <cuda_runtime_api.h>

#include <stdexcept>

#include "PPO_Discrete.h"

using ppo_host::check_rc;

torch::Tensor PPOTensorTrainer::arenaView(const char* name, c10::ScalarType dtype, std::vector<int64_t> shape) const {
    int64_t bytes = 0;
    void* p = ppo_trainer_buffer(handle(), name, &bytes);
    if (!p) return torch::Tensor();
    int64_t n = 1;
    for (int64_t d : shape) n *= d;
    if (n * (int64_t)c10::elementSize(dtype) > bytes) throw std::runtime_error(std::string("arena buffer too small: ") + name);
    return torch::from_blob(p, shape, torch::TensorOptions().dtype(dtype).device(*m_device));
}

void PPOTensorTrainer::bindTensors() {
    int dev = 0;
    cudaGetDevice(&dev);
    m_device = std::make_shared<torch::Device>(torch::kCUDA, (c10::DeviceIndex)dev);   // PPO_Discrete.cpp:59-73
    const int64_t T = m_num_steps, N = m_num_envs, O = m_obs_size;
    std::vector<int64_t> nvec = m_config.nvec.empty() || !multidiscrete() ? std::vector<int64_t>{m_action_size} : m_config.nvec;
    int64_t K = 0;
    for (int64_t k : nvec) K += k;
    // storage (:90-95): views of the arena the rollout kernels write
    m_obs = arenaView("obs", torch::kFloat32, {T, N, O});
    m_actions = arenaView("actions", torch::kInt32, {T, N, (int64_t)nvec.size()});
    m_logprobs = arenaView("logprobs", torch::kFloat32, {T, N});
    m_rewards = arenaView("rewards", torch::kFloat32, {T, N});
    m_dones = arenaView("dones", torch::kFloat32, {T, N});
    m_values = arenaView("values", torch::kFloat32, {T, N});
    if (multidiscrete()) m_action_masks = arenaView("masks", torch::kUInt8, {T, N, K});
    // the agent: same class as the reference's, its parameters aliasing the trainer's slab (current weights included)
    torch::manual_seed((uint64_t)m_seed);
    m_agent = std::make_shared<Agent>(m_obs_size, m_action_size, m_device, m_config.hidden,
                                      multidiscrete() ? m_config.nvec : std::vector<int64_t>{});
    m_agent->m_exact_entropy = m_config.exact_entropy;
    int64_t pbytes = 0;
    void* pp = ppo_trainer_buffer(handle(), "params", &pbytes);
    torch::Tensor slab = torch::from_blob(pp, {pbytes / 4}, torch::TensorOptions().dtype(torch::kFloat32).device(*m_device));
    {   // keep the trainer's weights: alias WITHOUT overwriting them with the fresh agent's init
        torch::NoGradGuard ng;
        torch::Tensor keep = slab.clone();
        m_agent->aliasParameters(slab);
        slab.copy_(keep);
    }
}

AgentOutput PPOTensorTrainer::computeActionLogic(const torch::Tensor& next_obs) const {   // PPO_Discrete.cpp:262-271
    torch::NoGradGuard no_grad;
    return m_agent->getActionAndValueDiscrete(next_obs);
}
AgentOutput PPOTensorTrainer::computeActionLogic(const torch::Tensor& next_obs, const torch::Tensor& action_mask,
                                                 const torch::Tensor& input_action) const {   // PPO_MultiDiscrete.cpp:270-281
    torch::NoGradGuard no_grad;
    return m_agent->getActionAndValueMasked(next_obs, action_mask, input_action);
}

// PPO_Discrete.cpp:274-331: bootstrap value of next_obs, then GAE (or plain returns) over m_rewards / m_values / m_dones
std::array<torch::Tensor, 2> PPOTensorTrainer::calcAdvantage(const torch::Tensor& next_obs, const torch::Tensor& next_done) const {
    torch::NoGradGuard no_grad;
    const int64_t T = m_num_steps, N = m_num_envs;
    torch::Tensor next_value = m_agent->getValue(next_obs).reshape({N}).contiguous();
    torch::Tensor nd = next_done.to(*m_device).to(torch::kFloat32).reshape({N}).contiguous();
    torch::Tensor advantages = torch::empty({T, N}, m_rewards.options()), returns = torch::empty({T, N}, m_rewards.options());
    check_rc(ppo_gae((int)T, (int)N, m_rewards.data_ptr<float>(), m_values.data_ptr<float>(), m_dones.data_ptr<float>(),
                     next_value.data_ptr<float>(), nd.data_ptr<float>(), m_gamma, m_gae_lambda, m_use_gae ? 1 : 0,
                     advantages.data_ptr<float>(), returns.data_ptr<float>(), ppo_host::stream_of(advantages)),
             "ppo_gae");
    return {returns, advantages};
}

// PPO_Discrete.cpp:343-356.  train() takes both numbers from the fused update kernel's diagnostics; this method serves
// callers that hold ratio / logratio tensors of their own (plain LibTorch reductions, one host sync like the reference).
torch::Tensor PPOTensorTrainer::getApproxKLAndClippedObj(const torch::Tensor& ratio, const torch::Tensor& logratio) {
    torch::NoGradGuard no_grad;
    torch::Tensor clippedRatios = torch::gt((ratio - 1.0).abs(), m_clip_coef);
    m_clipfracs.push_back(static_cast<float>(clippedRatios.count_nonzero().item<int64_t>()) /
                          static_cast<float>(clippedRatios.size(0)));
    return ((ratio - 1) - logratio).mean();
}

// PPO_Discrete.cpp:365-402: (re)initialise every env, return the observations [N, O]
torch::Tensor PPOTensorTrainer::initEnvs() const {
    check_rc(ppo_trainer_reset_envs(handle()), "ppo_trainer_reset_envs");
    torch::Tensor soa = arenaView("state", torch::kFloat32, {m_obs_size, m_num_envs});
    return soa.t().contiguous();
}
torch::Tensor PPOTensorTrainer::initEnvs(const torch::Tensor& action_mask) {   // PPO_MultiDiscrete.cpp:384-432
    torch::Tensor obs = initEnvs();
    if (action_mask.defined() && action_mask.numel()) action_mask.fill_(1);    // every env of the built-ins allows every action
    return obs;
}

// PPO_Discrete.cpp:413-483: one vectorised transition with truncation and auto-reset; finished episodes go into
// m_episode_stats in env order (:474-480).  Returns obs [N, O], reward [N, 1], done int32 [N, 1].
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> PPOTensorTrainer::stepEnvs(const torch::Tensor& action) {
    const int64_t N = m_num_envs, O = m_obs_size;
    torch::Tensor act = action.to(*m_device).to(torch::kInt32).reshape({N, -1}).select(1, 0).contiguous();
    auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(*m_device);
    torch::Tensor obs = torch::empty({N, O}, f32), reward = torch::empty({N, 1}, f32), done = torch::empty({N, 1}, f32);
    torch::Tensor fin_len = torch::empty({N}, f32.dtype(torch::kInt32)), fin_ret = torch::empty({N}, f32);
    int64_t b = 0;
    float* state = static_cast<float*>(ppo_trainer_buffer(handle(), "state", &b));
    int32_t* ep_len = static_cast<int32_t*>(ppo_trainer_buffer(handle(), "ep_len", &b));
    float* ep_ret = static_cast<float*>(ppo_trainer_buffer(handle(), "ep_ret", &b));
    check_rc(ppo_env_step(envKind(), (int)N, state, ep_len, ep_ret, act.data_ptr<int32_t>(), (int)m_max_episode_steps, nullptr,
                          (uint64_t)m_seed ^ 0x5745505354455053ull, m_step_calls++, obs.data_ptr<float>(),
                          reward.data_ptr<float>(), done.data_ptr<float>(), fin_len.data_ptr<int32_t>(),
                          fin_ret.data_ptr<float>(), ppo_host::stream_of(obs)),
             "ppo_env_step");
    torch::Tensor fl = fin_len.cpu(), fr = fin_ret.cpu();   // small N on this per-step path: one D2H like the reference's loop
    const int32_t* pl = fl.data_ptr<int32_t>();
    const float* pr = fr.data_ptr<float>();
    for (int64_t i = 0; i < N; i++)
        if (pl[i] > 0) m_episode_stats->add(pr[i], static_cast<int64_t>(pl[i]));
    return {obs, reward, done.to(torch::kInt32)};
}
