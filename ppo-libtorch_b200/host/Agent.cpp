// Agent.cpp — see Agent.h.
#include "Agent.h"

#include <iostream>
#include <numeric>
#include <stdexcept>

#include "../../include/ppo_b200.h"

using ppo_host::check_rc;
using ppo_host::stream_of;

Agent::Agent(int64_t obsSize, int64_t actionSize, std::shared_ptr<torch::Device> device, int64_t hidden,
             std::vector<int64_t> actionSpace) {
    m_actionSpace = actionSpace.empty() ? std::vector<int64_t>{actionSize} : actionSpace;   // Agent.cpp:16
    m_actionSpaceSum = std::accumulate(m_actionSpace.begin(), m_actionSpace.end(), static_cast<int64_t>(0));
    m_actionSpaceSize = static_cast<int64_t>(m_actionSpace.size());
    m_obsSize = obsSize;
    m_hidden = hidden;
    // same construction order as Agent.cpp:25-37 => the same draws from torch's generator for the orthogonal init
    torch::nn::Linear criticInputLayer = ppoLayerInit(torch::nn::Linear(obsSize, hidden));
    torch::nn::Linear criticMiddleLayer = ppoLayerInit(torch::nn::Linear(hidden, hidden));
    torch::nn::Linear criticOutputLayer = ppoLayerInit(torch::nn::Linear(hidden, 1), 1.0);
    torch::nn::Linear actorInputLayer = ppoLayerInit(torch::nn::Linear(obsSize, hidden));
    torch::nn::Linear actorMiddleLayer = ppoLayerInit(torch::nn::Linear(hidden, hidden));
    torch::nn::Linear actorOutputLayer = ppoLayerInit(torch::nn::Linear(hidden, m_actionSpaceSum), 0.01);
    torch::nn::Sequential critic({{"criticInputLayer", criticInputLayer},
                                  {"Tanh1", torch::nn::Tanh()},
                                  {"criticMiddleLayer", criticMiddleLayer},
                                  {"Tanh2", torch::nn::Tanh()},
                                  {"criticOutputLayer", criticOutputLayer}});
    torch::nn::Sequential action_mean({{"actorInputLayer", actorInputLayer},
                                       {"Tanh1", torch::nn::Tanh()},
                                       {"actorMiddleLayer", actorMiddleLayer},
                                       {"Tanh2", torch::nn::Tanh()},
                                       {"actorOutputLayer", actorOutputLayer}});
    m_Critic = critic;
    m_Actor = action_mean;
    m_Critic = register_module("m_Critic", m_Critic);
    m_Actor = register_module("m_Actor", m_Actor);
    m_device = device;
}
Agent::~Agent() {}

torch::nn::Linear Agent::ppoLayerInit(torch::nn::Linear layer, const double stdDev, const double bias_const) {
    torch::NoGradGuard noGrad;                                   // Agent.cpp:91-99
    torch::nn::init::orthogonal_(layer->weight, stdDev);
    torch::nn::init::constant_(layer->bias, bias_const);
    return layer;
}

void Agent::aliasParameters(torch::Tensor flat) {
    torch::NoGradGuard noGrad;
    int64_t total = 0;
    for (const auto& p : parameters()) total += p.numel();
    if (!flat.defined() || flat.numel() != total || flat.scalar_type() != torch::kFloat32 || !flat.is_contiguous())
        throw std::runtime_error("Agent::aliasParameters: the slab must be a contiguous fp32 tensor of the parameter count");
    int64_t off = 0;
    for (auto& p : parameters()) {
        torch::Tensor view = flat.narrow(0, off, p.numel()).view(p.sizes());
        if (view.data_ptr() != p.data_ptr()) {
            view.copy_(p.detach().to(flat.device()));
            p.set_data(view);
        }
        off += p.numel();
    }
    m_flat = flat;
}

torch::Tensor Agent::flatParameters() {
    if (!m_device || !m_device->is_cuda())
        throw std::runtime_error("Agent: the B200 path needs a CUDA device (there is no CPU fallback)");
    bool aliased = m_flat.defined();
    if (aliased) {
        int64_t off = 0;
        for (const auto& p : parameters()) {
            if (p.data_ptr() != static_cast<void*>(m_flat.data_ptr<float>() + off)) { aliased = false; break; }
            off += p.numel();
        }
    }
    if (!aliased) {
        int64_t total = 0;
        for (const auto& p : parameters()) total += p.numel();
        torch::Tensor flat = m_flat.defined() && m_flat.numel() == total
                                 ? m_flat
                                 : torch::empty({total}, torch::TensorOptions().dtype(torch::kFloat32).device(*m_device));
        aliasParameters(flat);
    }
    return m_flat;
}

AgentOutput Agent::forward_fused(const torch::Tensor& x, const torch::Tensor* mask, torch::Tensor action, bool masked_api) {
    if (!x.defined() || !x.is_cuda()) throw std::runtime_error("Agent: observations must be a CUDA tensor");
    torch::Tensor params = flatParameters();
    torch::Tensor obs = x.detach().to(torch::kFloat32).reshape({-1, m_obsSize}).contiguous();
    const int64_t n = obs.size(0), heads = m_actionSpaceSize;
    ppo_dims d{};
    d.obs_dim = (int)m_obsSize; d.hidden = (int)m_hidden; d.heads = (int)heads;
    for (int64_t h = 0; h < heads; h++) d.nvec[h] = (int)m_actionSpace[h];
    auto opt = obs.options();
    torch::Tensor acts = torch::empty({n, heads}, opt.dtype(torch::kInt32));
    torch::Tensor logp = torch::empty({n}, opt), ent = torch::empty({n}, opt), val = torch::empty({n, 1}, opt);
    torch::Tensor ain, m8;
    if (action.defined() && action.numel()) {
        // Discrete: [n] (or [n, 1]); Masked: [heads, n] as the reference indexes action[i] per head (Agent.cpp:160)
        torch::Tensor a = action.to(obs.device()).to(torch::kInt32);
        ain = (masked_api ? a.reshape({heads, n}).t() : a.reshape({n, heads})).contiguous();
    }
    if (mask) m8 = mask->to(obs.device()).to(torch::kBool).to(torch::kUInt8).reshape({n, m_actionSpaceSum}).contiguous();
    uint64_t seed = 0, counter = 0;
    if (!ain.defined()) ppo_host::next_sampling_key(&seed, &counter);
    const int emode = (mask || m_exact_entropy) ? PPO_ENTROPY_EXACT : PPO_ENTROPY_REFERENCE;
    check_rc(ppo_policy_forward(params.data_ptr<float>(), &d, (int)n, obs.data_ptr<float>(),
                                m8.defined() ? m8.data_ptr<uint8_t>() : nullptr,
                                ain.defined() ? ain.data_ptr<int32_t>() : nullptr, nullptr, seed, counter, emode,
                                acts.data_ptr<int32_t>(), logp.data_ptr<float>(), ent.data_ptr<float>(),
                                val.data_ptr<float>(), nullptr, stream_of(obs)),
             "ppo_policy_forward");
    torch::Tensor action_out;
    if (ain.defined()) action_out = masked_api ? action.t() : action;             // Agent.cpp:127 / :169
    else action_out = masked_api ? acts.to(torch::kLong) : acts.to(torch::kLong).reshape({n});
    return {action_out, logp, ent, val};
}

torch::Tensor Agent::getValue(const torch::Tensor& x) {                           // Agent.cpp:107-109
    if (!x.defined() || !x.is_cuda()) throw std::runtime_error("Agent: observations must be a CUDA tensor");
    torch::Tensor params = flatParameters();
    torch::Tensor obs = x.detach().to(torch::kFloat32).reshape({-1, m_obsSize}).contiguous();
    ppo_dims d{};
    d.obs_dim = (int)m_obsSize; d.hidden = (int)m_hidden; d.heads = (int)m_actionSpaceSize;
    for (int64_t h = 0; h < m_actionSpaceSize; h++) d.nvec[h] = (int)m_actionSpace[h];
    torch::Tensor val = torch::empty({obs.size(0), 1}, obs.options());
    check_rc(ppo_value_forward(params.data_ptr<float>(), &d, (int)obs.size(0), obs.data_ptr<float>(), val.data_ptr<float>(),
                               stream_of(obs)),
             "ppo_value_forward");
    return val;
}
AgentOutput Agent::getActionAndValueDiscrete(const torch::Tensor& x, torch::Tensor action) {   // Agent.cpp:117-128
    return forward_fused(x, nullptr, action, false);
}
AgentOutput Agent::getActionAndValueMasked(const torch::Tensor& x, const torch::Tensor& mask, torch::Tensor action) {
    return forward_fused(x, &mask, action, true);                                               // Agent.cpp:137-170
}

void Agent::printAgent() {
    for (auto& p : named_parameters()) std::cout << p.key() << "\n" << p.value() << "\n";
}
