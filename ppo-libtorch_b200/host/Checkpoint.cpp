// Checkpoint.cpp — see Checkpoint.h.  The only LibTorch translation unit of the host facade.
#include "Checkpoint.h"

#include <torch/torch.h>

#include <cmath>
#include <cstring>

namespace ppo_host {

namespace {
// Agent::ppoLayerInit (PPO/Agent.cpp:91-99): orthogonal weight with the given gain, constant bias
torch::nn::Linear layer_init(torch::nn::Linear l, double gain, double bias = 0.0) {
    torch::NoGradGuard ng;
    torch::nn::init::orthogonal_(l->weight, gain);
    torch::nn::init::constant_(l->bias, bias);
    return l;
}
struct AgentShellImpl : torch::nn::Module {
    torch::nn::Sequential m_Critic{nullptr}, m_Actor{nullptr};
    AgentShellImpl(int64_t obs, int64_t hidden, int64_t actions) {
        const double g = std::sqrt(2.0);
        m_Critic = torch::nn::Sequential({{"criticInputLayer", layer_init(torch::nn::Linear(obs, hidden), g)},
                                          {"Tanh1", torch::nn::Tanh()},
                                          {"criticMiddleLayer", layer_init(torch::nn::Linear(hidden, hidden), g)},
                                          {"Tanh2", torch::nn::Tanh()},
                                          {"criticOutputLayer", layer_init(torch::nn::Linear(hidden, 1), 1.0)}});
        m_Actor = torch::nn::Sequential({{"actorInputLayer", layer_init(torch::nn::Linear(obs, hidden), g)},
                                         {"Tanh1", torch::nn::Tanh()},
                                         {"actorMiddleLayer", layer_init(torch::nn::Linear(hidden, hidden), g)},
                                         {"Tanh2", torch::nn::Tanh()},
                                         {"actorOutputLayer", layer_init(torch::nn::Linear(hidden, actions), 0.01)}});
        m_Critic = register_module("m_Critic", m_Critic);
        m_Actor = register_module("m_Actor", m_Actor);
    }
};
TORCH_MODULE(AgentShell);
}  // namespace

struct CheckpointShell {
    AgentShell agent{nullptr};
    std::unique_ptr<torch::optim::AdamW> opt;
};

CheckpointShell* checkpoint_shell_create(int64_t obs, int64_t hidden, int64_t actions, double lr, int64_t seed) {
    torch::manual_seed((uint64_t)seed);   // PPO_Discrete.cpp:44
    auto* s = new CheckpointShell();
    s->agent = AgentShell(obs, hidden, actions);
    s->opt = std::make_unique<torch::optim::AdamW>(s->agent->parameters(), torch::optim::AdamWOptions(lr).eps(1e-5f));   // :76-78 (float literal, like the reference)
    return s;
}
void checkpoint_shell_destroy(CheckpointShell* s) { delete s; }

int64_t checkpoint_num_params(const CheckpointShell* s) {
    int64_t n = 0;
    for (const auto& p : s->agent->parameters()) n += p.numel();
    return n;
}
void checkpoint_get_params(const CheckpointShell* s, float* flat) {
    for (const auto& p : s->agent->parameters()) {
        auto v = p.detach().to(torch::kCPU, torch::kFloat32).contiguous();
        std::memcpy(flat, v.data_ptr<float>(), (size_t)v.numel() * 4);
        flat += v.numel();
    }
}
void checkpoint_set_params(CheckpointShell* s, const float* flat) {
    torch::NoGradGuard ng;
    for (auto& p : s->agent->parameters()) {
        p.copy_(torch::from_blob(const_cast<float*>(flat), p.sizes(), torch::kFloat32));
        flat += p.numel();
    }
}
void checkpoint_set_opt_state(CheckpointShell* s, const float* m, const float* v, int64_t step, double lr) {
    auto& groups = s->opt->param_groups();
    static_cast<torch::optim::AdamWOptions&>(groups[0].options()).lr(lr);
    auto& state = s->opt->state();
    for (auto& p : groups[0].params()) {
        auto st = std::make_unique<torch::optim::AdamWParamState>();
        st->step(step);
        st->exp_avg(torch::from_blob(const_cast<float*>(m), p.sizes(), torch::kFloat32).clone());
        st->exp_avg_sq(torch::from_blob(const_cast<float*>(v), p.sizes(), torch::kFloat32).clone());
        state[p.unsafeGetTensorImpl()] = std::move(st);
        m += p.numel();
        v += p.numel();
    }
}
bool checkpoint_get_opt_state(const CheckpointShell* s, float* m, float* v, int64_t* step, double* lr) {
    auto& groups = s->opt->param_groups();
    *lr = static_cast<torch::optim::AdamWOptions&>(groups[0].options()).lr();
    auto& state = s->opt->state();
    *step = 0;
    bool any = false;
    for (auto& p : groups[0].params()) {
        auto it = state.find(p.unsafeGetTensorImpl());
        const int64_t n = p.numel();
        if (it == state.end()) {   // a parameter that never stepped: zero moments (fresh optimiser)
            std::memset(m, 0, (size_t)n * 4);
            std::memset(v, 0, (size_t)n * 4);
        } else {
            auto& st = static_cast<torch::optim::AdamWParamState&>(*it->second);
            auto a = st.exp_avg().to(torch::kCPU, torch::kFloat32).contiguous();
            auto b = st.exp_avg_sq().to(torch::kCPU, torch::kFloat32).contiguous();
            std::memcpy(m, a.data_ptr<float>(), (size_t)n * 4);
            std::memcpy(v, b.data_ptr<float>(), (size_t)n * 4);
            *step = st.step();
            any = true;
        }
        m += n;
        v += n;
    }
    return any;
}
void checkpoint_save_agent(const CheckpointShell* s, const std::string& file) { torch::save(s->agent, file); }
void checkpoint_save_optimizer(const CheckpointShell* s, const std::string& file) { torch::save(*s->opt, file); }
void checkpoint_load_agent(CheckpointShell* s, const std::string& file) { torch::load(s->agent, file); }
void checkpoint_load_optimizer(CheckpointShell* s, const std::string& file) { torch::load(*s->opt, file); }

}  // namespace ppo_host
