// api_check.cpp — exercises every public method / member of the reference's class surface as re-hosted here (Agent,
// Categorical, CategoricalMasked, PPO_Discrete, PPO_MultiDiscrete) on cuda:0 and checks each result against LibTorch's own
// arithmetic for the same module / tensors (the operations the reference runs: Sequential::forward = addmm + tanh,
// log_softmax, gather; PPO/Agent.cpp:107-170, Distributions/*.cpp).  Prints one "ok <what>" line per check; exits 1 on the
// first failure.  Driven by tests/test_gpu_facade.py; needs ./PPOConfig.toml in the working directory for the trainers.
#include <cfloat>
#include <cmath>
#include <iostream>
#include <stdexcept>

#include "PPO_MultiDiscrete.h"

static void expect(bool cond, const std::string& what) {
    if (!cond) throw std::runtime_error("FAILED " + what);
    std::cout << "ok " << what << std::endl;
}
static double max_abs_diff(const torch::Tensor& a, const torch::Tensor& b) {
    return (a.to(torch::kFloat64).cpu().reshape(-1) - b.to(torch::kFloat64).cpu().reshape(-1)).abs().max().item<double>();
}

static void check_distributions(std::shared_ptr<torch::Device> dev) {
    torch::manual_seed(7);
    const int64_t n = 4096, K = 5;
    torch::Tensor logits = torch::randn({n, K}, torch::TensorOptions().device(*dev)) * 2.0;
    Categorical c(logits, dev);
    expect(c.m_num_events == K, "Categorical.m_num_events");
    expect(max_abs_diff(c.m_probs, torch::softmax(logits, -1)) < 2e-6, "Categorical.m_probs = softmax");
    expect(max_abs_diff(c.m_logits, torch::log_softmax(logits, -1)) < 2e-5, "Categorical.m_logits = logits - logsumexp");
    expect(max_abs_diff(c.logits_to_probs(logits), c.m_probs) == 0.0, "Categorical.logits_to_probs");
    torch::Tensor a = c.sample();
    expect(a.scalar_type() == torch::kLong && a.dim() == 1 && a.size(0) == n, "Categorical.sample shape / dtype");
    expect(a.min().item<int64_t>() >= 0 && a.max().item<int64_t>() < K, "Categorical.sample range");
    torch::Tensor a2 = c.sample();
    expect(!torch::equal(a, a2), "Categorical.sample draws a fresh stream per call");
    {   // empirical frequencies of 64 x 4096 draws against the mean probabilities (binomial sd ~ 1e-3)
        torch::Tensor freq = torch::zeros({K}, torch::kFloat64);
        for (int r = 0; r < 64; r++) freq += torch::bincount(c.sample().cpu(), {}, K).to(torch::kFloat64);
        freq /= (double)(64 * n);
        expect(max_abs_diff(freq, c.m_probs.mean(0)) < 6e-3, "Categorical.sample frequencies follow m_probs");
    }
    torch::Tensor lp = c.log_prob(a);
    expect(max_abs_diff(lp, torch::log_softmax(logits, -1).gather(-1, a.unsqueeze(-1)).squeeze(-1)) < 2e-5,
           "Categorical.log_prob = gather(log_softmax)");
    torch::Tensor ent = c.entropy();   // the reference's clamp(min = FLT_MIN) quirk: -FLT_MIN * sum(p)  
    expect(ent.size(0) == n && (ent + FLT_MIN).abs().max().item<float>() < 1e-44f, "Categorical.entropy (reference quirk)");
    expect(torch::equal(c.mode(), torch::argmax(c.m_probs, -1)), "Categorical.mode");
    expect(c.mean().isnan().all().item<bool>() && c.variance().isnan().all().item<bool>(), "Categorical.mean / variance are NaN");
    expect(c.enumerate_support().size(0) == K && c.enumerate_support().size(1) == n, "Categorical.enumerate_support");

    torch::Tensor mask = torch::rand({n, K}, torch::TensorOptions().device(*dev)) < 0.6;
    mask.index_put_({torch::indexing::Slice(), 0}, true);
    CategoricalMasked m(logits, mask, dev);
    torch::Tensor ml = torch::where(mask, logits, torch::full_like(logits, -1e8f));
    expect(max_abs_diff(m.m_probs, torch::softmax(ml, -1)) < 2e-6, "CategoricalMasked.m_probs = softmax(where(mask, x, -1e8))");
    expect(m.m_probs.masked_select(~mask).abs().max().item<float>() == 0.0f, "CategoricalMasked: masked probabilities are 0");
    torch::Tensor lsm = torch::log_softmax(ml, -1);
    expect(max_abs_diff(m.m_logits.masked_select(mask), lsm.masked_select(mask)) < 2e-5, "CategoricalMasked.m_logits");
    torch::Tensor ma = m.sample();
    expect(mask.gather(-1, ma.unsqueeze(-1)).all().item<bool>(), "CategoricalMasked.sample never picks a masked action");
    expect(max_abs_diff(m.log_prob(ma), lsm.gather(-1, ma.unsqueeze(-1)).squeeze(-1)) < 2e-5, "CategoricalMasked.log_prob");
    torch::Tensor want_ent = -torch::where(mask, lsm * torch::softmax(ml, -1), torch::zeros_like(lsm)).sum(-1);
    expect(max_abs_diff(m.entropy(), want_ent) < 2e-5, "CategoricalMasked.entropy = -sum(where(mask, logp * p, 0))");
    expect(torch::equal(m.m_masks, mask), "CategoricalMasked.m_masks");
}

static void check_agent(std::shared_ptr<torch::Device> dev) {
    torch::manual_seed(3);
    const int64_t n = 1000, O = 4, A = 2;
    auto agent = std::make_shared<Agent>(O, A, dev);
    agent->to(*dev);
    torch::NoGradGuard ng;
    expect(agent->named_parameters().size() == 12 && agent->named_parameters().contains("m_Critic.criticInputLayer.weight") &&
               agent->named_parameters().contains("m_Actor.actorOutputLayer.bias"),
           "Agent registers the reference's 12 parameters under its names");
    for (auto& p : agent->parameters())
        if (p.dim() == 1) p.copy_(torch::randn_like(p) * 0.1);   // non-zero biases
    torch::Tensor x = torch::randn({n, O}, torch::TensorOptions().device(*dev));
    torch::Tensor v = agent->getValue(x);
    torch::Tensor v_ref = agent->m_Critic->forward(x);            // LibTorch: what Agent.cpp:107-109 computes
    expect(v.sizes() == v_ref.sizes() && max_abs_diff(v, v_ref) < 2e-5, "Agent.getValue = m_Critic->forward (LibTorch)");
    AgentOutput o = agent->getActionAndValueDiscrete(x);
    expect(o.action.scalar_type() == torch::kLong && o.action.dim() == 1 && o.action.size(0) == n &&
               o.action.min().item<int64_t>() >= 0 && o.action.max().item<int64_t>() < A,
           "Agent.getActionAndValueDiscrete samples int64 [n] actions");
    torch::Tensor lsm = torch::log_softmax(agent->m_Actor->forward(x), -1);
    expect(max_abs_diff(o.logprob, lsm.gather(-1, o.action.unsqueeze(-1)).squeeze(-1)) < 2e-5,
           "Agent.getActionAndValueDiscrete log-prob = LibTorch log_softmax(m_Actor(x))[action]");
    expect(max_abs_diff(o.value, v_ref) < 2e-5, "Agent.getActionAndValueDiscrete value");
    expect((o.entropy + FLT_MIN).abs().max().item<float>() < 1e-44f, "Agent entropy keeps the reference quirk");
    AgentOutput o2 = agent->getActionAndValueDiscrete(x, o.action);
    expect(torch::equal(o2.logprob, o.logprob) && torch::equal(o2.action, o.action), "Agent: evaluate mode reproduces the sampled log-probs");
    // parameters rebound behind the agent's back (what torch::load does) are picked up again
    for (auto& p : agent->parameters()) p.set_data(p.detach().clone() * 1.5);
    expect(max_abs_diff(agent->getValue(x), agent->m_Critic->forward(x)) < 5e-5, "Agent follows rebound parameters (torch::load)");

    // MultiDiscrete: nvec {3, 3, 2} with masks (Agent.cpp:137-170)
    auto multi = std::make_shared<Agent>(O, 8, dev, 64, std::vector<int64_t>{3, 3, 2});
    multi->to(*dev);
    torch::Tensor mask = torch::rand({n, 8}, torch::TensorOptions().device(*dev)) < 0.7;
    for (int64_t c : {0, 3, 6}) mask.index_put_({torch::indexing::Slice(), c}, true);
    AgentOutput mo = multi->getActionAndValueMasked(x, mask);
    expect(mo.action.size(0) == n && mo.action.size(1) == 3, "Agent.getActionAndValueMasked returns [n, heads] actions");
    torch::Tensor logits = multi->m_Actor->forward(x);
    torch::Tensor ml = torch::where(mask, logits, torch::full_like(logits, -1e8f));
    torch::Tensor want_lp = torch::zeros({n}, logits.options()), want_ent = torch::zeros({n}, logits.options());
    int64_t off = 0, h = 0;
    for (int64_t k : {3, 3, 2}) {
        torch::Tensor l = torch::log_softmax(ml.narrow(1, off, k), -1), mk = mask.narrow(1, off, k);
        torch::Tensor ah = mo.action.select(1, h);
        expect(mk.gather(-1, ah.unsqueeze(-1)).all().item<bool>(), "masked head " + std::to_string(h) + " samples valid actions only");
        want_lp += l.gather(-1, ah.unsqueeze(-1)).squeeze(-1);
        want_ent += -torch::where(mk, l * l.exp(), torch::zeros_like(l)).sum(-1);
        off += k; h++;
    }
    expect(max_abs_diff(mo.logprob, want_lp) < 5e-5, "Agent.getActionAndValueMasked log-prob summed over heads");
    expect(max_abs_diff(mo.entropy, want_ent) < 5e-5, "Agent.getActionAndValueMasked entropy summed over heads");
    AgentOutput mo2 = multi->getActionAndValueMasked(x, mask, mo.action.t());   // the reference passes [heads, n]
    expect(max_abs_diff(mo2.logprob, mo.logprob) == 0.0, "Agent.getActionAndValueMasked evaluate mode");
}

static void check_trainer() {
    PPO_Discrete algo;   // ./PPOConfig.toml
    const int64_t N = algo.m_num_envs, T = algo.m_num_steps, O = algo.m_obs_size;
    expect(algo.m_device && algo.m_device->is_cuda() && algo.m_agent != nullptr, "PPO_Discrete.m_device / m_agent");
    expect(algo.m_obs.sizes() == torch::IntArrayRef({T, N, O}) && algo.m_logprobs.sizes() == torch::IntArrayRef({T, N}) &&
               algo.m_actions.size(0) == T && algo.m_rewards.size(1) == N && algo.m_dones.is_cuda() && algo.m_values.is_cuda(),
           "PPO_Discrete storage tensors m_obs .. m_values are [T, N, ..] device views");
    torch::Tensor obs = algo.initEnvs();
    expect(obs.sizes() == torch::IntArrayRef({N, O}) && obs.abs().max().item<float>() <= 0.05f, "PPO_Discrete.initEnvs");
    AgentOutput out = algo.computeActionLogic(obs);
    expect(out.action.size(0) == N && out.value.size(0) == N, "PPO_Discrete.computeActionLogic");
    auto [o2, r, d] = algo.stepEnvs(out.action);
    expect(o2.sizes() == obs.sizes() && r.sizes() == torch::IntArrayRef({N, 1}) && d.scalar_type() == torch::kInt32 &&
               r.abs().max().item<float>() == 1.0f,
           "PPO_Discrete.stepEnvs returns (obs [N,O], reward [N,1], done int32 [N,1])");
    expect(max_abs_diff(o2.select(1, 0), obs.select(1, 0) + 0.02 * obs.select(1, 1)) < 1e-6, "stepEnvs advances x by tau * x_dot");
    // one update with update_epochs as configured; then calcAdvantage on the stored rollout reproduces the trainer's GAE
    ppo_iter_result res = algo.iterate(algo.m_learning_rate);
    expect(std::isfinite(res.loss), "PPO_Discrete.iterate");
    int64_t b = 0;
    float* st = static_cast<float*>(ppo_trainer_buffer(algo.handle(), "state", &b));
    torch::Tensor next_obs = torch::from_blob(st, {O, N}, torch::TensorOptions().device(*algo.m_device)).t().contiguous();
    float* ndp = static_cast<float*>(ppo_trainer_buffer(algo.handle(), "next_done", &b));
    torch::Tensor next_done = torch::from_blob(ndp, {N}, torch::TensorOptions().device(*algo.m_device));
    auto ra = algo.calcAdvantage(next_obs, next_done);
    float* advp = static_cast<float*>(ppo_trainer_buffer(algo.handle(), "adv", &b));
    float* retp = static_cast<float*>(ppo_trainer_buffer(algo.handle(), "ret", &b));
    torch::Tensor adv = torch::from_blob(advp, {T, N}, torch::TensorOptions().device(*algo.m_device));
    torch::Tensor ret = torch::from_blob(retp, {T, N}, torch::TensorOptions().device(*algo.m_device));
    if (algo.m_update_epochs == 0) {   // parameters unchanged since the rollout: same bootstrap value
        expect(max_abs_diff(ra[1], adv) < 1e-4 && max_abs_diff(ra[0], ret) < 1e-4, "PPO_Discrete.calcAdvantage = the trainer's GAE");
    } else {
        expect(max_abs_diff(ra[0] - ra[1], algo.m_values) < 1e-5, "PPO_Discrete.calcAdvantage: returns - advantages = values");
    }
    torch::Tensor ratio = torch::full({8}, 1.0f, torch::TensorOptions().device(*algo.m_device));
    ratio.index_put_({0}, 1.5f);
    const size_t before = algo.m_clipfracs.size();
    torch::Tensor kl = algo.getApproxKLAndClippedObj(ratio, ratio.log());
    expect(algo.m_clipfracs.size() == before + 1 && std::fabs(algo.m_clipfracs.back() - 0.125f) < 1e-6f &&
               std::fabs(kl.item<float>() - (0.5f - std::log(1.5f)) / 8.0f) < 1e-6f,
           "PPO_Discrete.getApproxKLAndClippedObj");
    // m_agent aliases the trainer's parameters
    std::vector<float> flat = algo.getParameters();
    torch::Tensor first = algo.m_agent->parameters()[0].reshape(-1).cpu();
    expect(std::fabs(first[0].item<float>() - flat[0]) == 0.0f && std::fabs(first[5].item<float>() - flat[5]) == 0.0f,
           "PPO_Discrete.m_agent parameters alias the trainer's slab");
}

int main(int argc, char** argv) {
    try {
        auto dev = std::make_shared<torch::Device>(torch::kCUDA, 0);
        check_distributions(dev);
        check_agent(dev);
        if (argc > 1 && std::string(argv[1]) == "trainer") check_trainer();
        std::cout << "api_check: all ok" << std::endl;
    } catch (const std::exception& e) {
        std::cout << e.what() << std::endl;
        return 1;
    }
    return 0;
}
