// Distributions.cpp — see Distributions.h.
#include "Distributions.h"

#include <c10/cuda/CUDAStream.h>

#include <atomic>
#include <limits>
#include <stdexcept>
#include <string>

#include "../../include/ppo_b200.h"

namespace ppo_host {

static std::atomic<uint64_t> g_sample_calls{0};

void next_sampling_key(uint64_t* seed, uint64_t* counter) {
    *seed = at::globalContext().defaultGenerator(at::kCPU).current_seed();   // what torch::manual_seed() set
    *counter = g_sample_calls.fetch_add(1);
}
void check_rc(int rc, const char* what) {
    if (rc != PPO_OK) throw std::runtime_error(std::string(what) + ": " + ppo_last_error());
}
void* stream_of(const torch::Tensor& t) {
    return reinterpret_cast<void*>(c10::cuda::getCurrentCUDAStream(t.device().index()).stream());
}

namespace {
// [.., K] logits -> contiguous fp32 [n, K] on the device the kernels run on
torch::Tensor as_rows(const torch::Tensor& logits, std::vector<int64_t>* batch_shape) {
    if (!logits.defined() || !logits.is_cuda())
        throw std::runtime_error("Categorical: logits must be a CUDA tensor (the B200 path has no CPU fallback)");
    const int64_t K = logits.sizes().back();
    batch_shape->assign(logits.sizes().begin(), logits.sizes().end() - 1);
    return logits.detach().to(torch::kFloat32).reshape({-1, K}).contiguous();
}
torch::Tensor shaped(const torch::Tensor& flat, const std::vector<int64_t>& batch_shape) {
    return batch_shape.size() == 1 ? flat : flat.reshape(batch_shape);
}
struct Eval {
    torch::Tensor actions, logprob, entropy;
};
// one ppo_categorical launch: sample (actions_in undefined) or evaluate
Eval evaluate(const torch::Tensor& raw, const torch::Tensor& mask8, const torch::Tensor& actions_in, int entropy_mode) {
    const int64_t n = raw.size(0), K = raw.size(1);
    const int32_t nvec[1] = {(int32_t)K};
    auto opt = raw.options();
    Eval e;
    e.actions = torch::empty({n}, opt.dtype(torch::kInt32));
    e.logprob = torch::empty({n}, opt);
    e.entropy = torch::empty({n}, opt);
    torch::Tensor ain;
    uint64_t seed = 0, counter = 0;
    if (actions_in.defined()) ain = actions_in.to(raw.device()).to(torch::kInt32).reshape({-1}).contiguous();
    else next_sampling_key(&seed, &counter);
    if (ain.defined() && ain.size(0) != n) throw std::runtime_error("Categorical::log_prob: value has the wrong batch size");
    check_rc(ppo_categorical((int)n, 1, nvec, raw.data_ptr<float>(), mask8.defined() ? mask8.data_ptr<uint8_t>() : nullptr,
                             ain.defined() ? ain.data_ptr<int32_t>() : nullptr, nullptr, seed, counter, entropy_mode,
                             e.actions.data_ptr<int32_t>(), e.logprob.data_ptr<float>(), e.entropy.data_ptr<float>(),
                             nullptr, stream_of(raw)),
             "ppo_categorical");
    return e;
}
void normalise(const torch::Tensor& raw, const torch::Tensor& mask8, torch::Tensor* logits, torch::Tensor* probs) {
    const int64_t n = raw.size(0), K = raw.size(1);
    const int32_t nvec[1] = {(int32_t)K};
    *logits = torch::empty_like(raw);
    *probs = torch::empty_like(raw);
    check_rc(ppo_categorical_normalize((int)n, 1, nvec, raw.data_ptr<float>(),
                                       mask8.defined() ? mask8.data_ptr<uint8_t>() : nullptr, logits->data_ptr<float>(),
                                       probs->data_ptr<float>(), stream_of(raw)),
             "ppo_categorical_normalize");
}
torch::Tensor nan_like_rows(const torch::Tensor& probs, const torch::Device& dev) {
    return torch::full({probs.sizes()[0]}, std::numeric_limits<float>::quiet_NaN(),
                       torch::TensorOptions().device(dev).dtype(probs.dtype()));
}
}  // namespace
}  // namespace ppo_host

using namespace ppo_host;

// ---------------------------------------------------------------------------------------------- Categorical
Categorical::Categorical() { m_num_events = 0; }
Categorical::Categorical(const torch::Tensor& logits, std::shared_ptr<torch::Device> device) {
    m_raw = as_rows(logits, &m_batch_shape);
    torch::Tensor ln, pr;
    normalise(m_raw, torch::Tensor(), &ln, &pr);             // Categorical.cpp:33-35
    m_logits = ln.reshape(logits.sizes());
    m_probs = pr.reshape(logits.sizes());
    m_num_events = static_cast<int64_t>(m_probs.sizes().back());
    m_device = device;
}
Categorical::~Categorical() {}
torch::Tensor Categorical::logits_to_probs(torch::Tensor logits, bool is_binary) {
    if (is_binary) return torch::sigmoid(logits);            // :66 (not on the PPO path)
    std::vector<int64_t> bs;
    torch::Tensor raw = as_rows(logits, &bs), ln, pr;
    normalise(raw, torch::Tensor(), &ln, &pr);
    return pr.reshape(logits.sizes());
}
torch::Tensor Categorical::sample() {                        // :73-79 -> int64 [n]
    return shaped(evaluate(m_raw, torch::Tensor(), torch::Tensor(), PPO_ENTROPY_REFERENCE).actions.to(torch::kLong),
                  m_batch_shape);
}
torch::Tensor Categorical::log_prob(torch::Tensor value) {   // :92-101
    return shaped(evaluate(m_raw, torch::Tensor(), value, PPO_ENTROPY_REFERENCE).logprob, m_batch_shape);
}
torch::Tensor Categorical::entropy() {                       // :112-119, the clamp(min = FLT_MIN) quirk included
    torch::Tensor zeros = torch::zeros({m_raw.size(0)}, m_raw.options().dtype(torch::kInt32));
    return shaped(evaluate(m_raw, torch::Tensor(), zeros, PPO_ENTROPY_REFERENCE).entropy, m_batch_shape);
}
torch::Tensor Categorical::mean() { return nan_like_rows(m_probs, *m_device); }
torch::Tensor Categorical::mode() { return torch::argmax(m_probs, -1); }
torch::Tensor Categorical::variance() { return nan_like_rows(m_probs, *m_device); }
torch::Tensor Categorical::enumerate_support() {
    torch::Tensor values = torch::arange(m_num_events, torch::TensorOptions().dtype(torch::kLong).device(*m_device));
    return values.view({-1, 1}).expand({-1, static_cast<int64_t>(m_probs.sizes()[0])});
}

// ---------------------------------------------------------------------------------------------- CategoricalMasked
CategoricalMasked::CategoricalMasked() { m_num_events = 0; }
CategoricalMasked::CategoricalMasked(const torch::Tensor& logits, const torch::Tensor& masks,
                                     std::shared_ptr<torch::Device> device) {
    m_masks = masks.to(torch::kBool).to(*device);            // CategoricalMasked.cpp:32
    m_raw = as_rows(logits, &m_batch_shape);
    m_mask8 = m_masks.to(torch::kUInt8).reshape({-1, m_raw.size(1)}).contiguous();
    torch::Tensor ln, pr;
    normalise(m_raw, m_mask8, &ln, &pr);                     // :35-40: where(mask, x, -1e8), then as Categorical
    m_logits = ln.reshape(logits.sizes());
    m_probs = pr.reshape(logits.sizes());
    m_num_events = m_probs.sizes().back();
    m_device = device;
}
CategoricalMasked::~CategoricalMasked() {}
torch::Tensor CategoricalMasked::logits_to_probs(torch::Tensor logits, bool is_binary) {
    if (is_binary) return torch::sigmoid(logits);
    std::vector<int64_t> bs;
    torch::Tensor raw = as_rows(logits, &bs), ln, pr;
    normalise(raw, torch::Tensor(), &ln, &pr);
    return pr.reshape(logits.sizes());
}
torch::Tensor CategoricalMasked::sample() {
    return shaped(evaluate(m_raw, m_mask8, torch::Tensor(), PPO_ENTROPY_EXACT).actions.to(torch::kLong), m_batch_shape);
}
torch::Tensor CategoricalMasked::log_prob(torch::Tensor value) {
    return shaped(evaluate(m_raw, m_mask8, value, PPO_ENTROPY_EXACT).logprob, m_batch_shape);
}
torch::Tensor CategoricalMasked::entropy() {                 // :127-144: -sum(where(mask, logp * p, 0))
    torch::Tensor zeros = torch::zeros({m_raw.size(0)}, m_raw.options().dtype(torch::kInt32));
    return shaped(evaluate(m_raw, m_mask8, zeros, PPO_ENTROPY_EXACT).entropy, m_batch_shape);
}
torch::Tensor CategoricalMasked::mean() { return nan_like_rows(m_probs, *m_device); }
torch::Tensor CategoricalMasked::mode() { return torch::argmax(m_probs, -1); }
torch::Tensor CategoricalMasked::variance() { return nan_like_rows(m_probs, *m_device); }
torch::Tensor CategoricalMasked::enumerate_support() {
    torch::Tensor values = torch::arange(m_num_events, torch::TensorOptions().dtype(torch::kLong).device(*m_device));
    return values.view({-1, 1}).expand({-1, static_cast<int64_t>(m_probs.sizes()[0])});
}
