// PPOConfig.h — the reference's hyper-parameter file, same schema and same fall-back behaviour
// (PPO/PPO_Discrete.cpp:7-34 defaults, :107-255 getArgs): ./PPOConfig.toml in the working directory, three sections
// ([environment], [general], [ppo]), every key optional; a missing file or a parse error prints a message and keeps
// the defaults.  The reference parses with the vendored toml++; this schema is flat (section headers, key = scalar),
// so a ~100-line reader of that TOML subset is all the host needs (strings, integers with '_' separators, floats,
// booleans, comments).  Keys outside the reference schema live in an optional [b200] section and default to the
// reference's hard-coded choices (CartPole / MountainCar by trainer class, hidden width 64).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace ppo_host {

struct Config {
    // [environment]
    int64_t obs_size = 2;            // :7
    int64_t action_size = 1;         // :8
    int64_t max_episode_steps = 500; // :29 (PPO_MultiDiscrete: 200, PPO_MultiDiscrete.cpp:31)
    float action_high = 1.0f, action_low = -1.0f;   // PPO_MultiDiscrete only (unused there as well)
    // [general]
    int64_t seed = 1;                // :10
    int64_t total_timesteps = 100000;// :11
    bool use_cuda = true;            // :12
    bool torch_deterministic = true; // :13
    int64_t checkpoint_updates = 5;  // :28
    // [ppo]
    float learning_rate = 3e-4f;     // :9
    int64_t num_envs = 1;            // :14
    int64_t num_steps = 2048;        // :15
    bool anneal_lr = false;          // :16
    bool use_gae = true;             // :17
    float gamma = 0.99f;             // :18
    float gae_lambda = 0.95f;        // :19
    int64_t num_minibatches = 64;    // :20
    int64_t update_epochs = 10;      // :21
    bool norm_adv = true;            // :22
    float clip_coef = 0.2f;          // :23
    bool clip_vloss = true;          // :24
    float ent_coef = 0.01f;          // :25
    float vf_coef = 0.5f;            // :26
    float max_grad_norm = 0.5f;      // :27
    // [b200] (new, optional)
    std::string env;                 // "" = the trainer class's own env ("cartpole" | "mountaincar")
    int64_t hidden = 64;             // PPO/Agent.cpp:25-37 hard-codes 64
    std::vector<int64_t> nvec;       // MultiDiscrete action space (PPO/Agent.cpp:16 keeps { 3,3,3,3,3,2,2,2 } commented); empty = { action_size }
    bool use_cuda_graph = true;
    bool exact_entropy = false;      // false = reproduce Distributions/Categorical.cpp:114-117
    // derived (:246-247)
    int64_t batch_size = 0, minibatch_size = 0;
    void derive() {
        batch_size = num_envs * num_steps;
        minibatch_size = num_minibatches > 0 ? batch_size / num_minibatches : batch_size;
    }
};

namespace detail {
inline std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}
// strips a trailing comment that is not inside a quoted string
inline std::string strip_comment(const std::string& s) {
    bool in_str = false;
    char q = 0;
    for (size_t i = 0; i < s.size(); i++) {
        const char c = s[i];
        if (in_str) {
            if (c == '\\' && q == '"') i++;
            else if (c == q) in_str = false;
        } else if (c == '"' || c == '\'') {
            in_str = true;
            q = c;
        } else if (c == '#') {
            return s.substr(0, i);
        }
    }
    if (in_str) throw std::runtime_error("unterminated string");
    return s;
}
struct Value {
    enum Kind { INT, FLOAT, BOOL, STRING, INT_ARRAY } kind;
    std::vector<int64_t> arr;   // INT_ARRAY: [3, 3, 2]
    int64_t i = 0;
    double f = 0.0;
    bool b = false;
    std::string s;
};
inline Value parse_value(const std::string& raw) {
    Value v;
    if (raw.empty()) throw std::runtime_error("missing value");
    if (raw == "true" || raw == "false") { v.kind = Value::BOOL; v.b = raw == "true"; return v; }
    if (raw[0] == '[') {   // one-line array of integers (the only array the [b200] table uses: nvec)
        if (raw.back() != ']') throw std::runtime_error("bad array " + raw);
        v.kind = Value::INT_ARRAY;
        std::string item;
        std::istringstream ss(raw.substr(1, raw.size() - 2));
        while (std::getline(ss, item, ',')) {
            item = trim(item);
            if (item.empty()) continue;
            char* e = nullptr;
            const long long x = std::strtoll(item.c_str(), &e, 0);
            if (e == item.c_str() || *e != '\0') throw std::runtime_error("bad array element '" + item + "'");
            v.arr.push_back(x);
        }
        return v;
    }
    if (raw[0] == '"' || raw[0] == '\'') {
        if (raw.size() < 2 || raw.back() != raw[0]) throw std::runtime_error("bad string " + raw);
        v.kind = Value::STRING;
        v.s = raw.substr(1, raw.size() - 2);
        return v;
    }
    std::string t;
    for (char c : raw)
        if (c != '_') t += c;
    if (t == "inf" || t == "+inf" || t == "-inf" || t == "nan") { v.kind = Value::FLOAT; v.f = std::strtod(t.c_str(), nullptr); return v; }
    const bool is_float = t.find_first_of(".eE") != std::string::npos && t.compare(0, 2, "0x") != 0;
    char* end = nullptr;
    if (is_float) {
        v.kind = Value::FLOAT;
        v.f = std::strtod(t.c_str(), &end);
    } else {
        v.kind = Value::INT;
        v.i = std::strtoll(t.c_str(), &end, 0);
        v.f = (double)v.i;
    }
    if (end == t.c_str() || *end != '\0') throw std::runtime_error("bad value '" + raw + "'");
    return v;
}
}  // namespace detail

// section.key -> value; throws std::runtime_error on anything outside the subset
inline std::map<std::string, detail::Value> parse_toml_subset(std::istream& in) {
    std::map<std::string, detail::Value> out;
    std::string line, section;
    int ln = 0;
    while (std::getline(in, line)) {
        ln++;
        try {
            std::string s = detail::trim(detail::strip_comment(line));
            if (s.empty()) continue;
            if (s.front() == '[') {
                if (s.back() != ']' || s.size() < 3 || s[1] == '[') throw std::runtime_error("bad table header");
                section = detail::trim(s.substr(1, s.size() - 2));
                continue;
            }
            const size_t eq = s.find('=');
            if (eq == std::string::npos) throw std::runtime_error("expected key = value");
            const std::string key = detail::trim(s.substr(0, eq));
            if (key.empty()) throw std::runtime_error("empty key");
            out[section.empty() ? key : section + "." + key] = detail::parse_value(detail::trim(s.substr(eq + 1)));
        } catch (const std::exception& e) {
            throw std::runtime_error("line " + std::to_string(ln) + ": " + e.what());
        }
    }
    return out;
}

// getArgs (PPO_Discrete.cpp:107-255): missing file -> message + defaults (:111-115); parse error -> message + defaults
// (:250-254); wrong-typed or absent keys are skipped like toml++'s value<T>() returning nullopt.
// Returns 0 = loaded, 1 = file missing, 2 = parse error (defaults kept in both).
inline int load_config(const std::string& path, Config& c, std::ostream& log = std::cout) {
    std::ifstream f(path);
    if (!f.good()) {
        log << "Could not find " << path << ", using default hyperparameters." << std::endl;
        c.derive();
        return 1;
    }
    std::map<std::string, detail::Value> kv;
    try {
        kv = parse_toml_subset(f);
    } catch (const std::exception& e) {
        log << "Error parsing " << path << " (" << e.what() << "), using default hyperparameters." << std::endl;
        c = Config();
        c.derive();
        return 2;
    }
    using V = detail::Value;
    auto geti = [&](const char* k, int64_t& dst) { auto it = kv.find(k); if (it != kv.end() && it->second.kind == V::INT) dst = it->second.i; };
    auto getf = [&](const char* k, float& dst) {
        auto it = kv.find(k);
        if (it != kv.end() && (it->second.kind == V::FLOAT || it->second.kind == V::INT)) dst = (float)it->second.f;
    };
    auto getb = [&](const char* k, bool& dst) { auto it = kv.find(k); if (it != kv.end() && it->second.kind == V::BOOL) dst = it->second.b; };
    auto gets = [&](const char* k, std::string& dst) { auto it = kv.find(k); if (it != kv.end() && it->second.kind == V::STRING) dst = it->second.s; };
    geti("environment.obs_size", c.obs_size);
    geti("environment.action_size", c.action_size);
    geti("environment.max_episode_steps", c.max_episode_steps);
    getf("environment.action_high", c.action_high);
    getf("environment.action_low", c.action_low);
    geti("general.seed", c.seed);
    geti("general.total_timesteps", c.total_timesteps);
    getb("general.use_cuda", c.use_cuda);
    getb("general.torch_deterministic", c.torch_deterministic);
    geti("general.checkpoint_updates", c.checkpoint_updates);
    getf("ppo.learning_rate", c.learning_rate);
    geti("ppo.num_envs", c.num_envs);
    geti("ppo.num_steps", c.num_steps);
    getb("ppo.anneal_lr", c.anneal_lr);
    getb("ppo.use_gae", c.use_gae);
    getf("ppo.gamma", c.gamma);
    getf("ppo.gae_lambda", c.gae_lambda);
    geti("ppo.num_minibatches", c.num_minibatches);
    geti("ppo.update_epochs", c.update_epochs);
    getb("ppo.norm_adv", c.norm_adv);
    getf("ppo.clip_coef", c.clip_coef);
    getb("ppo.clip_vloss", c.clip_vloss);
    getf("ppo.ent_coef", c.ent_coef);
    getf("ppo.vf_coef", c.vf_coef);
    getf("ppo.max_grad_norm", c.max_grad_norm);
    gets("b200.env", c.env);
    geti("b200.hidden", c.hidden);
    { auto it = kv.find("b200.nvec"); if (it != kv.end() && it->second.kind == V::INT_ARRAY) c.nvec = it->second.arr; }
    getb("b200.use_cuda_graph", c.use_cuda_graph);
    getb("b200.exact_entropy", c.exact_entropy);
    c.derive();
    return 0;
}

}  // namespace ppo_host
