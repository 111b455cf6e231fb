// include/compat/ppo_nn.h -- the reference's PPO interface (include/ppo_nn.h) on top of libb200ppo.so.
//
// Keeps the names and call shapes src/main.cpp uses -- TRAIN_STATUS, MemoryNN::{push_reward,merge,clear},
// PPO(state_dim, action_dim, action_std, lr, betas, gamma, K_epochs, eps_clip), PPO::update(MemoryNN),
// getAction(observation, dim, ppo, memory) -- with torch tensors replaced by plain float vectors and every
// computation delegated to the CUDA engine through the C ABI (include/b200ppo.h).  Header-only, C++17, no
// libtorch, no CPU fallback: constructing a PPO without a CUDA device throws.
//
// Differences from the reference that a caller can observe (all deliberate):
//   * MemoryNN's tensor vectors (include/ppo_nn.h:22) are flat float stores: actions/logprobs one float per
//     record, states a StateStore (state_dim floats per record) whose size() counts RECORDS like the vector of
//     tensors it replaces, so the size guard of src/main.cpp:281-285 compiles and holds unmodified.
//   * errors are std::runtime_error carrying b2_last_error() instead of prints / exit(1) / torch exceptions.
//   * nothing is printed (the reference prints every tensor it touches, e.g. src/ppo_nn.cpp:68,292-297).
//   * action noise comes from the engine's Philox stream, not torch's mt19937.
#ifndef B200PPO_COMPAT_PPO_NN_H
#define B200PPO_COMPAT_PPO_NN_H
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "../b200ppo.h"

enum TRAIN_STATUS { NORMAL = 0, TERMINATE, TIMEUP, START };  // include/ppo_nn.h:15

// Conveniences src/main.cpp gets transitively from <torch/torch.h> and uses unqualified: isnan() in the global
// namespace (src/main.cpp:50) and stream output of std::vector (c10/util/Logging.h puts one in namespace std;
// src/main.cpp:52,176).  Provided here so that the reference's driver compiles unmodified against this header.
#include <cmath>
#include <iomanip>   // std::setprecision at src/main.cpp:142, also reached through torch upstream
#include <iostream>
using std::isnan;
namespace std {
template <class T>
ostream& operator<<(ostream& os, const vector<T>& v) {
  for (size_t i = 0; i < v.size(); i++) os << (i ? " " : "") << v[i];
  return os;
}
}  // namespace std

namespace b2compat {
inline void check(int rc, const char* what) {
  if (rc != B2_OK) throw std::runtime_error(std::string(what) + ": " + b2_last_error());
}
struct Handle {  // shared by every copy of a PPO, like the reference's module shared_ptrs (include/ppo_nn.h:183)
  b2_handle h = nullptr;
  ~Handle() { if (h) b2_destroy(h); }
};
// observations of the recorded transitions, `dim` floats each; size() == number of records
struct StateStore {
  std::vector<float> data;
  size_t dim = 6;
  size_t size() const { return dim ? data.size() / dim : 0; }
  bool empty() const { return data.empty(); }
  void clear() { data.clear(); }
  void append(const float* p, size_t n_floats) { data.insert(data.end(), p, p + n_floats); }
  void append(const StateStore& o) { data.insert(data.end(), o.data.begin(), o.data.end()); }
};
}  // namespace b2compat

// Trajectory store, one record per chosen action: states/actions/logprobs are appended when
// the action is chosen (src/ppo_nn.cpp:118-123), reward and flags when the env's next message arrives
// (src/main.cpp:256-262).
class MemoryNN {
 public:
  std::vector<float> actions, logprobs;  // [n]
  b2compat::StateStore states;           // [n][state_dim]
  std::vector<double> rewards;
  std::vector<bool> is_terminals;
  std::vector<bool> is_timeups;
  void push_reward(double reward, bool terminate, bool timeup) {  // src/ppo_nn.cpp:67-72
    rewards.push_back(reward);
    is_terminals.push_back(terminate);
    is_timeups.push_back(timeup);
  }
  void merge(MemoryNN& r) {  // src/ppo_nn.cpp:74-88: append r's records after ours, r is left untouched
    actions.insert(actions.end(), r.actions.begin(), r.actions.end());
    states.append(r.states);
    logprobs.insert(logprobs.end(), r.logprobs.begin(), r.logprobs.end());
    rewards.insert(rewards.end(), r.rewards.begin(), r.rewards.end());
    is_terminals.insert(is_terminals.end(), r.is_terminals.begin(), r.is_terminals.end());
    is_timeups.insert(is_timeups.end(), r.is_timeups.begin(), r.is_timeups.end());
  }
  void clear() {  // src/ppo_nn.cpp:89-97
    actions.clear(); states.clear(); logprobs.clear();
    rewards.clear(); is_terminals.clear(); is_timeups.clear();
  }
  size_t size() const { return rewards.size(); }
};

class PPO {
 public:
  // include/ppo_nn.h:146-173.  `device` and `seed` are engine extensions with defaults.
  PPO(int64_t state_dim, int64_t action_dim, double action_std, double lr, std::tuple<double, double> betas,
      double gamma, int64_t K_epochs, double eps_clip, int device = 0, uint64_t seed = 0x5EED)
      : lr(lr), gamma(gamma), eps_clip(eps_clip), betas(betas), K_epochs(K_epochs),
        state_dim(state_dim), action_dim(action_dim), engine(std::make_shared<b2compat::Handle>()) {
    b2_config c;
    b2compat::check(b2_config_default(&c), "b2_config_default");
    c.obs_vars = (int)state_dim;
    c.control_vars = (int)action_dim;
    c.action_std = action_std;
    c.lr = lr;
    c.beta1 = std::get<0>(betas);
    c.beta2 = std::get<1>(betas);
    c.gamma = gamma;
    c.k_epochs = (int)K_epochs;
    c.eps_clip = eps_clip;
    c.n_envs = 1;      // this interface keeps the environments on the caller's side
    c.horizon = 1;     // the trajectory buffer grows to whatever update() is handed
    c.device = device;
    c.seed = seed;
    b2compat::check(b2_create(&c, &engine->h), "PPO");
  }

  // PPO::select_action -> ActorCritic::act (src/ppo_nn.cpp:196-203,99-125): sample a = mu + sigma eps, record
  // (state, action, logprob) in `memory`, return (action, logprob).
  std::tuple<std::vector<float>, float> select_action(const std::vector<float>& state, MemoryNN& memory) {
    if ((int64_t)state.size() != state_dim) throw std::runtime_error("select_action: observation size != state_dim");
    float a = 0.f, lp = 0.f;
    b2compat::check(b2_act(engine->h, state.data(), 1, nullptr, &a, &lp, nullptr), "select_action");
    memory.states.dim = (size_t)state_dim;
    memory.states.append(state.data(), state.size());
    memory.actions.push_back(a);
    memory.logprobs.push_back(lp);
    return {std::vector<float>{a}, lp};
  }

  // batched action query: n observations at once (what the learner's per-env loop becomes on a GPU)
  void select_actions(const float* obs, int64_t n, float* action, float* logp) {
    b2compat::check(b2_act(engine->h, obs, n, nullptr, action, logp, nullptr), "select_actions");
  }

  // value estimate: policy.critic->forward(state) (src/ppo_nn.cpp:189,227)
  std::vector<float> value(const std::vector<float>& states) {
    const int64_t n = (int64_t)states.size() / state_dim;
    std::vector<float> v((size_t)n);
    b2compat::check(b2_value(engine->h, states.data(), n, v.data()), "value");
    return v;
  }

  // PPO::update(MemoryNN) (src/ppo_nn.cpp:205-307): returns with the TIMEUP bootstrap, K_epochs full-batch
  // epochs of clipped surrogate + 0.5 MSE value loss (+ the entropy constant), Adam.  By value, as upstream.
  void update(MemoryNN memory) {
    const size_t n = memory.rewards.size();
    if (n == 0) return;
    if (memory.actions.size() != n || memory.logprobs.size() != n || memory.states.data.size() != n * (size_t)state_dim ||
        memory.is_terminals.size() != n || memory.is_timeups.size() != n)
      throw std::runtime_error("PPO::update: MemoryNN fields have inconsistent sizes (cf. src/main.cpp:281-289)");
    std::vector<float> reward(n);
    std::vector<uint8_t> flags(n);
    for (size_t i = 0; i < n; i++) {
      reward[i] = (float)memory.rewards[i];
      flags[i] = (uint8_t)((memory.is_terminals[i] ? B2_FLAG_TERMINAL : 0u) | (memory.is_timeups[i] ? B2_FLAG_TIMEUP : 0u));
    }
    b2_batch b;
    b.n_envs = 1;  // the merged memory is one flat sequence (src/main.cpp:290-303)
    b.horizon = (int64_t)n;
    b.obs = memory.states.data.data();
    b.action = memory.actions.data();
    b.logp = memory.logprobs.data();
    b.reward = reward.data();
    b.flags = flags.data();
    b2compat::check(b2_update(engine->h, &b, &last_stats), "PPO::update");
  }

  // parameter access in policy.parameters() order (checkpointing, importing reference weights)
  int64_t n_parameters() const { return b2_n_params(engine->h); }
  std::vector<float> parameters() const {
    std::vector<float> p((size_t)n_parameters());
    b2compat::check(b2_get_params(engine->h, p.data(), (int64_t)p.size()), "parameters");
    return p;
  }
  void load_parameters(const std::vector<float>& p) {
    b2compat::check(b2_set_params(engine->h, p.data(), (int64_t)p.size()), "load_parameters");
  }
  std::vector<float> gradients() const {
    std::vector<float> g((size_t)n_parameters());
    b2compat::check(b2_get_grads(engine->h, g.data(), (int64_t)g.size()), "gradients");
    return g;
  }
  b2_handle handle() const { return engine->h; }

  double lr, gamma, eps_clip;
  std::tuple<double, double> betas;
  int64_t K_epochs;
  int64_t state_dim, action_dim;
  b2_update_stats last_stats{};

 private:
  std::shared_ptr<b2compat::Handle> engine;
};

// src/ppo_nn.cpp:308-331 -- same signature; PPO is taken by value as upstream (copies share the engine).
inline std::tuple<std::vector<float>, float> getAction(std::vector<float> observation, int dim, PPO ppo,
                                                       MemoryNN& memoryNN) {
  (void)dim;
  return ppo.select_action(observation, memoryNN);
}

#endif  // B200PPO_COMPAT_PPO_NN_H
