// tools/compat_check.cpp -- exercises include/compat/{ppo_nn.h,cart_pole.h} the way src/main.cpp uses the
// reference headers, against golden vectors minted from the unmodified reference (tests/golden, made by
// oracle/ref_harness.cpp).  Prints one JSON line of relative errors; exit code 0 iff all are inside the gates.
//   usage: compat_check <tests/golden dir>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>

#include "compat/cart_pole.h"
#include "compat/ppo_nn.h"

template <typename T>
static std::vector<T> load(const std::string& dir, const std::string& name) {
  std::ifstream f(dir + "/" + name, std::ios::binary | std::ios::ate);
  if (!f) throw std::runtime_error("cannot open " + dir + "/" + name);
  const std::streamsize bytes = f.tellg();
  f.seekg(0);
  std::vector<T> v((size_t)bytes / sizeof(T));
  f.read(reinterpret_cast<char*>(v.data()), bytes);
  return v;
}
template <typename A, typename B>
static double relerr(const std::vector<A>& a, const std::vector<B>& b) {
  double num = 0, den = 0;
  for (size_t i = 0; i < b.size(); i++) {
    const double d = (double)a[i] - (double)b[i];
    num += d * d;
    den += (double)b[i] * (double)b[i];
  }
  return std::sqrt(num / (den > 0 ? den : 1e-300));
}

int main(int argc, char** argv) {
  if (argc < 2) { std::fprintf(stderr, "usage: %s <golden dir>\n", argv[0]); return 2; }
  const std::string dir = argv[1];
  try {
    // ---------------- PPO / MemoryNN / getAction, as NN_run does (src/main.cpp:196-206, 256-304)
    auto betas = std::make_tuple(0.9, 0.999);
    PPO ppo = PPO(6, 1, 0.5, 0.0003, betas, 0.99, /*K_epochs=*/1, 0.2);
    ppo.load_parameters(load<float>(dir, "init_params.f32"));
    const auto obs = load<float>(dir, "c1_obs.f32");
    const auto act = load<float>(dir, "c1_action.f32");
    const auto lp = load<float>(dir, "c1_logp.f32");
    const auto rew = load<float>(dir, "c1_reward.f32");
    const auto term = load<uint8_t>(dir, "c1_terminal.u8");
    const auto tup = load<uint8_t>(dir, "c1_timeup.u8");
    const size_t B = rew.size();

    // value estimate on the recorded states
    const double err_value = relerr(ppo.value(obs), load<float>(dir, "c1_value.f32"));

    // two per-env memories filled the way act()/push_reward() fill them, then merged in rank order
    MemoryNN memNN[2], merged;
    for (size_t k = 0; k < B; k++) {
      MemoryNN& m = memNN[k < B / 2 ? 0 : 1];
      m.states.append(obs.data() + 6 * k, 6);
      m.actions.push_back(act[k]);
      m.logprobs.push_back(lp[k]);
      m.push_reward(rew[k], term[k] != 0, tup[k] != 0);
    }
    for (auto& m : memNN) { merged.merge(m); m.clear(); }
    ppo.update(merged);
    const double err_grads = relerr(ppo.gradients(), load<float>(dir, "c1_k1_grads.f32"));
    const double err_params = relerr(ppo.parameters(), load<float>(dir, "c1_k1_params.f32"));
    const double loss_ref = load<double>(dir, "c1_k1_loss.f64").back();
    const double err_loss = std::fabs(ppo.last_stats.loss - loss_ref) / std::fabs(loss_ref);

    // action query: one record appended per call, PPO passed by value shares the engine
    MemoryNN live;
    std::vector<float> o0(obs.begin(), obs.begin() + 6);
    bool act_ok = true;
    for (int i = 0; i < 3; i++) {
      auto [action, logprob] = getAction(o0, 1, ppo, live);
      act_ok = act_ok && action.size() == 1 && std::isfinite(action[0]) && std::isfinite(logprob);
    }
    act_ok = act_ok && live.actions.size() == 3 && live.states.size() == 3 && live.states.data.size() == 18 && live.logprobs.size() == 3;

    // ---------------- CartPole, as env_run does (src/main.cpp:20-79), against the header's own known answers
    double err_state = 0, err_obs = 0;
    bool reward_ok = true;
    for (const char* kat : {"kat_a", "kat_b", "kat_c"}) {
      const auto F = load<double>(dir, std::string(kat) + "_force.f64");
      const auto st = load<double>(dir, std::string(kat) + "_state.f64");
      const auto ob = load<double>(dir, std::string(kat) + "_obs.f64");
      const auto rw = load<double>(dir, std::string(kat) + "_reward.f64");
      CartPole env;
      env.reset();
      for (size_t k = 0; k < F.size(); k++) {
        const bool terminate = env.advance({F[k]});
        (void)terminate;
        const std::vector<double> u = {env.u.y1, env.u.y2, env.u.y3, env.u.y4};
        const std::vector<double> uref(st.begin() + 4 * k, st.begin() + 4 * k + 4);
        const std::vector<double> oref(ob.begin() + 6 * k, ob.begin() + 6 * k + 6);
        err_state = std::fmax(err_state, relerr(u, uref));
        err_obs = std::fmax(err_obs, relerr(env.getobs(), oref));
        reward_ok = reward_ok && env.getReward() == rw[k];
      }
    }
    const bool ok = err_value < 1e-6 && err_grads < 1e-4 && err_params < 1e-5 && err_loss < 1e-4 && act_ok &&
                    err_state < 1e-9 && err_obs < 1e-9 && reward_ok;
    std::printf("{\"ok\": %s, \"err_value\": %.3e, \"err_grads\": %.3e, \"err_params\": %.3e, \"err_loss\": %.3e, "
                "\"act_ok\": %s, \"err_state\": %.3e, \"err_obs\": %.3e, \"reward_ok\": %s}\n",
                ok ? "true" : "false", err_value, err_grads, err_params, err_loss, act_ok ? "true" : "false", err_state,
                err_obs, reward_ok ? "true" : "false");
    return ok ? 0 : 1;
  } catch (const std::exception& e) {
    std::printf("{\"ok\": false, \"error\": \"%s\"}\n", e.what());
    return 3;
  }
}
