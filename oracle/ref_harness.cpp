// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Drives the UNMODIFIED reference sources where they lie under /root/reference
// (include/cart_pole.h, include/ppo_nn.h, src/ppo_nn.cpp are #included, never
// copied) to (1) mint the golden vectors under tests/golden/ and (2) time the
// reference's own CPU path for bench.py's `--impl reference` / `cpu_baseline`.
//
// Why #include the .cpp: ActorCritic::act, PPO::select_action and
// multivariateLogProb have deduced `auto` return types (include/ppo_nn.h:131,175;
// src/ppo_nn.cpp:8,99,196) and are therefore only usable in the TU that defines
// them.  src/main.cpp cannot be built here (no mpi.h), so this file replays the
// request/response protocol of env_run (src/main.cpp:12-148) and NN_run
// (src/main.cpp:190-354) in-process: same calls, same order, no transport.
//
// Built by oracle/Makefile into oracle/_ref/ref_harness (git-ignored).
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "src/ppo_nn.cpp"        // -I/root/reference -I/root/reference/include
#include "cart_pole.h"

namespace {

struct NullBuf : std::streambuf {
  int overflow(int c) override { return c; }
  std::streamsize xsputn(const char*, std::streamsize n) override { return n; }
};

struct Quiet {  // reference prints O(n) debug text per call; send it nowhere
  std::streambuf* old;
  NullBuf nb;
  Quiet() : old(std::cout.rdbuf(&nb)) {}
  ~Quiet() { std::cout.rdbuf(old); }
};

struct Capture {  // or keep it, to read the LOSSx lines update() prints
  std::streambuf* old;
  std::ostringstream ss;
  std::streamsize oldprec;
  Capture() : old(std::cout.rdbuf(ss.rdbuf())), oldprec(std::cout.precision(17)) {}
  ~Capture() { std::cout.rdbuf(old); std::cout.precision(oldprec); }
};

template <class T>
void wr(const std::string& dir, const std::string& name, const std::vector<T>& v) {
  std::ofstream f(dir + "/" + name, std::ios::binary);
  f.write(reinterpret_cast<const char*>(v.data()), sizeof(T) * v.size());
  if (!f) { fprintf(stderr, "write failed: %s\n", name.c_str()); exit(2); }
}

std::vector<float> flat(const torch::Tensor& t) {
  auto c = t.detach().to(torch::kFloat32).contiguous().flatten();
  return std::vector<float>(c.data_ptr<float>(), c.data_ptr<float>() + c.numel());
}

std::vector<float> params_of(PPO& p) {  // policy.parameters(): actor l1.w,l1.b,l2.w,l2.b,l3.w,l3.b then critic
  std::vector<float> out;
  for (auto& q : p.policy.parameters()) { auto v = flat(q); out.insert(out.end(), v.begin(), v.end()); }
  return out;
}
std::vector<float> grads_of(PPO& p) {
  std::vector<float> out;
  for (auto& q : p.policy.parameters()) { auto v = flat(q.grad()); out.insert(out.end(), v.begin(), v.end()); }
  return out;
}
void adam_of(PPO& p, std::vector<float>& m, std::vector<float>& v, std::vector<int64_t>& step) {
  auto& st = p.optimizer->state();
  for (auto& q : p.policy.parameters()) {
    auto it = st.find(q.unsafeGetTensorImpl());
    if (it == st.end()) { fprintf(stderr, "no adam state\n"); exit(2); }
    auto& s = static_cast<torch::optim::AdamParamState&>(*it->second);
    auto a = flat(s.exp_avg()), b = flat(s.exp_avg_sq());
    m.insert(m.end(), a.begin(), a.end());
    v.insert(v.end(), b.begin(), b.end());
    step.push_back(s.step());
  }
}

// observation width: 6 (config:2, the shipped default) or 4 (getobs(4), include/cart_pole.h:120-127 -- the cos/sin
// pair is dropped and the nets are built with state_dim 4); only the golden_obs4 command changes it
static int g_obs_dim = 6;

PPO make_ppo(uint64_t seed, int64_t K) {
  torch::manual_seed(seed);
  Quiet q;
  // hyper-parameters: src/main.cpp:196-206
  return PPO(g_obs_dim, 1, 0.5, 0.0003, std::make_tuple(0.9, 0.999), 0.99, K, 0.2);
}

// One recorded transition (SURVEY 3.6 record alignment).
struct Trace {
  std::vector<float> obs, action, logp, mu, value, reward, force;
  std::vector<uint8_t> timeup, terminal;
  std::vector<double> state_after;  // fp64 (y1,y2,y3,y4) after advance
};

// In-process replay of env_run + NN_run for `nenv` env ranks, `episodes` each.
// Poll order = rank order each round, like the MPI_Test loop (src/main.cpp:227).
// Returns the merged memory in rank order (src/main.cpp:290-303) and per-env traces.
MemoryNN rollout(PPO& ppo, int nenv, int episodes, int n_timestep,
                 std::vector<Trace>& tr, std::vector<double>* ep_rewards) {
  std::vector<MemoryNN> mem(nenv);
  std::vector<CartPole> env(nenv);
  tr.assign(nenv, Trace());
  const double hi = 10, lo = -10;  // src/main.cpp:18
  for (int ep = 0; ep < episodes; ep++) {
    std::vector<std::vector<float>> obs(nenv);
    std::vector<double> epr(nenv, 0.0);
    for (int e = 0; e < nenv; e++) {
      env[e].reset();                                   // src/main.cpp:27
      auto raw = env[e].getobs(g_obs_dim);              // :29
      obs[e] = std::vector<float>(raw.begin(), raw.end());  // :34  (START message, no reward pushed :256)
    }
    for (int j = 0; j < n_timestep; j++) {
      for (int e = 0; e < nenv; e++) {
        Trace& T = tr[e];
        // learner side: respond_to_env -> getAction (src/main.cpp:172-174)
        auto [act, lp] = getAction(obs[e], 1, ppo, mem[e]);
        {
          torch::NoGradGuard ng;
          auto st = torch::from_blob(obs[e].data(), {1, g_obs_dim}, torch::kFloat32).clone();
          T.mu.push_back(ppo.policy.actor->forward(st).item<float>());
          T.value.push_back(ppo.policy.critic->forward(st).item<float>());
        }
        T.obs.insert(T.obs.end(), obs[e].begin(), obs[e].end());
        T.action.push_back(act[0]);
        T.logp.push_back(lp);
        // env side: scale (src/main.cpp:57-60) -- float vector, double-promoted arithmetic
        std::vector<float> action(act);
        for (size_t k = 0; k < action.size(); k++)
          action[k] = (action[k] + 1) * (hi - lo) / 2 + lo;
        T.force.push_back(action[0]);
        std::vector<double> action_double(action.begin(), action.end());  // :69
        bool terminate = env[e].advance(action_double);                    // :71
        auto raw = env[e].getobs(g_obs_dim);                               // :72
        double reward = env[e].getReward();                                // :73
        epr[e] += reward;
        obs[e] = std::vector<float>(raw.begin(), raw.end());               // :79
        float reward_wire = (float)reward;                                 // :80
        bool timeup = !terminate && (j == n_timestep - 1);                 // :81-95
        mem[e].push_reward(reward_wire, terminate, timeup);                // :258-259
        T.reward.push_back(reward_wire);
        T.timeup.push_back(timeup);
        T.terminal.push_back(terminate);
        T.state_after.push_back(env[e].u.y1); T.state_after.push_back(env[e].u.y2);
        T.state_after.push_back(env[e].u.y3); T.state_after.push_back(env[e].u.y4);
      }
    }
    if (ep_rewards) for (int e = 0; e < nenv; e++) ep_rewards->push_back(epr[e]);
  }
  MemoryNN merged;
  for (int e = 0; e < nenv; e++) merged.merge(mem[e]);
  return merged;
}

// libtorch restatement of src/ppo_nn.cpp:207-236 (update()'s `Rewards` local is
// not observable); cross-checked against the printed LOSS2 by the caller.
torch::Tensor returns_of(PPO& ppo, MemoryNN& m) {
  torch::NoGradGuard ng;
  auto r = m.rewards; auto te = m.is_terminals; auto tu = m.is_timeups; auto st = m.states;
  std::reverse(r.begin(), r.end()); std::reverse(te.begin(), te.end());
  std::reverse(tu.begin(), tu.end()); std::reverse(st.begin(), st.end());
  torch::Tensor d = torch::tensor({0.0});
  std::vector<torch::Tensor> out(r.size());
  for (size_t i = 0; i < r.size(); i++) {
    if (tu[i]) d = ppo.policy.critic->forward(st[i].squeeze());
    else if (te[i]) d = torch::tensor({0.0});
    d = r[i] + (ppo.gamma * d);
    out[r.size() - 1 - i] = d;
  }
  return torch::cat(out);
}

// parse "LOSS1 is: <v>\n[ CPUFloatType{} ]" lines from update()'s stdout (src/ppo_nn.cpp:292-297)
void parse_losses(const std::string& s, std::vector<double>& l1, std::vector<double>& l2,
                  std::vector<double>& l3, std::vector<double>& lt) {
  std::istringstream in(s);
  std::string line;
  while (std::getline(in, line)) {
    auto grab = [&](const char* key, std::vector<double>& dst) {
      size_t n = strlen(key);
      if (line.compare(0, n, key) == 0) {
        const char* p = line.c_str() + n;
        char* end = nullptr;
        double v = strtod(p, &end);
        if (end != p) { dst.push_back(v); return true; }
      }
      return false;
    };
    if (grab("LOSS1 is: ", l1)) continue;
    if (grab("LOSS2 is: ", l2)) continue;
    if (grab("LOSS3 is: ", l3)) continue;
    grab("LOSS is: ", lt);  // first of the three "LOSS is:" lines is numeric; the others fail strtod or are 0/1 flags
  }
}

void dump_update(const std::string& dir, const std::string& tag, uint64_t seed, int64_t K,
                 MemoryNN& mem, int n_calls = 1) {
  PPO p = make_ppo(seed, K);
  std::vector<double> l1, l2, l3, lt;
  for (int c = 0; c < n_calls; c++) {
    Capture cap;
    p.update(mem);
    parse_losses(cap.ss.str(), l1, l2, l3, lt);
  }
  // "LOSS is:" appears 3x per epoch: value, requires_grad (1), grad_fn name. keep every first.
  std::vector<double> ltv;
  for (size_t i = 0; i + 1 < lt.size() + 1 && i < lt.size(); i += 2) ltv.push_back(lt[i]);
  std::vector<float> m, v; std::vector<int64_t> step;
  adam_of(p, m, v, step);
  wr(dir, tag + "_params.f32", params_of(p));
  wr(dir, tag + "_grads.f32", grads_of(p));
  wr(dir, tag + "_adam_m.f32", m);
  wr(dir, tag + "_adam_v.f32", v);
  wr(dir, tag + "_adam_step.i64", step);
  wr(dir, tag + "_loss1.f64", l1);
  wr(dir, tag + "_loss2.f64", l2);
  wr(dir, tag + "_loss3.f64", l3);
  wr(dir, tag + "_loss.f64", ltv);
  fprintf(stderr, "[%s] K=%ld calls=%d epochs_logged=%zu loss1[0]=%.9g loss2[0]=%.9g loss3[0]=%.9g loss[0]=%.9g step=%ld\n",
          tag.c_str(), (long)K, n_calls, l1.size(), l1.empty() ? 0 : l1[0], l2.empty() ? 0 : l2[0],
          l3.empty() ? 0 : l3[0], ltv.empty() ? 0 : ltv[0], (long)step[0]);
}

void dump_traces(const std::string& dir, const std::string& tag, std::vector<Trace>& tr) {
  // merged in rank order, like MemoryNN::merge
  Trace all;
  for (auto& t : tr) {
    all.obs.insert(all.obs.end(), t.obs.begin(), t.obs.end());
    all.action.insert(all.action.end(), t.action.begin(), t.action.end());
    all.logp.insert(all.logp.end(), t.logp.begin(), t.logp.end());
    all.mu.insert(all.mu.end(), t.mu.begin(), t.mu.end());
    all.value.insert(all.value.end(), t.value.begin(), t.value.end());
    all.reward.insert(all.reward.end(), t.reward.begin(), t.reward.end());
    all.force.insert(all.force.end(), t.force.begin(), t.force.end());
    all.timeup.insert(all.timeup.end(), t.timeup.begin(), t.timeup.end());
    all.terminal.insert(all.terminal.end(), t.terminal.begin(), t.terminal.end());
    all.state_after.insert(all.state_after.end(), t.state_after.begin(), t.state_after.end());
  }
  wr(dir, tag + "_obs.f32", all.obs);
  wr(dir, tag + "_action.f32", all.action);
  wr(dir, tag + "_logp.f32", all.logp);
  wr(dir, tag + "_mu.f32", all.mu);
  wr(dir, tag + "_value.f32", all.value);
  wr(dir, tag + "_reward.f32", all.reward);
  wr(dir, tag + "_force.f32", all.force);
  wr(dir, tag + "_timeup.u8", all.timeup);
  wr(dir, tag + "_terminal.u8", all.terminal);
  wr(dir, tag + "_state.f64", all.state_after);
}

// ---- env known-answer vectors straight from the unmodified header ----
void env_kat(const std::string& dir) {
  auto run = [&](const std::string& tag, const std::vector<double>& forces) {
    CartPole env; env.reset();
    std::vector<double> st, ob, rw;
    for (double F : forces) {
      env.advance({F});
      st.push_back(env.u.y1); st.push_back(env.u.y2); st.push_back(env.u.y3); st.push_back(env.u.y4);
      auto o = env.getobs(); ob.insert(ob.end(), o.begin(), o.end());
      rw.push_back(env.getReward());
    }
    wr(dir, tag + "_force.f64", forces);
    wr(dir, tag + "_state.f64", st);
    wr(dir, tag + "_obs.f64", ob);
    wr(dir, tag + "_reward.f64", rw);
    fprintf(stderr, "[%s] last state % .17g % .17g % .17g % .17g r=%g\n", tag.c_str(),
            st[st.size() - 4], st[st.size() - 3], st[st.size() - 2], st[st.size() - 1], rw.back());
  };
  run("kat_a", {1, 1, -10, 10, 2.5});
  run("kat_b", std::vector<double>(100, 10.0));
  // pseudo-random forces in [-15, 15], fixed LCG, 300 steps (pole falls and keeps being integrated, Q1)
  std::vector<double> f; uint64_t s = 0x9E3779B97F4A7C15ull;
  for (int i = 0; i < 300; i++) { s = s * 6364136223846793005ull + 1442695040888963407ull;
    f.push_back(((double)(s >> 11) / 9007199254740992.0) * 30.0 - 15.0); }
  run("kat_c", f);
  // forces the reference does not clip (src/main.cpp:57-60; `bounded` unused :17): square waves strong enough that
  // |thetadot| crosses 16 rad/s repeatedly, the range where the CUDA engine leaves its carried-rotation stage for the
  // reference-shaped substep (csrc/cartpole.cuh kOmegaFast)
  std::vector<double> fa, fb;
  for (int i = 0; i < 100; i++) { fa.push_back(((i / 4) % 2 == 0) ? 150.0 : -150.0); fb.push_back(((i / 10) % 2 == 0) ? 40.0 : -40.0); }
  run("kat_fast_a", fa);
  run("kat_fast_b", fb);
}

int cmd_golden(const std::string& dir) {
  env_kat(dir);
  const uint64_t seed = 0;
  {  // initial weights
    PPO p = make_ppo(seed, 1);
    wr(dir, "init_params.f32", params_of(p));
    std::vector<int64_t> shapes;
    for (auto& q : p.policy.named_parameters()) {
      fprintf(stderr, "param %s [", q.key().c_str());
      for (auto d : q.value().sizes()) fprintf(stderr, "%ld ", (long)d);
      fprintf(stderr, "]\n");
    }
  }
  // C1: 1 env x 20 episodes x 100 steps (B=2000).  C2: 3 envs x 7 episodes (B=2100).
  struct Case { const char* tag; int nenv, episodes; };
  for (Case c : {Case{"c1", 1, 20}, Case{"c2", 3, 7}}) {
    PPO p = make_ppo(seed, 1);
    std::vector<Trace> tr; MemoryNN mem;
    { Quiet q; torch::manual_seed(1234); mem = rollout(p, c.nenv, c.episodes, 100, tr, nullptr); }
    dump_traces(dir, c.tag, tr);
    auto R = returns_of(p, mem);
    wr(dir, std::string(c.tag) + "_returns.f32", flat(R));
    {  // cross-check vs unmodified update(): LOSS2 = 0.5*mean((V-R)^2) at epoch 0
      torch::NoGradGuard ng;
      auto S = torch::squeeze(torch::stack(mem.states));
      auto V = p.policy.critic->forward(S).squeeze();
      double l2 = 0.5 * torch::mse_loss(V, R).item<double>();
      fprintf(stderr, "[%s] B=%ld restated loss2=%.9g\n", c.tag, (long)R.numel(), l2);
    }
    std::string t = c.tag;
    dump_update(dir, t + "_k1", seed, 1, mem);
    dump_update(dir, t + "_k5", seed, 5, mem);
    dump_update(dir, t + "_k80", seed, 80, mem);
    if (c.nenv == 1) {
      dump_update(dir, t + "_k2x2", seed, 2, mem, 2);  // Adam step counter persists across update() calls
      // synthetic TERMINATE flags mid-episode (never produced by the shipped env, Q1; exercises :229-230)
      MemoryNN m2 = mem;
      for (size_t i = 37; i < m2.is_terminals.size(); i += 173) { m2.is_terminals[i] = true; }
      auto R2 = returns_of(p, m2);
      wr(dir, t + "_term_returns.f32", flat(R2));
      std::vector<uint8_t> te(m2.is_terminals.begin(), m2.is_terminals.end());
      wr(dir, t + "_term_terminal.u8", te);
      dump_update(dir, t + "_term_k1", seed, 1, m2);
      // perturbed old log-probs so that a large share of ratios leaves [1-eps, 1+eps] (exercises :285-287)
      MemoryNN m3 = mem;
      std::vector<float> lp3;
      for (size_t i = 0; i < m3.logprobs.size(); i++) {
        m3.logprobs[i] = (m3.logprobs[i].detach() + 0.35 * std::sin((double)i)).to(torch::kFloat32);
        lp3.push_back(m3.logprobs[i].item<float>());
      }
      wr(dir, t + "_clip_logp.f32", lp3);
      dump_update(dir, t + "_clip_k1", seed, 1, m3);
      dump_update(dir, t + "_clip_k5", seed, 5, m3);
    }
  }
  return 0;
}

// ---- the 4-observation configuration (obs_vars = 4): PPO(4, ...) + getobs(4), 1 env x 10 episodes x 100 steps ----
int cmd_golden_obs4(const std::string& dir) {
  g_obs_dim = 4;
  const uint64_t seed = 0;
  { PPO p = make_ppo(seed, 1); wr(dir, "o4_init_params.f32", params_of(p)); }
  PPO p = make_ppo(seed, 1);
  std::vector<Trace> tr; MemoryNN mem;
  { Quiet q; torch::manual_seed(4321); mem = rollout(p, 1, 10, 100, tr, nullptr); }
  dump_traces(dir, "o4", tr);
  wr(dir, "o4_returns.f32", flat(returns_of(p, mem)));
  dump_update(dir, "o4_k1", seed, 1, mem);
  dump_update(dir, "o4_k5", seed, 5, mem);
  return 0;
}

// ---- timing of the reference's own CPU path (bench.py --impl reference) ----
int cmd_bench(int nenv, int episodes, int K, int threads, int updates) {
  if (threads > 0) torch::set_num_threads(threads);
  PPO p = make_ppo(0, K);
  Quiet q;
  using clk = std::chrono::steady_clock;
  double t_roll = 0, t_upd = 0; long steps = 0, samples = 0;
  for (int u = 0; u < updates; u++) {
    std::vector<Trace> tr;
    auto t0 = clk::now();
    MemoryNN mem = rollout(p, nenv, episodes, 100, tr, nullptr);
    auto t1 = clk::now();
    p.update(mem);
    auto t2 = clk::now();
    t_roll += std::chrono::duration<double>(t1 - t0).count();
    t_upd += std::chrono::duration<double>(t2 - t1).count();
    steps += (long)mem.rewards.size();
    samples += (long)mem.rewards.size() * K;
  }
  printf(
          "{\"env_steps\": %ld, \"rollout_s\": %.6f, \"update_s\": %.6f, \"sample_epochs\": %ld, "
          "\"torch_threads\": %d, \"nenv\": %d, \"episodes\": %d, \"K\": %d, \"updates\": %d}\n",
          steps, t_roll, t_upd, samples, torch::get_num_threads(), nenv, episodes, K, updates);
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  if (argc >= 3 && std::string(argv[1]) == "golden") return cmd_golden(argv[2]);
  if (argc >= 3 && std::string(argv[1]) == "golden_obs4") return cmd_golden_obs4(argv[2]);
  if (argc >= 7 && std::string(argv[1]) == "bench")
    return cmd_bench(atoi(argv[2]), atoi(argv[3]), atoi(argv[4]), atoi(argv[5]), atoi(argv[6]));
  fprintf(stderr, "usage: %s golden <outdir> | golden_obs4 <outdir> | bench <nenv> <episodes> <K> <threads> <updates>\n", argv[0]);
  return 1;
}
