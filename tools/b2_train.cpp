// tools/b2_train.cpp -- host orchestrator: the reference's driver semantics (src/main.cpp) on the
// device-resident engine.  C++ host code over the C ABI only (include/b200ppo.h); no MPI, no libtorch.
//
// What it keeps from the reference:
//   * `b2_train [config]` -- same positional config file, default "../config" (src/main.cpp:390-394, config:1-6)
//   * hyper-parameters hard-coded at src/main.cpp:196-203 (defaults of b2_config)
//   * one learner + (nprocs-1) environments: --envs N plays the role of `mpirun -n N+1`
//   * every env runs Nepisodes/(nprocs-1) episodes of N_timestep steps (src/main.cpp:23,44)
//   * update trigger: at an episode boundary once >= updateTimestep transitions are stored (src/main.cpp:264-307);
//     the envs step in lockstep here, so that is after ceil(updateTimestep / (N * N_timestep)) whole episodes
//     per env: 2,000 records for 1 env, 2,100 for 3 envs, exactly the reference's batch sizes
//   * per-env log ./proc<rank>_log.txt, rank = env + 1, one line per episode in the reference's format
//     (src/main.cpp:15-16,139-142) so learning_curve_.m still parses it
// What it deliberately changes:
//   * Max_timestep counts TOTAL training transitions; upstream compares it with a counter that every update
//     resets (src/main.cpp:305,349), so it never fires there
//   * the learner stops when the envs are done (upstream can spin forever)
//   * multi-GPU: one process per GPU (RANK / WORLD_SIZE / LOCAL_RANK), envs sharded contiguously, gradients
//     all-reduced inside b2_update_resident; the NCCL id travels through a file (--rendezvous)
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <string>
#include <thread>
#include <vector>

#include "b200ppo.h"

namespace {

struct Options {
  std::string config = "../config";
  long envs = 1;
  long iters = -1;  // cap on update cycles (-1: run to Nepisodes / Max_timestep)
  int precision = B2_PREC_F64;
  unsigned long long seed = 0x5EED;
  std::string log_dir = ".";
  long log_envs = 16;
  int minibatches = 1;
  int epochs = -1;
  std::string save, resume, rendezvous = "/tmp/b2_train_nccl_id";
  bool quiet = false;
};

int envi(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

[[noreturn]] void die(const char* what) {
  std::fprintf(stderr, "b2_train: %s: %s\n", what, b2_last_error());
  std::exit(1);
}
#define CK(call) do { if ((call) != B2_OK) die(#call); } while (0)

void usage() {
  std::puts("usage: b2_train [config] [--envs N] [--iters I] [--precision f64|f32] [--seed S] [--log-dir D]\n"
            "                [--log-envs K] [--minibatches M] [--epochs K] [--save ckpt] [--resume ckpt]\n"
            "                [--rendezvous file] [--quiet]");
}

Options parse(int argc, char** argv) {
  Options o;
  for (int i = 1; i < argc; i++) {
    const std::string a = argv[i];
    auto next = [&]() -> const char* {
      if (i + 1 >= argc) { usage(); std::exit(2); }
      return argv[++i];
    };
    if (a == "--envs") o.envs = atol(next());
    else if (a == "--iters") o.iters = atol(next());
    else if (a == "--precision") { const std::string v = next(); o.precision = v == "f32" ? B2_PREC_F32 : (v == "mixed" ? B2_PREC_MIXED : B2_PREC_F64); }
    else if (a == "--seed") o.seed = strtoull(next(), nullptr, 0);
    else if (a == "--log-dir") o.log_dir = next();
    else if (a == "--log-envs") o.log_envs = atol(next());
    else if (a == "--minibatches") o.minibatches = atoi(next());
    else if (a == "--epochs") o.epochs = atoi(next());
    else if (a == "--save") o.save = next();
    else if (a == "--resume") o.resume = next();
    else if (a == "--rendezvous") o.rendezvous = next();
    else if (a == "--quiet") o.quiet = true;
    else if (a == "-h" || a == "--help") { usage(); std::exit(0); }
    else if (a.rfind("--", 0) == 0) { usage(); std::exit(2); }
    else o.config = a;
  }
  return o;
}

// rank 0 publishes the 128-byte NCCL id through the filesystem; the others wait for it
void exchange_id(const Options& o, int rank, char id[B2_COMM_ID_BYTES]) {
  const std::string tmp = o.rendezvous + ".tmp";
  if (rank == 0) {
    CK(b2_comm_unique_id(id));
    std::ofstream f(tmp, std::ios::binary);
    f.write(id, B2_COMM_ID_BYTES);
    f.close();
    std::rename(tmp.c_str(), o.rendezvous.c_str());
  } else {
    for (int tries = 0; tries < 6000; tries++) {
      std::ifstream f(o.rendezvous, std::ios::binary);
      if (f && f.read(id, B2_COMM_ID_BYTES)) return;
      std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    std::fprintf(stderr, "b2_train: rank %d timed out waiting for %s\n", rank, o.rendezvous.c_str());
    std::exit(1);
  }
}

}  // namespace

int main(int argc, char** argv) {
  const Options o = parse(argc, argv);
  const int rank = envi("RANK", 0), world = envi("WORLD_SIZE", 1), local = envi("LOCAL_RANK", rank);

  b2_config c;
  CK(b2_config_default(&c));
  CK(b2_config_read(o.config.c_str(), &c));  // src/main.cpp:410
  if (o.envs < 1 || o.envs % world != 0) {
    std::fprintf(stderr, "b2_train: --envs must be a positive multiple of WORLD_SIZE\n");
    return 2;
  }
  const long n_global = o.envs, n_local = n_global / world;
  const long episodes_per_env = c.n_episodes / n_global;  // src/main.cpp:23
  // episodes per env between updates: the first episode boundary with >= updateTimestep records (src/main.cpp:268)
  const long per_round = n_global * (long)c.n_timestep;
  const long eps_per_update = (c.update_timestep + per_round - 1) / per_round;
  const int T = (int)(eps_per_update * c.n_timestep);
  c.n_envs = n_local;
  c.horizon = T;
  c.env_offset = (uint64_t)rank * (uint64_t)n_local;
  c.precision = o.precision;
  c.seed = o.seed;
  c.device = local;
  c.n_minibatches = o.minibatches;
  if (o.epochs > 0) c.k_epochs = o.epochs;

  b2_handle h = nullptr;
  CK(b2_create(&c, &h));
  CK(b2_init_params(h, o.seed));  // same seed on every rank: replicas start identical
  if (!o.resume.empty()) CK(b2_load_checkpoint(h, o.resume.c_str()));
  if (world > 1) {
    char id[B2_COMM_ID_BYTES];
    exchange_id(o, rank, id);
    CK(b2_comm_init(h, rank, world, id));
  }
  if (!o.quiet && rank == 0)
    std::printf("control_vars: %d\nobs_vars: %d\nNepisodes: %d\nN_timestep: %d\nMax_timestep: %d\nupdateTimestep: %d\n"
                "envs: %ld (x%d GPUs)  records per update: %ld  episodes per env: %ld\n",
                c.control_vars, c.obs_vars, c.n_episodes, c.n_timestep, c.max_timestep, c.update_timestep, n_global,
                world, n_global * (long)T, episodes_per_env);

  // per-env episode logs in the reference's format (src/main.cpp:15-16)
  const long n_logged = std::min(o.log_envs, n_local);
  std::vector<std::ofstream> logs((size_t)n_logged);
  for (long e = 0; e < n_logged; e++) {
    const long myid = (long)c.env_offset + e + 1;  // env ranks start at 1 (rank 0 is the learner)
    logs[(size_t)e].open(o.log_dir + "/proc" + std::to_string(myid) + "_log.txt");
  }

  const int max_eps = (int)eps_per_update + 1;
  std::vector<int32_t> counts((size_t)n_local);
  std::vector<float> ep_returns((size_t)n_local * max_eps);
  long episodes_done = 0;  // per env
  long long total_steps = 0;
  const auto t_start = std::chrono::steady_clock::now();
  for (long it = 0; (o.iters < 0 || it < o.iters) && episodes_done < episodes_per_env && total_steps < c.max_timestep; it++) {
    // the last cycle may be shorter: never run more episodes than Nepisodes/(nprocs-1)
    const long eps_now = std::min(eps_per_update, episodes_per_env - episodes_done);
    const int T_now = (int)(eps_now * c.n_timestep);
    b2_update_stats st;
    CK(b2_rollout(h, T_now, nullptr));
    CK(b2_update_resident(h, &st));                        // ppo.update(mergedMemory), src/main.cpp:304
    CK(b2_episode_stats(h, counts.data(), ep_returns.data(), max_eps, n_local));
    double mean = 0;
    long n_ep = 0;
    for (long e = 0; e < n_local; e++)
      for (int i = 0; i < counts[(size_t)e] && i < max_eps; i++) {
        const float r = ep_returns[(size_t)e * max_eps + i];
        mean += r;
        n_ep++;
        if (e < n_logged) {                                // src/main.cpp:139-142
          std::ofstream& f = logs[(size_t)e];
          f << "myid: " << ((long)c.env_offset + e + 1) << "\t episode:" << (episodes_done + i) << "\t reward:"
            << std::fixed << std::setprecision(8) << (double)r << std::endl;
        }
      }
    episodes_done += eps_now;
    total_steps += (long long)n_global * T_now;
    if (!o.quiet && rank == 0) {
      const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
      std::printf("update %ld  episodes/env %ld  mean episode reward %.3f  loss %.6f (surrogate %.6f, value %.6f)  "
                  "clip %.3f  adam step %lld  %.3e env-steps/s\n",
                  it + 1, episodes_done, n_ep ? mean / n_ep : 0.0, st.loss, st.loss1, st.loss2, st.clip_frac,
                  (long long)st.adam_step, (double)total_steps / secs);
    }
  }
  if (!o.save.empty() && rank == 0) CK(b2_save_checkpoint(h, o.save.c_str()));
  if (!o.quiet && rank == 0) std::printf("done: %lld transitions, %lld kernel launches\n", total_steps, (long long)b2_kernel_launches(h));
  CK(b2_destroy(h));
  if (world > 1 && rank == 0) std::remove(o.rendezvous.c_str());
  return 0;
}
