/* oracle/oracle.h -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (libb200ppo.so) never links, loads or calls it.
 *
 * Parity status: PINNED.  The reference ships no golden vectors, so the pins were
 * minted from the reference's own unmodified sources by oracle/ref_harness.cpp (tests/golden/*)
 * and tests/test_oracle_vs_golden.py checks every function below against them.
 *
 * All file:line citations are relative to /root/reference.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- network shape: two MLPs S->H1->H2->{A,1}, include/ppo_nn.h:80-140 ---- */
typedef struct {
  int S, H1, H2, A;
} orc_dims;
int orc_net_params(const orc_dims* d, int out); /* floats in one net                         */
int orc_n_params(const orc_dims* d);            /* actor + critic, policy.parameters() order */

/* ---- environment: include/cart_pole.h ---- */
void orc_cartpole_diff(const double u[4], double F, double du[4]);   /* :149-170 (SWINGUP 0)  */
void orc_rk46_nl(double u[4], double dt, double F);                  /* :23-45                */
void orc_cartpole_advance(double u[4], double F);                    /* :107-118, 50 substeps */
void orc_cartpole_obs(const double u[4], float obs[6]);              /* :120-133 + main.cpp:79 */
double orc_cartpole_reward(const double u[4]);                       /* :135-147              */
double orc_scale_action(float a);                                    /* src/main.cpp:57-60,69 */

/* ---- environment variants the header documents but ships disabled, opt-in:
 *   ORC_ENV_SWINGUP      the `#if SWINGUP` branches: Diff :154-159, getReward :136-139, is_over :100-101,
 *                        reset range +-0.1 :80-81
 *   ORC_ENV_RANDOM_INIT  reset draws u ~ U(-r, r)^4, r = 0.05 (0.1 SWINGUP): the commented-out line :85
 *                        (live in archived/single_environment/cart_pole.h:87).  The reference seeds mt19937 from
 *                        std::random_device (not reproducible); here the draw is Philox, keyed like orc_gauss.
 *   ORC_ENV_EARLY_TERM   advance returns 1 as soon as is_over() holds after a substep: the commented-out
 *                        line :115 (live in archived/.../cart_pole.h:119); env_run turns it into TERMINATE
 *                        (src/main.cpp:81-85) and the learner restarts the return there (src/ppo_nn.cpp:229-230) */
#define ORC_ENV_SWINGUP 1
#define ORC_ENV_RANDOM_INIT 2
#define ORC_ENV_EARLY_TERM 4
void orc_cartpole_diff_v(int env_flags, const double u[4], double F, double du[4]);
int orc_cartpole_is_over(int env_flags, const double u[4], int step);             /* :99-105 */
int orc_cartpole_advance_v(int env_flags, double u[4], int step, double F);       /* step = value after step++ (:109) */
double orc_cartpole_reward_v(int env_flags, const double u[4]);
void orc_cartpole_init_state(int env_flags, uint64_t seed, uint64_t env, uint64_t tick, double u[4]);

/* ---- policy: src/ppo_nn.cpp ---- */
/* one net forward; h1/h2 may be NULL.  final_tanh=1 for the actor (ppo_nn.h:98-102). */
void orc_mlp_forward(const orc_dims* d, const float* net, int out, int final_tanh, const float* obs,
                     float* h1, float* h2, float* y);
float orc_logprob(float a, float mu, double sigma);                  /* :8-54, k = 1          */
float orc_entropy_const(double sigma);                               /* :55-65 (Q8)           */
void orc_act(const orc_dims* d, const float* params, const float* obs, int n, const float* eps,
             double sigma, float* mu, float* action, float* logp);   /* :99-125               */
void orc_value(const orc_dims* d, const float* params, const float* obs, int n, float* v); /* :189 */

/* ---- returns, src/ppo_nn.cpp:207-236.  flat order = merged order; flags: bit0 terminal,
 * bit1 timeup, bit2 horizon-truncated (engine extension, bootstraps like timeup). ---- */
void orc_returns(const orc_dims* d, const float* params, const float* obs, const float* reward,
                 const uint8_t* flags, long B, double gamma, float* R);

/* ---- loss + gradient of one epoch over samples idx[0..nb) (idx NULL = 0..nb), :245-302.
 * loss terms are means over nb; grads = d(mean loss)/d params, P floats. ---- */
typedef struct {
  double loss1, loss2, loss3, loss, clip_frac;
} orc_loss;
void orc_loss_grad(const orc_dims* d, const float* params, const float* obs, const float* action,
                   const float* old_logp, const float* R, const long* idx, long nb, double sigma,
                   double eps_clip, orc_loss* out, float* grads);

/* ---- torch::optim::Adam::step (libtorch C++ adam.cpp), options include/ppo_nn.h:166-170 ---- */
void orc_adam(float* params, const float* grads, float* m, float* v, long n, long step /*1-based*/,
              double lr, double b1, double b2, double eps);

/* ---- whole PPO::update on a flat batch, :205-307 (full batch, K epochs) ---- */
void orc_update(const orc_dims* d, float* params, float* m, float* v, long* step, const float* obs,
                const float* action, const float* old_logp, const float* reward,
                const uint8_t* flags, long B, int K, double sigma, double eps_clip, double gamma,
                double lr, double b1, double b2, orc_loss* per_epoch /*K or NULL*/,
                float* last_grads /*P or NULL*/, float* returns_out /*B or NULL*/);

/* ---- counter-based noise (engine design, no reference analogue: torch::normal uses mt19937).
 * Philox4x32-10, key=(seed lo,hi), counter=(env lo, env hi, step lo, step hi); eps = Box-Muller
 * of the first two words.  The CUDA engine must reproduce these draws. ---- */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
float orc_gauss(uint64_t seed, uint64_t env, uint64_t step);

/* ---- batched rollout in [T][N] layout (Appendix A of SURVEY.md), OpenMP over envs.
 * state: N*4 doubles (in/out), ep_step: N ints (in/out), ep_ret: N doubles (in/out).
 * forced_action NULL => sample with orc_gauss(seed, env0+e, step0+t).
 * Any output pointer may be NULL. ---- */
void orc_rollout(const orc_dims* d, const float* params, long N, int T, int n_timestep, double sigma,
                 uint64_t seed, uint64_t env0, uint64_t step0, const float* forced_action,
                 double* state, int* ep_step, double* ep_ret, float* obs, float* action,
                 float* logp, float* mu, float* reward, uint8_t* flags, double* state_trace);

/* orc_rollout with the env variants above.  TERMINATE takes precedence over TIMEUP (src/main.cpp:81-88); after either
 * the env is reset (zero state, or orc_cartpole_init_state(.., env0 + e, step0 + t + 1) with RANDOM_INIT). */
void orc_rollout_v(const orc_dims* d, const float* params, long N, int T, int n_timestep, double sigma,
                   uint64_t seed, uint64_t env0, uint64_t step0, const float* forced_action, int env_flags,
                   double* state, int* ep_step, double* ep_ret, float* obs, float* action,
                   float* logp, float* mu, float* reward, uint8_t* flags, double* state_trace);

#ifdef __cplusplus
}
#endif
#endif
