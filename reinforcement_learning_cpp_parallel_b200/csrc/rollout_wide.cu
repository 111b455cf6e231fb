// rollout_wide.cu -- device-resident rollout for WIDE policies (hidden widths the fused k_rollout is not built for,
// e.g. the 2 x 256 nets of BASELINE configs[4]), so that such a policy can collect its own data on the device.
//
// Same semantics as k_rollout (rollout.cu: env_run src/main.cpp:12-148, getAction src/ppo_nn.cpp:308-331,99-125), same
// Philox streams (counter = global env id, step), same trajectory layout -- but one environment step is three launches
// instead of a register-resident loop, because a 256-wide actor does not fit a thread's registers:
//   k_rw_observe   state -> f32 observation planes of step t                        (getobs, cart_pole.h:120-133)
//   k_wide_forward actor(obs) -> mu[t]                                              (wide_update.cu, FP32, 16 rows per block)
//   k_rw_step      noise, action, log-prob, force scaling, advance, reward, TIMEUP / TRUNC / TERMINATE, reset, logging
// It is a functional path, not a throughput path (the physics is ~1 % of a 2 x 256 forward pass).
#include "cartpole.cuh"
#include "engine.h"
#include "wide.h"

namespace b2 {

namespace {

template <typename Real>
__global__ void __launch_bounds__(128) k_rw_observe(const RolloutArgs a, int t) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= a.n_envs) return;
  const int64_t NE = a.n_envs, k = (int64_t)t * NE + e;
  using S = typename StateOf<Real>::type;   // double for B2_PREC_MIXED
  S u[4];
#pragma unroll
  for (int j = 0; j < 4; j++) u[j] = (S)a.state[j * NE + e];
  float o[6];
  CartPoleDyn<Real>::observe(u, o);
#pragma unroll
  for (int c = 0; c < 6; c++) a.tr.obs[c * a.tr.cap + k] = o[c];
}

template <typename Real, bool VAR>
__global__ void __launch_bounds__(128) k_rw_step(const RolloutArgs a, int t, int trace) {
  using Dyn = CartPoleDyn<Real>;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= a.n_envs) return;
  const int64_t NE = a.n_envs, k = (int64_t)t * NE + e;
  using S = typename StateOf<Real>::type;   // double for B2_PREC_MIXED
  S u[4];
#pragma unroll
  for (int j = 0; j < 4; j++) u[j] = (S)a.state[j * NE + e];
  int es = a.ep_step[e];
  float er = a.ep_ret[e];
  int ecount = (t == 0) ? 0 : a.ep_count[e];
  const float mu = a.tr.mu[k];                                            // actor output of this step (k_wide_forward)
  float act;
  if (a.forced_action) {
    act = a.forced_action[k];
  } else {
    const float eps = gauss_noise(a.seed, a.env_offset + (uint64_t)e, a.step0 + (uint64_t)t);
    act = (float)__dadd_rn((double)mu, __dmul_rn(a.hp.sigma_d, (double)eps));   // src/ppo_nn.cpp:107-108
  }
  const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);    // :115
  if (!isfinite(act)) atomicOr(a.status, 1);                                    // src/main.cpp:50-55
  bool terminal = false;
  float r;
  if constexpr (VAR) {
    const CartPoleVar::Adv adv = CartPoleVar::advance(make_double4((double)u[0], (double)u[1], (double)u[2], (double)u[3]),
                                                      scale_action(act), a.env_flags, es + 1);
    u[0] = (S)adv.u.x; u[1] = (S)adv.u.y; u[2] = (S)adv.u.z; u[3] = (S)adv.u.w;
    terminal = adv.term != 0;                                             // src/main.cpp:81-85
    const double ud[4] = {adv.u.x, adv.u.y, adv.u.z, adv.u.w};
    r = CartPoleVar::reward(a.env_flags, ud);
  } else {
    Dyn::advance(u, (S)scale_action(act));                             // src/main.cpp:57-71
    r = Dyn::reward(u);                                                   // :73
  }
  es += 1;
  er += r;
  const bool timeup = !terminal && (es == a.n_timestep);                  // :86-88
  const uint8_t fl = terminal ? (uint8_t)B2_FLAG_TERMINAL
                              : (timeup ? (uint8_t)B2_FLAG_TIMEUP : ((t == a.T - 1) ? (uint8_t)B2_FLAG_TRUNC : (uint8_t)0));
  a.tr.action[k] = act;
  a.tr.logp[k] = lp;
  a.tr.reward[k] = r;
  a.tr.flags[k] = fl;
  if (trace) *reinterpret_cast<double4*>(a.tr.state_trace + 4 * k) = make_double4((double)u[0], (double)u[1], (double)u[2], (double)u[3]);
  if (timeup || terminal) {                                               // :137-142 log, :27 reset
    if (ecount < a.max_eps) a.ep_log[(int64_t)ecount * NE + e] = er;
    ecount++;
    if constexpr (VAR) {
      double u0[4];
      CartPoleVar::init_state(a.env_flags, a.seed, a.env_offset + (uint64_t)e, a.step0 + (uint64_t)t + 1, u0);
      u[0] = (S)u0[0]; u[1] = (S)u0[1]; u[2] = (S)u0[2]; u[3] = (S)u0[3];
    } else {
      u[0] = u[1] = u[2] = u[3] = (S)0;                                // cart_pole.h:86
    }
    es = 0;
    er = 0.0f;
  }
#pragma unroll
  for (int j = 0; j < 4; j++) a.state[j * NE + e] = (double)u[j];
  a.ep_step[e] = es;
  a.ep_ret[e] = er;
  a.ep_count[e] = ecount;
}

}  // namespace

// steps [a.t_begin, a.t_end) of a T-step rollout with a wide actor; *launches += kernels launched
cudaError_t launch_rollout_wide(const RolloutArgs& a, const NetDims& d, int precision, bool trace, cudaStream_t st,
                                int64_t* launches) {
  if (a.env_flags && precision != B2_PREC_F64) return cudaErrorInvalidValue;
  if (a.obs_aos || a.obs4_aos || a.h_action || a.h_logp || a.h_mu || a.h_reward || a.h_flags) return cudaErrorInvalidValue;   // k_rollout only
  const unsigned grid = (unsigned)((a.n_envs + 127) / 128);
  cudaError_t e;
  for (int t = a.t_begin; t < a.t_end; t++) {
    if (precision != B2_PREC_F32) k_rw_observe<double><<<grid, 128, 0, st>>>(a, t);   // (mixed: double state, same observe)
    else k_rw_observe<float><<<grid, 128, 0, st>>>(a, t);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    const int64_t k0 = (int64_t)t * a.n_envs;
    if ((e = launch_wide_forward(d, a.params, a.tr.obs + k0, a.tr.cap, 1, nullptr, a.n_envs, 1, a.tr.mu + k0, st)) != cudaSuccess) return e;
    if (a.env_flags) k_rw_step<double, true><<<grid, 128, 0, st>>>(a, t, trace ? 1 : 0);
    else if (precision == B2_PREC_F64) k_rw_step<double, false><<<grid, 128, 0, st>>>(a, t, trace ? 1 : 0);
    else if (precision == B2_PREC_MIXED) k_rw_step<MixedPrec, false><<<grid, 128, 0, st>>>(a, t, trace ? 1 : 0);
    else k_rw_step<float, false><<<grid, 128, 0, st>>>(a, t, trace ? 1 : 0);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    *launches += 3;
  }
  return cudaSuccess;
}

}  // namespace b2
