// rollout.cu -- K1: fused batched cart-pole rollout.
//
// One thread owns one environment for the whole T-step horizon: the 4-component state, the episode
// counters and the RK accumulators never leave registers.  Per step the thread (1) forms the f32
// observation, (2) runs the actor MLP out of shared-memory weights, (3) draws its Philox Gaussian,
// (4) evaluates the log-density, (5) integrates 50 x 6 RK46-NL stages, (6) emits the transition with
// one coalesced store per field (SoA, time-major), (7) applies the TIMEUP/reset rule.
// There is no per-step message and no host involvement: this replaces env_run (src/main.cpp:12-148),
// respond_to_env/getAction (src/main.cpp:165-188, src/ppo_nn.cpp:308-331,99-125) and the MPI traffic.
//
// Roofline: FP64 (B2_PREC_F64) or FP32 FMA issue, not HBM -- 41 B are written per env-step against
// ~12k-24k arithmetic instructions.
#include "cartpole.cuh"
#include "engine.h"
#include "mlp.cuh"

namespace b2 {

constexpr int ROLLOUT_BLOCK = 64;   // 65,536 envs -> 1024 CTAs: 6.9 per SM, all resident at <= 144 regs

template <typename Real, class N, bool TRACE>
__global__ void __launch_bounds__(ROLLOUT_BLOCK, 7) k_rollout(const RolloutArgs a) {
  extern __shared__ __align__(16) float sm[];
  float* hs = sm + NetSmem<N>::FLOATS;  // hidden-1 scratch, one column per thread
  stage_net<N>(sm, a.params);  // actor only: V is not needed while acting (src/ppo_nn.cpp:100)
  __syncthreads();
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= a.n_envs) return;

  using Dyn = CartPoleDyn<Real>;
  const int64_t NE = a.n_envs;
  Real u[4];
#pragma unroll
  for (int j = 0; j < 4; j++) u[j] = (Real)a.state[j * NE + e];
  int es = a.ep_step[e];
  float er = a.ep_ret[e];
  int ecount = 0;
  bool bad = false;
  const int64_t cap = a.tr.cap;

  for (int t = 0; t < a.T; t++) {
    const int64_t k = (int64_t)t * NE + e;
    float o[6];
    Dyn::observe(u, o);
    const float mu = act_tanh(mlp_row<N>(sm, hs, o));                       // ppo_nn.h:98-102
    float act;
    if (a.forced_action) {
      act = a.forced_action[k];
    } else {
      const float eps = gauss_noise(a.seed, a.env_offset + (uint64_t)e, a.step0 + (uint64_t)t);
      act = (float)__dadd_rn((double)mu, __dmul_rn(a.hp.sigma_d, (double)eps));  // src/ppo_nn.cpp:107-108
    }
    const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);   // :115
    bad |= !isfinite(act);                                               // src/main.cpp:50-55

    Dyn::advance(u, (Real)scale_action(act));                            // src/main.cpp:57-71
    const float r = Dyn::reward(u);                                      // :73
    es += 1;
    er += r;
    const bool timeup = (es == a.n_timestep);                            // :86-88 (TERMINATE never fires, Q1)
    const uint8_t fl = timeup ? (uint8_t)B2_FLAG_TIMEUP : ((t == a.T - 1) ? (uint8_t)B2_FLAG_TRUNC : (uint8_t)0);

#pragma unroll
    for (int c = 0; c < 6; c++) a.tr.obs[c * cap + k] = o[c];
    a.tr.action[k] = act;
    a.tr.logp[k] = lp;
    a.tr.mu[k] = mu;
    a.tr.reward[k] = r;
    a.tr.flags[k] = fl;
    if (TRACE) {
      double4 s4 = make_double4((double)u[0], (double)u[1], (double)u[2], (double)u[3]);
      *reinterpret_cast<double4*>(a.tr.state_trace + 4 * k) = s4;
    }
    if (timeup) {                                                        // :137-142 log, :27 reset
      if (ecount < a.max_eps) a.ep_log[(int64_t)ecount * NE + e] = er;
      ecount++;
      u[0] = u[1] = u[2] = u[3] = (Real)0;                               // cart_pole.h:86
      es = 0;
      er = 0.0f;
    }
  }
#pragma unroll
  for (int j = 0; j < 4; j++) a.state[j * NE + e] = (double)u[j];
  a.ep_step[e] = es;
  a.ep_ret[e] = er;
  a.ep_count[e] = ecount;
  if (bad) atomicOr(a.status, 1);
}

template <typename Real, bool TRACE>
static cudaError_t launch_t(const RolloutArgs& a, cudaStream_t st) {
  using N = NetDefault;
  int grid = (int)((a.n_envs + ROLLOUT_BLOCK - 1) / ROLLOUT_BLOCK);
  k_rollout<Real, N, TRACE><<<grid, ROLLOUT_BLOCK, mlp_smem_bytes<N>(ROLLOUT_BLOCK), st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_rollout(const RolloutArgs& a, int precision, bool trace, cudaStream_t st, int h1, int h2) {
  if (h1 != NetDefault::H1 || h2 != NetDefault::H2) return cudaErrorInvalidValue;
  if (precision == B2_PREC_F64) return trace ? launch_t<double, true>(a, st) : launch_t<double, false>(a, st);
  return trace ? launch_t<float, true>(a, st) : launch_t<float, false>(a, st);
}

}  // namespace b2
