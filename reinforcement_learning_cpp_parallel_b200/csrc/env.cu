// env.cu -- the cart_pole.h-shaped batched environment entry points: advance every resident env by one
// step under caller-supplied forces, and read observations/rewards back.  No policy, no trajectory.
// Replaces CartPole::advance (include/cart_pole.h:107-118), getobs (:120-133) and getReward (:135-147) for
// callers that drive the simulator themselves (include/compat/cart_pole.h); the fused training path uses
// k_rollout instead.
#include "cartpole.cuh"
#include "engine.h"

namespace b2 {

template <typename Real>
__global__ void __launch_bounds__(128) k_env_step(double* __restrict__ state, int32_t* __restrict__ ep_step,
                                                  const double* __restrict__ force, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  using S = typename StateOf<Real>::type;   // double for B2_PREC_MIXED
  S u[4];
#pragma unroll
  for (int j = 0; j < 4; j++) u[j] = (S)state[j * n + e];
  CartPoleDyn<Real>::advance(u, (S)force[e]);   // F = action[0]; step++ (:108-109)
#pragma unroll
  for (int j = 0; j < 4; j++) state[j * n + e] = (double)u[j];
  ep_step[e] += 1;
}

// the same under the opt-in variants (B2_ENV_*): reference-shaped double physics, advance() may stop early and return 1
__global__ void __launch_bounds__(128) k_env_step_var(double* __restrict__ state, int32_t* __restrict__ ep_step,
                                                      const double* __restrict__ force, int64_t n, int env_flags,
                                                      uint8_t* __restrict__ terminated) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const int step = ep_step[e] + 1;                                      // step++ (:109)
  const CartPoleVar::Adv r = CartPoleVar::advance(make_double4(state[e], state[n + e], state[2 * n + e], state[3 * n + e]),
                                                  force[e], env_flags, step);
  state[e] = r.u.x; state[n + e] = r.u.y; state[2 * n + e] = r.u.z; state[3 * n + e] = r.u.w;
  ep_step[e] = step;
  terminated[e] = (uint8_t)r.term;
}

// CartPole::reset on every env (:77-90): zero state, or the random initial state of B2_ENV_RANDOM_INIT
__global__ void __launch_bounds__(128) k_env_init(double* __restrict__ state, int64_t n, int env_flags, uint64_t seed,
                                                  uint64_t env_offset, uint64_t tick) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  double u[4];
  CartPoleVar::init_state(env_flags, seed, env_offset + (uint64_t)e, tick, u);
#pragma unroll
  for (int j = 0; j < 4; j++) state[j * n + e] = u[j];
}

// obs[e][6] = x, xdot, thetadot, theta, cos, sin in double (getobs :123-130); reward[e] (:142)
__global__ void __launch_bounds__(128) k_env_observe(const double* __restrict__ state, double* __restrict__ obs,
                                                     double* __restrict__ reward, int64_t n, int env_flags) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const double u[4] = {state[e], state[n + e], state[2 * n + e], state[3 * n + e]};
  if (obs) {
    double s, c;
    sincos(u[2], &s, &c);
    double* o = obs + 6 * e;
    o[0] = u[0]; o[1] = u[1]; o[2] = u[3]; o[3] = u[2]; o[4] = c; o[5] = s;
  }
  if (reward) reward[e] = (double)CartPoleVar::reward(env_flags, u);
}

cudaError_t launch_env_step(double* state, int32_t* ep_step, const double* force, int64_t n, int precision,
                            int env_flags, uint8_t* terminated, cudaStream_t st) {
  const unsigned grid = (unsigned)((n + 127) / 128);
  if (env_flags) {
    k_env_step_var<<<grid, 128, 0, st>>>(state, ep_step, force, n, env_flags, terminated);
    return cudaGetLastError();
  }
  cudaError_t e = cudaMemsetAsync(terminated, 0, (size_t)n, st);        // the shipped advance() returns 0 (:117)
  if (e != cudaSuccess) return e;
  if (precision == B2_PREC_F64) k_env_step<double><<<grid, 128, 0, st>>>(state, ep_step, force, n);
  else if (precision == B2_PREC_MIXED) k_env_step<MixedPrec><<<grid, 128, 0, st>>>(state, ep_step, force, n);
  else k_env_step<float><<<grid, 128, 0, st>>>(state, ep_step, force, n);
  return cudaGetLastError();
}
cudaError_t launch_env_observe(const double* state, double* obs, double* reward, int64_t n, int env_flags, cudaStream_t st) {
  k_env_observe<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(state, obs, reward, n, env_flags);
  return cudaGetLastError();
}
cudaError_t launch_env_init(double* state, int64_t n, int env_flags, uint64_t seed, uint64_t env_offset, uint64_t tick,
                            cudaStream_t st) {
  k_env_init<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(state, n, env_flags, seed, env_offset, tick);
  return cudaGetLastError();
}

}  // namespace b2
