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
  Real u[4];
#pragma unroll
  for (int j = 0; j < 4; j++) u[j] = (Real)state[j * n + e];
  CartPoleDyn<Real>::advance(u, (Real)force[e]);   // F = action[0]; step++ (:108-109)
#pragma unroll
  for (int j = 0; j < 4; j++) state[j * n + e] = (double)u[j];
  ep_step[e] += 1;
}

// obs[e][6] = x, xdot, thetadot, theta, cos, sin in double (getobs :123-130); reward[e] (:142)
__global__ void __launch_bounds__(128) k_env_observe(const double* __restrict__ state, double* __restrict__ obs,
                                                     double* __restrict__ reward, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const double u[4] = {state[e], state[n + e], state[2 * n + e], state[3 * n + e]};
  if (obs) {
    double s, c;
    sincos(u[2], &s, &c);
    double* o = obs + 6 * e;
    o[0] = u[0]; o[1] = u[1]; o[2] = u[3]; o[3] = u[2]; o[4] = c; o[5] = s;
  }
  if (reward) reward[e] = (double)CartPoleDyn<double>::reward(u);
}

cudaError_t launch_env_step(double* state, int32_t* ep_step, const double* force, int64_t n, int precision,
                            cudaStream_t st) {
  const unsigned grid = (unsigned)((n + 127) / 128);
  if (precision == B2_PREC_F64) k_env_step<double><<<grid, 128, 0, st>>>(state, ep_step, force, n);
  else k_env_step<float><<<grid, 128, 0, st>>>(state, ep_step, force, n);
  return cudaGetLastError();
}
cudaError_t launch_env_observe(const double* state, double* obs, double* reward, int64_t n, cudaStream_t st) {
  k_env_observe<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(state, obs, reward, n);
  return cudaGetLastError();
}

}  // namespace b2
