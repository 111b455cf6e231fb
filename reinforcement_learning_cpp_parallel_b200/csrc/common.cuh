// common.cuh -- shared device/host helpers of the B200 PPO engine (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef __CUDACC__
#define __host__
#define __device__
#define __forceinline__ inline
#endif

namespace b2 {

// ---------------------------------------------------------------------------------------------
// Parameter layout == policy.parameters() of the reference (include/ppo_nn.h:83-123):
//   actor : W1[H1][S] b1[H1] W2[H2][H1] b2[H2] W3[1][H2] b3[1]      critic: the same, no final tanh
// ---------------------------------------------------------------------------------------------
template <int S_, int H1_, int H2_>
struct Net {
  static constexpr int S = S_, H1 = H1_, H2 = H2_;
  static constexpr int oW1 = 0;
  static constexpr int ob1 = oW1 + H1 * S;
  static constexpr int oW2 = ob1 + H1;
  static constexpr int ob2 = oW2 + H2 * H1;
  static constexpr int oW3 = ob2 + H2;
  static constexpr int ob3 = oW3 + H2;
  static constexpr int SIZE = ob3 + 1;  // one net, A = 1 (src/ppo_nn.cpp:107: action_dim == 1 only)
  static constexpr int P = 2 * SIZE;    // actor + critic
};
using NetDefault = Net<6, 64, 32>;  // 2561 per net, 5122 total

constexpr int OBS = 6;  // include/cart_pole.h:120-131

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, counter = (env id, step), key = seed.  One call per env-step gives the action
// noise; streams are a pure function of the GLOBAL env id, so any sharding reproduces them.
// (The CPU checker under oracle/ restates the same generator so tests can compare draw for draw.)
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t out[4]) {
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
  for (int r = 0; r < 10; r++) {
#ifdef __CUDA_ARCH__
    uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
#else
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t h0 = (uint32_t)(p0 >> 32), l0 = (uint32_t)p0, h1 = (uint32_t)(p1 >> 32), l1 = (uint32_t)p1;
#endif
    uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
    c0 = n0; c1 = l1; c2 = n2; c3 = l0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

#ifdef __CUDACC__
__device__ __forceinline__ float gauss_noise(uint64_t seed, uint64_t env, uint64_t step) {
  uint32_t x[4];
  philox4x32_10((uint32_t)env, (uint32_t)(env >> 32), (uint32_t)step, (uint32_t)(step >> 32),
                (uint32_t)seed, (uint32_t)(seed >> 32), x);
  float u1 = ((float)(x[0] >> 9) + 0.5f) * 1.1920928955078125e-07f;  // (0,1), exact in f32
  float u2 = ((float)(x[1] >> 9) + 0.5f) * 1.1920928955078125e-07f;
  float r = sqrtf(-2.0f * logf(u1));
  return r * cospif(2.0f * u2);  // Box-Muller, first variate
}

// Gaussian log-density for k = 1 exactly as multivariateLogProb evaluates it (src/ppo_nn.cpp:20-49):
//   numer = -0.5 * diff * (inv_var * diff);  denom = (log(var) + log(2 pi)) / 2
__device__ __forceinline__ float gauss_logprob(float a, float mu, float inv_var, float half_log_norm) {
  float diff = __fsub_rn(a, mu);
  float numer = __fmul_rn(-0.5f, __fmul_rn(diff, __fmul_rn(inv_var, diff)));
  return __fsub_rn(numer, half_log_norm);
}
#endif

struct LogProbConst {
  float inv_var, half_log_norm, sigma, entropy_term;  // entropy_term = loss3 = -0.01 * H (Q8)
};

}  // namespace b2

#define B2_CUDA_CHECK(expr)                                                              \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) return b2::set_error(B2_ERR_CUDA, "%s: %s (%s:%d)", #expr,    \
                                                cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)
