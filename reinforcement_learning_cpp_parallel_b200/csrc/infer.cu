// infer.cu -- K6: stand-alone batched action query / value estimate on caller-supplied observations,
// plus the AoS<->SoA observation transposes used at the host boundary.
// Replaces getAction -> select_action -> act (src/ppo_nn.cpp:308-331,196-203,99-125) and
// policy.critic->forward (src/ppo_nn.cpp:189,227) for callers that keep their own environments.
#include "engine.h"
#include "mlp.cuh"

namespace b2 {

constexpr int INFER_BLOCK = 128;

template <class N>
__global__ void __launch_bounds__(INFER_BLOCK, 4) k_act(const float* __restrict__ params, const float* __restrict__ obs,
                                                     const float* __restrict__ eps_in, int64_t n, uint64_t seed,
                                                     uint64_t call, float* __restrict__ action,
                                                     float* __restrict__ logp, float* __restrict__ mu_out,
                                                     HyperDev hp, int* status) {
  extern __shared__ __align__(16) float sm[];
  float* hs = sm + NetSmem<N>::FLOATS;
  stage_net<N>(sm, params);
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x[N::S];
#pragma unroll
    for (int c = 0; c < N::S; c++) x[c] = obs[i * N::S + c];
    const float mu = act_tanh(mlp_row<N>(sm, hs, x));
    const float eps = eps_in ? eps_in[i] : gauss_noise(seed ^ 0xA5A5A5A5A5A5A5A5ull, (uint64_t)i, call);
    const float a = (float)__dadd_rn((double)mu, __dmul_rn(hp.sigma_d, (double)eps));  // src/ppo_nn.cpp:107-108
    action[i] = a;
    logp[i] = gauss_logprob(a, mu, hp.inv_var, hp.half_log_norm);                      // :115
    if (mu_out) mu_out[i] = mu;
    if (!isfinite(a)) atomicOr(status, 1);
  }
}

template <class N>
__global__ void __launch_bounds__(INFER_BLOCK, 4) k_value(const float* __restrict__ params, const float* __restrict__ obs,
                                                       int64_t n, float* __restrict__ v) {
  extern __shared__ __align__(16) float sm[];
  float* hs = sm + NetSmem<N>::FLOATS;
  stage_net<N>(sm, params + N::SIZE);  // critic block follows the actor block
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x[N::S];
#pragma unroll
    for (int c = 0; c < N::S; c++) x[c] = obs[i * N::S + c];
    v[i] = mlp_row<N>(sm, hs, x);
  }
}

__global__ void k_aos_to_soa(const float* __restrict__ aos, float* __restrict__ soa, int64_t cap, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * OBS) return;
  int64_t k = i / OBS;
  int c = (int)(i % OBS);
  soa[c * cap + k] = aos[i];
}
__global__ void k_soa_to_aos(const float* __restrict__ soa, float* __restrict__ aos, int64_t cap, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * OBS) return;
  int64_t k = i / OBS;
  int c = (int)(i % OBS);
  aos[i] = soa[c * cap + k];
}
__global__ void k_flush(float* buf, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    buf[i] = (float)i;
}

static int grid_for(int64_t n, int block) {
  int64_t g = (n + block - 1) / block;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

cudaError_t launch_act(const float* params, const float* obs_aos, const float* eps, int64_t n, uint64_t seed,
                       uint64_t call, float* action, float* logp, float* mu, const HyperDev& hp, int* status,
                       cudaStream_t st) {
  using N = NetDefault;
  k_act<N><<<grid_for(n, INFER_BLOCK), INFER_BLOCK, mlp_smem_bytes<N>(INFER_BLOCK), st>>>(params, obs_aos, eps, n, seed, call,
                                                                               action, logp, mu, hp, status);
  return cudaGetLastError();
}
cudaError_t launch_value(const float* params, const float* obs_aos, int64_t n, float* v, cudaStream_t st) {
  using N = NetDefault;
  k_value<N><<<grid_for(n, INFER_BLOCK), INFER_BLOCK, mlp_smem_bytes<N>(INFER_BLOCK), st>>>(params, obs_aos, n, v);
  return cudaGetLastError();
}
cudaError_t launch_aos_to_soa(const float* a, float* s, int64_t cap, int64_t n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_aos_to_soa<<<(unsigned)((n * OBS + 255) / 256), 256, 0, st>>>(a, s, cap, n);
  return cudaGetLastError();
}
cudaError_t launch_soa_to_aos(const float* s, float* a, int64_t cap, int64_t n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_soa_to_aos<<<(unsigned)((n * OBS + 255) / 256), 256, 0, st>>>(s, a, cap, n);
  return cudaGetLastError();
}
cudaError_t launch_flush(float* buf, int64_t n, cudaStream_t st) {
  k_flush<<<148 * 8, 256, 0, st>>>(buf, n);
  return cudaGetLastError();
}

}  // namespace b2
