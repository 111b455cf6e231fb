// returns.cu -- K2: discounted returns as a reverse segmented scan over the trajectory buffer.
//
// Reference semantics (src/ppo_nn.cpp:207-236): walking the records backwards,
//   carry = TIMEUP ? V(obs_k) : TERMINAL ? 0 : carry ;  R_k = r_k + gamma * carry      (all f32)
// i.e. GAE(lambda=1) whose segments end at flagged records and whose bootstrap uses the critic at the
// PRE-step observation with the weights the update starts from.  Two launches:
//   k_bootstrap : V(obs_k) for the (rare) flagged records, one thread per record, critic out of smem
//   k_returns   : one thread per environment walks t = T-1..0; loads are coalesced across envs and
//                 batched 8 deep, the recurrence itself is two dependent f32 ops per record and is
//                 evaluated in exactly the reference's order (mul, then add, no contraction), so the
//                 result is bit-identical to a sequential f32 evaluation.
// Roofline: HBM, 13 B/record (r 4 + flag 1 + vboot 4 (flagged only) + R 4).
#include "engine.h"
#include "mlp.cuh"

namespace b2 {

template <class N>
__global__ void __launch_bounds__(128, 4) k_bootstrap(const float* __restrict__ params, Traj tr, int64_t B) {
  extern __shared__ __align__(16) float sm[];
  float* hs = sm + NetSmem<N>::FLOATS;
  stage_net<N>(sm, params + N::SIZE);
  __syncthreads();
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < B; k += (int64_t)gridDim.x * blockDim.x) {
    if (tr.flags[k] & (B2_FLAG_TIMEUP | B2_FLAG_TRUNC)) {
      float x[N::S];
#pragma unroll
      for (int c = 0; c < N::S; c++) x[c] = tr.obs[c * tr.cap + k];
      tr.vboot[k] = mlp_row<N>(sm, hs, x);  // src/ppo_nn.cpp:227
    }
  }
}

__global__ void __launch_bounds__(128) k_returns(Traj tr, float gamma) {
  const int64_t NE = tr.n_envs;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= NE) return;
  float carry = 0.0f;  // src/ppo_nn.cpp:215
  int64_t t = tr.horizon - 1;
  constexpr int U = 8;
  for (; t >= U - 1; t -= U) {
    float r[U];
    uint8_t f[U];
#pragma unroll
    for (int i = 0; i < U; i++) {
      const int64_t k = (t - i) * NE + e;
      r[i] = tr.reward[k];
      f[i] = tr.flags[k];
    }
#pragma unroll
    for (int i = 0; i < U; i++) {
      const int64_t k = (t - i) * NE + e;
      if (f[i] & (B2_FLAG_TIMEUP | B2_FLAG_TRUNC)) carry = tr.vboot[k];  // :226-228
      else if (f[i] & B2_FLAG_TERMINAL) carry = 0.0f;                    // :229-230
      carry = __fadd_rn(r[i], __fmul_rn(gamma, carry));                  // :231
      tr.ret[k] = carry;
    }
  }
  for (; t >= 0; t--) {
    const int64_t k = t * NE + e;
    const uint8_t f = tr.flags[k];
    if (f & (B2_FLAG_TIMEUP | B2_FLAG_TRUNC)) carry = tr.vboot[k];
    else if (f & B2_FLAG_TERMINAL) carry = 0.0f;
    carry = __fadd_rn(tr.reward[k], __fmul_rn(gamma, carry));
    tr.ret[k] = carry;
  }
}

cudaError_t launch_returns(const float* params, const Traj& tr, const HyperDev& hp, cudaStream_t st, int* n_launches,
                           bool skip_bootstrap) {
  using N = NetDefault;
  const int64_t B = tr.n_envs * tr.horizon;
  if (B == 0) return cudaSuccess;
  int64_t g = (B + 127) / 128;
  if (g > 148 * 16) g = 148 * 16;
  cudaError_t e = cudaSuccess;
  if (!skip_bootstrap) {   // wide nets bootstrap through wide_bootstrap (wide_update.cu)
    k_bootstrap<N><<<(int)g, 128, mlp_smem_bytes<N>(128), st>>>(params, tr, B);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (n_launches) *n_launches += 1;
  }
  k_returns<<<(int)((tr.n_envs + 127) / 128), 128, 0, st>>>(tr, hp.gamma);
  if (n_launches) *n_launches += 1;
  return cudaGetLastError();
}

}  // namespace b2
