// mlp.cuh -- thread-per-row MLP forward with weights staged in shared memory.
//
// Hidden widths 64/32 are far too small for tcgen05 tiles to pay off (K = 6/64/32, N = 64/32/1),
// and the parity gate (1e-6 on mu/V) rules out TF32/BF16 inputs, so inference runs as FP32 FMA:
// every lane owns one observation row, weights sit in shared memory in k-major (transposed) order
// so that one broadcast LDS.128 feeds four FFMAs, activations stay in registers.
// Reference: ActorCritic actor/critic Sequential, include/ppo_nn.h:83-123.
#pragma once
#include "common.cuh"

namespace b2 {

// smem image of one net, k-major:  W1T[S][H1] b1[H1] W2T[H1][H2] b2[H2] W3[H2] b3[1](+pad)
template <class N>
struct NetSmem {
  static constexpr int oW1T = 0;
  static constexpr int ob1 = oW1T + N::S * N::H1;
  static constexpr int oW2T = ob1 + N::H1;
  static constexpr int ob2 = oW2T + N::H1 * N::H2;
  static constexpr int oW3 = ob2 + N::H2;
  static constexpr int ob3 = oW3 + N::H2;
  static constexpr int FLOATS = ((ob3 + 1 + 3) / 4) * 4;
  static constexpr int BYTES = FLOATS * 4;
};

// cooperative load: global row-major params of one net -> k-major smem image
template <class N>
__device__ __forceinline__ void stage_net(float* sm, const float* __restrict__ g) {
  using L = NetSmem<N>;
  for (int i = threadIdx.x; i < N::H1 * N::S; i += blockDim.x) {
    int j = i / N::S, s = i % N::S;
    sm[L::oW1T + s * N::H1 + j] = g[N::oW1 + i];
  }
  for (int i = threadIdx.x; i < N::H1; i += blockDim.x) sm[L::ob1 + i] = g[N::ob1 + i];
  for (int i = threadIdx.x; i < N::H2 * N::H1; i += blockDim.x) {
    int j = i / N::H1, k = i % N::H1;
    sm[L::oW2T + k * N::H2 + j] = g[N::oW2 + i];
  }
  for (int i = threadIdx.x; i < N::H2; i += blockDim.x) {
    sm[L::ob2 + i] = g[N::ob2 + i];
    sm[L::oW3 + i] = g[N::oW3 + i];
  }
  if (threadIdx.x == 0) sm[L::ob3] = g[N::ob3];
}

__device__ __forceinline__ float act_tanh(float x) { return tanhf(x); }

// one row through one net; returns the pre-final-activation output (caller applies tanh for the actor)
template <class N>
__device__ __forceinline__ float mlp_row(const float* __restrict__ sm, const float (&x)[N::S]) {
  using L = NetSmem<N>;
  float h1[N::H1];
#pragma unroll
  for (int j = 0; j < N::H1; j += 4) {
    float4 b = *reinterpret_cast<const float4*>(sm + L::ob1 + j);
    h1[j] = b.x; h1[j + 1] = b.y; h1[j + 2] = b.z; h1[j + 3] = b.w;
  }
#pragma unroll
  for (int s = 0; s < N::S; s++) {
#pragma unroll
    for (int j = 0; j < N::H1; j += 4) {
      float4 w = *reinterpret_cast<const float4*>(sm + L::oW1T + s * N::H1 + j);
      h1[j] = fmaf(w.x, x[s], h1[j]);
      h1[j + 1] = fmaf(w.y, x[s], h1[j + 1]);
      h1[j + 2] = fmaf(w.z, x[s], h1[j + 2]);
      h1[j + 3] = fmaf(w.w, x[s], h1[j + 3]);
    }
  }
#pragma unroll
  for (int j = 0; j < N::H1; j++) h1[j] = act_tanh(h1[j]);
  float h2[N::H2];
#pragma unroll
  for (int j = 0; j < N::H2; j += 4) {
    float4 b = *reinterpret_cast<const float4*>(sm + L::ob2 + j);
    h2[j] = b.x; h2[j + 1] = b.y; h2[j + 2] = b.z; h2[j + 3] = b.w;
  }
#pragma unroll
  for (int k = 0; k < N::H1; k++) {
#pragma unroll
    for (int j = 0; j < N::H2; j += 4) {
      float4 w = *reinterpret_cast<const float4*>(sm + L::oW2T + k * N::H2 + j);
      h2[j] = fmaf(w.x, h1[k], h2[j]);
      h2[j + 1] = fmaf(w.y, h1[k], h2[j + 1]);
      h2[j + 2] = fmaf(w.z, h1[k], h2[j + 2]);
      h2[j + 3] = fmaf(w.w, h1[k], h2[j + 3]);
    }
  }
  float y = sm[L::ob3];
#pragma unroll
  for (int j = 0; j < N::H2; j += 4) {
    float4 w = *reinterpret_cast<const float4*>(sm + L::oW3 + j);
    y = fmaf(w.x, act_tanh(h2[j]), y);
    y = fmaf(w.y, act_tanh(h2[j + 1]), y);
    y = fmaf(w.z, act_tanh(h2[j + 2]), y);
    y = fmaf(w.w, act_tanh(h2[j + 3]), y);
  }
  return y;
}

}  // namespace b2
