// mlp.cuh -- thread-per-row MLP forward with weights staged in shared memory.
//
// Hidden widths 64/32 are far too small for tcgen05 tiles to pay off (K = 6/64/32, N = 64/32/1),
// and the parity gate (1e-6 on mu/V) rules out TF32/BF16 inputs, so inference runs as FP32 FMA:
// every lane owns one observation row, weights sit in shared memory in k-major (transposed) order
// so that one broadcast LDS.128 feeds four FFMAs, activations stay in registers.
// Reference: ActorCritic actor/critic Sequential, include/ppo_nn.h:83-123.
#pragma once
#include "common.cuh"

#ifndef B2_MLP_FFMA2
#define B2_MLP_FFMA2 1
#endif

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

// tanh(x) = 1 - 2 / (1 + 2^(2 x log2 e)): FMUL + MUFU.EX2 + FADD + MUFU.RCP + FFMA.  Absolute error
// <= ~2e-7 over the whole range (saturates cleanly to +-1), which keeps mu / V within the 1e-6 parity gate;
// libdevice tanhf costs ~4x the instructions and its two code paths bloat the unrolled layers.
__device__ __forceinline__ float act_tanh(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.8853900817779268f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
  return fmaf(-2.0f, r, 1.0f);
}

// One observation row through one net, thread-per-row.  Compact on purpose: the fully unrolled version
// (4.5k instructions per call site) made instruction fetch the top stall of the rollout kernel (ncu r1b:
// 61 % of the MLP phase's samples were stall_no_inst).  Hidden layer 1 goes through a per-thread column
// of shared memory (hs[k * blockDim.x + tid]) so that layer 2 can be a rolled loop over k with its 32
// accumulators in registers; every loop body is a few hundred bytes of SASS and stays in the L0 I-cache.
// Returns the pre-final-activation output (caller applies tanh for the actor).
template <class N>
__device__ __forceinline__ float mlp_row(const float* __restrict__ sm, float* __restrict__ hs,
                                         const float (&x)[N::S]) {
  using L = NetSmem<N>;
  const int BS = blockDim.x;
  float* hcol = hs + threadIdx.x;
  // ---- layer 1: 8 outputs per iteration ----
#pragma unroll 2
  for (int j = 0; j < N::H1; j += 8) {
    float4 a0 = *reinterpret_cast<const float4*>(sm + L::ob1 + j);
    float4 a1 = *reinterpret_cast<const float4*>(sm + L::ob1 + j + 4);
#pragma unroll
    for (int s = 0; s < N::S; s++) {
      const float4 w0 = *reinterpret_cast<const float4*>(sm + L::oW1T + s * N::H1 + j);
      const float4 w1 = *reinterpret_cast<const float4*>(sm + L::oW1T + s * N::H1 + j + 4);
      a0.x = fmaf(w0.x, x[s], a0.x); a0.y = fmaf(w0.y, x[s], a0.y); a0.z = fmaf(w0.z, x[s], a0.z); a0.w = fmaf(w0.w, x[s], a0.w);
      a1.x = fmaf(w1.x, x[s], a1.x); a1.y = fmaf(w1.y, x[s], a1.y); a1.z = fmaf(w1.z, x[s], a1.z); a1.w = fmaf(w1.w, x[s], a1.w);
    }
    hcol[(j + 0) * BS] = act_tanh(a0.x); hcol[(j + 1) * BS] = act_tanh(a0.y);
    hcol[(j + 2) * BS] = act_tanh(a0.z); hcol[(j + 3) * BS] = act_tanh(a0.w);
    hcol[(j + 4) * BS] = act_tanh(a1.x); hcol[(j + 5) * BS] = act_tanh(a1.y);
    hcol[(j + 6) * BS] = act_tanh(a1.z); hcol[(j + 7) * BS] = act_tanh(a1.w);
  }
  // ---- layer 2: accumulators in registers, rolled over k ----
  float h2[N::H2];
#if B2_MLP_FFMA2
  // packed FP32 FMA (FFMA2, sm_100): the two halves of an LDS.128 are register pairs already, so a k step costs
  // H2/2 issue slots instead of H2.  Same IEEE fma per element as the scalar form: results are bit-identical.
  float2 acc[N::H2 / 2];
#pragma unroll
  for (int j = 0; j < N::H2; j += 4) {
    const float4 b = *reinterpret_cast<const float4*>(sm + L::ob2 + j);
    acc[j / 2] = make_float2(b.x, b.y); acc[j / 2 + 1] = make_float2(b.z, b.w);
  }
#pragma unroll 4
  for (int k = 0; k < N::H1; k++) {
    const float hk = hcol[k * BS];
    const float2 hk2 = make_float2(hk, hk);
#pragma unroll
    for (int j = 0; j < N::H2; j += 4) {
      const float4 w = *reinterpret_cast<const float4*>(sm + L::oW2T + k * N::H2 + j);
      acc[j / 2] = __ffma2_rn(make_float2(w.x, w.y), hk2, acc[j / 2]);
      acc[j / 2 + 1] = __ffma2_rn(make_float2(w.z, w.w), hk2, acc[j / 2 + 1]);
    }
  }
#pragma unroll
  for (int j = 0; j < N::H2; j += 2) { h2[j] = acc[j / 2].x; h2[j + 1] = acc[j / 2].y; }
#else
#pragma unroll
  for (int j = 0; j < N::H2; j += 4) {
    const float4 b = *reinterpret_cast<const float4*>(sm + L::ob2 + j);
    h2[j] = b.x; h2[j + 1] = b.y; h2[j + 2] = b.z; h2[j + 3] = b.w;
  }
#pragma unroll 4
  for (int k = 0; k < N::H1; k++) {
    const float hk = hcol[k * BS];
#pragma unroll
    for (int j = 0; j < N::H2; j += 4) {
      const float4 w = *reinterpret_cast<const float4*>(sm + L::oW2T + k * N::H2 + j);
      h2[j] = fmaf(w.x, hk, h2[j]);
      h2[j + 1] = fmaf(w.y, hk, h2[j + 1]);
      h2[j + 2] = fmaf(w.z, hk, h2[j + 2]);
      h2[j + 3] = fmaf(w.w, hk, h2[j + 3]);
    }
  }
#endif
  // ---- layer 3 ----
  float y = sm[L::ob3];
#pragma unroll
  for (int j = 0; j < N::H2; j += 4) {
    const float4 w = *reinterpret_cast<const float4*>(sm + L::oW3 + j);
    y = fmaf(w.x, act_tanh(h2[j]), y);
    y = fmaf(w.y, act_tanh(h2[j + 1]), y);
    y = fmaf(w.z, act_tanh(h2[j + 2]), y);
    y = fmaf(w.w, act_tanh(h2[j + 3]), y);
  }
  return y;
}

// dynamic shared memory a thread-per-row kernel needs: one net image + the hidden-1 scratch columns
template <class N>
constexpr int mlp_smem_bytes(int block) { return NetSmem<N>::BYTES + N::H1 * block * 4; }

}  // namespace b2
