// ppo_update.cu -- K3/K4/K5: one PPO optimizer step = evaluate + clipped-surrogate/value loss +
// backward (k_ppo_grad), deterministic cross-CTA reduction (k_reduce) and Adam (fused into k_reduce
// on one GPU, k_adam after the NCCL all-reduce on several).
//
// Replaces ActorCritic::evaluate (src/ppo_nn.cpp:126-194), the loss (:263-290), autograd backward
// (:301-302) and torch::optim::Adam::step (:303; options include/ppo_nn.h:166-170).
//
// k_ppo_grad design.  The two 6-64-32-1 MLPs are far below tensor-core tile sizes and the 1e-4
// gradient gate rules out TF32/BF16 operands, so the pass is FP32 FMA, organised as register-tiled
// small GEMMs over shared memory.  A CTA (256 threads) walks tiles of TS = 128 samples; all
// activations of a tile live in shared memory as [feature][sample] rows (pitch TS+4):
//     X(6+1) -> L1 -> H1(64+1) -> L2 -> H2(32+1) -> L3 -> loss head -> dz3
//     D2 = dz3 (x) W3 .* (1-H2^2);   dW2 += D2 . H1^T;   D1 = (W2^T D2) .* (1-H1^2) (in place on H1);
//     dW1 += D1 . X^T;  biases ride along as a row of ones.
// Every thread owns a 4x4 (feature x sample) register tile; lanes of a warp form an 8 (feature) x 4
// (sample) grid so that both operands of the k-loop are one shared-memory wavefront each
// (2 wavefronts per 16 FFMA per warp).  Feature tiles are strided (f = ft + 8 i) so tile stores are
// bank-conflict free; the staged weights are column-permuted to keep the strided tile contiguous.
// Weight-gradient accumulators stay in registers across ALL tiles of a CTA and are written once, to a
// per-CTA row of `partials`; k_reduce then sums rows in a fixed order (bit-reproducible, no atomics).
// The critic pass runs first and leaves V in `vcur` for the actor's advantage A = R - V (:274, Q7).
//
// Roofline: FP32 FMA issue; 28,032 flop per sample-epoch, 36 B read per sample-epoch.
#include <cooperative_groups.h>

#include "engine.h"
#include "mlp.cuh"

namespace b2 {

namespace {

using N = NetDefault;
constexpr int TS = 128;            // samples per tile
constexpr int NT = 256;            // threads per CTA
constexpr int PS = TS + 4;         // row pitch (floats): 4 r mod 32 bank offset per row
constexpr int H1 = N::H1, H2 = N::H2, S = N::S;
static_assert(H1 == 64 && H2 == 32 && S == 6, "k_ppo_grad is specialised for the reference MLP");

// shared-memory map (floats)
constexpr int XR = 8;              // rows 0..5 obs, 6 ones, 7 zero
constexpr int H1R = H1 + 4;        // row H1 ones
constexpr int H2R = H2 + 4;        // row H2 ones
constexpr int oX = 0;
constexpr int oH1 = oX + XR * PS;
constexpr int oH2 = oH1 + H1R * PS;
constexpr int oD2 = oH2 + H2R * PS;
constexpr int oVec = oD2 + H2 * PS;          // act, oldlp, R, V, dz3 : 5 x TS
constexpr int oW1T = oVec + 5 * TS;          // [8][H1]   rows 0..5 W1^T (permuted cols), 6 = b1, 7 = 0
constexpr int oW2T = oW1T + 8 * H1;          // [H1+1][H2] rows k<H1 W2^T (permuted cols), row H1 = b2
constexpr int oW2N = oW2T + (H1 + 1) * H2;   // [H2][H1]  W2 natural rows, permuted cols
constexpr int oW3 = oW2N + H2 * H1;          // [H2+4]    W3, b3, 0..
constexpr int oRed = oW3 + H2 + 4;           // 32 floats reduction scratch
constexpr int oX1 = oRed + 32;                // second X tile buffer (prefetch target)
constexpr int oVec1 = oX1 + XR * PS;          // second act/oldlp/R/V buffer (4 x TS)
constexpr int SMEM_FLOATS = oVec1 + 4 * TS;
constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
static_assert(2 * SMEM_BYTES <= 227 * 1024, "two CTAs per SM");
// end-of-pass scratch (aliases the activation rows, which are dead by then)
constexpr int oScrW2 = 0;                    // [2][H2*H1]
constexpr int oScrW1 = oScrW2 + 2 * H2 * H1; // [16][H1*8]
static_assert(oScrW1 + 16 * H1 * 8 <= oVec, "scratch fits in the activation area");

constexpr int PSTRIDE = ((N::P + 4 + 31) / 32) * 32;  // partial row: P grads + loss1, sum e^2, clipped, pad

__device__ __forceinline__ int perm32(int f) { return (f & 7) * 4 + (f >> 3); }           // f < 32
__device__ __forceinline__ int perm64(int f) { return (f & 32) + perm32(f & 31); }        // f < 64

__device__ __forceinline__ void stage_weights(float* sm, const float* __restrict__ g) {
  for (int i = threadIdx.x; i < 8 * H1; i += NT) {
    int r = i / H1, f = i % H1;
    float v = (r < S) ? g[N::oW1 + f * S + r] : (r == S ? g[N::ob1 + f] : 0.0f);
    sm[oW1T + r * H1 + perm64(f)] = v;
  }
  for (int i = threadIdx.x; i < (H1 + 1) * H2; i += NT) {
    int k = i / H2, j = i % H2;
    float v = (k < H1) ? g[N::oW2 + j * H1 + k] : g[N::ob2 + j];
    sm[oW2T + k * H2 + perm32(j)] = v;
  }
  for (int i = threadIdx.x; i < H2 * H1; i += NT) {
    int j = i / H1, k = i % H1;
    sm[oW2N + j * H1 + perm64(k)] = g[N::oW2 + i];
  }
  for (int i = threadIdx.x; i < H2 + 4; i += NT)
    sm[oW3 + i] = (i < H2) ? g[N::oW3 + i] : (i == H2 ? g[N::ob3] : 0.0f);
}

#ifndef B2_GRAD_FFMA2
#define B2_GRAD_FFMA2 1
#endif
#if B2_GRAD_FFMA2
// 4 x 4 rank-1 update with packed FP32 FMA (FFMA2, sm_100): 8 issue slots instead of 16; the column pairs of `a` are the
// halves of an LDS.128 and the accumulator rows live in aligned register pairs.  Same IEEE fma per element.
__device__ __forceinline__ void fma44(float (&acc)[4][4], const float4& b, const float4& a) {
  const float2 a01 = make_float2(a.x, a.y), a23 = make_float2(a.z, a.w);
  const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float2 bb = make_float2(bv[i], bv[i]);
    const float2 r0 = __ffma2_rn(bb, a01, make_float2(acc[i][0], acc[i][1]));
    const float2 r1 = __ffma2_rn(bb, a23, make_float2(acc[i][2], acc[i][3]));
    acc[i][0] = r0.x; acc[i][1] = r0.y; acc[i][2] = r1.x; acc[i][3] = r1.y;
  }
}
#else
__device__ __forceinline__ void fma44(float (&acc)[4][4], const float4& b, const float4& a) {
  acc[0][0] = fmaf(b.x, a.x, acc[0][0]); acc[0][1] = fmaf(b.x, a.y, acc[0][1]);
  acc[0][2] = fmaf(b.x, a.z, acc[0][2]); acc[0][3] = fmaf(b.x, a.w, acc[0][3]);
  acc[1][0] = fmaf(b.y, a.x, acc[1][0]); acc[1][1] = fmaf(b.y, a.y, acc[1][1]);
  acc[1][2] = fmaf(b.y, a.z, acc[1][2]); acc[1][3] = fmaf(b.y, a.w, acc[1][3]);
  acc[2][0] = fmaf(b.z, a.x, acc[2][0]); acc[2][1] = fmaf(b.z, a.y, acc[2][1]);
  acc[2][2] = fmaf(b.z, a.z, acc[2][2]); acc[2][3] = fmaf(b.z, a.w, acc[2][3]);
  acc[3][0] = fmaf(b.w, a.x, acc[3][0]); acc[3][1] = fmaf(b.w, a.y, acc[3][1]);
  acc[3][2] = fmaf(b.w, a.z, acc[3][2]); acc[3][3] = fmaf(b.w, a.w, acc[3][3]);
}
#endif
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void fma_dot4(float& acc, const float4& a, const float4& b) {  // 4 FFMA, no FADD
  acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); acc = fmaf(a.w, b.w, acc);
}
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ float sum4(const float4& a) { return (a.x + a.y) + (a.z + a.w); }
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace

__global__ void __launch_bounds__(NT, 2) k_ppo_grad(const UpdateArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ft = lane & 7;                 // feature-tile lane coordinate
  const int s0 = 4 * (4 * warp + (lane >> 3));  // first sample of this thread's 4-sample chunk
  const int64_t ntiles = (a.count + TS - 1) / TS;
  const float inv_denom = 1.0f / (float)a.denom;
  float* part = a.partials + (int64_t)blockIdx.x * PSTRIDE;
  const int64_t cap = a.tr.cap;

  if (blockIdx.x == 0 && tid == 0) *a.adam_step += 1;  // one launch == one optimizer step

  // dW2 tile ownership: 8 j-tiles x 16 k-tiles x 2 sample halves
  const int w2_half = warp >> 2;
  const int w2_kt = lane & 15;
  const int w2_jt = (warp & 3) * 2 + (lane >> 4);
  // dW1 tile ownership: 16 k-tiles x 16 sample slices of 8
  const int w1_kt = tid & 15;
  const int w1_sl = tid >> 4;

  float loss1_sum = 0.f, loss2_sum = 0.f, clip_sum = 0.f;  // accumulated by lanes < 16 of every warp

#pragma unroll 1
  for (int net = 1; net >= 0; net--) {   // critic first: the actor's advantage needs the current V (:274)
    __syncthreads();
    stage_weights(sm, a.params + net * N::SIZE);
    float gW2[4][4], gW1[4][8], gRow[9];
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
      for (int j = 0; j < 4; j++) gW2[i][j] = 0.f;
#pragma unroll
      for (int j = 0; j < 8; j++) gW1[i][j] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < 9; i++) gRow[i] = 0.f;

    // constant rows (bias columns) -- rewritten every pass: the end-of-pass scratch aliases them
    for (int i = tid; i < 2 * TS; i += NT) {
      const float v = (i < TS) ? 1.0f : 0.0f;           // row S: ones, row S+1: zeros
      sm[oX + S * PS + (i % TS) + (i / TS) * PS] = v;
      sm[oX1 + S * PS + (i % TS) + (i / TS) * PS] = v;
    }
    if (tid < TS) {
      sm[oH1 + H1 * PS + tid] = 1.0f;
      sm[oH2 + H2 * PS + tid] = 1.0f;
    }
    // Asynchronous tile fetch (LDGSTS): tile i+1 streams into the other buffer while tile i is being computed.
    auto prefetch = [&](int64_t tile, int b) {
      float* xb = sm + (b ? oX1 : oX);
      float* vb = sm + (b ? oVec1 : oVec);
      const int64_t p0 = a.first + tile * TS;
      for (int i = tid; i < (S + 4) * TS; i += NT) {
        const int r = i / TS, s = i % TS;
        const bool valid = (tile * TS + s) < a.count;
        float* dst = (r < S) ? (xb + r * PS + s) : (vb + (r - S) * TS + s);
        const float* src = nullptr;
        if (valid) {
          const int64_t k = a.idx ? (int64_t)a.idx[p0 + s] : (p0 + s);
          if (r < S) src = a.tr.obs + r * cap + k;
          else if (r == S) src = a.tr.action + k;
          else if (r == S + 1) src = a.tr.logp + k;
          else if (r == S + 2) src = a.tr.ret + k;
          else if (net == 0) src = a.tr.vcur + k;
        }
        if (src) cp_async4(dst, src);
        else *dst = 0.f;
      }
      cp_async_commit();
    };
    int cur = 0;
    if ((int64_t)blockIdx.x < ntiles) prefetch(blockIdx.x, 0);

#pragma unroll 1
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, cur ^= 1) {
      cp_async_wait_all();
      __syncthreads();  // this tile has landed; the previous tile's readers are done (and weights are staged)
      if (tile + gridDim.x < ntiles) prefetch(tile + gridDim.x, cur ^ 1);
      const float* xb = sm + (cur ? oX1 : oX);
      const float* vb = sm + (cur ? oVec1 : oVec);
      const int64_t p0 = a.first + tile * TS;

      // ---- L1: H1 = tanh(W1 x + b1), two strided feature groups per thread ----
      {
        float acc0[4][4] = {}, acc1[4][4] = {};
#pragma unroll
        for (int k = 0; k < S + 1; k++) {
          const float4 av = ld4(xb + k * PS + s0);
          fma44(acc0, ld4(sm + oW1T + k * H1 + 4 * ft), av);
          fma44(acc1, ld4(sm + oW1T + k * H1 + 32 + 4 * ft), av);
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
          st4(sm + oH1 + (ft + 8 * i) * PS + s0, act_tanh(acc0[i][0]), act_tanh(acc0[i][1]), act_tanh(acc0[i][2]),
              act_tanh(acc0[i][3]));
          st4(sm + oH1 + (32 + ft + 8 * i) * PS + s0, act_tanh(acc1[i][0]), act_tanh(acc1[i][1]),
              act_tanh(acc1[i][2]), act_tanh(acc1[i][3]));
        }
      }
      __syncwarp();  // H1 columns of this warp's 16 samples are produced and consumed by the same warp
      // ---- L2: H2 = tanh(W2 h1 + b2) ----
      {
        float acc[4][4] = {};
#pragma unroll 5
        for (int k = 0; k < H1 + 1; k++) fma44(acc, ld4(sm + oW2T + k * H2 + 4 * ft), ld4(sm + oH1 + k * PS + s0));
#pragma unroll
        for (int i = 0; i < 4; i++)
          st4(sm + oH2 + (ft + 8 * i) * PS + s0, act_tanh(acc[i][0]), act_tanh(acc[i][1]), act_tanh(acc[i][2]),
              act_tanh(acc[i][3]));
      }
      __syncwarp();
      // ---- L3 + loss head, warp-local: lane l and lane l+16 split the 33-term dot product of sample 16 warp + (l & 15)
      {
        const int sl = 16 * warp + (lane & 15);
        const int jb = (lane >> 4) * 16;
        float y0 = 0.f, y1 = 0.f;
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          y0 = fmaf(sm[oW3 + jb + j], sm[oH2 + (jb + j) * PS + sl], y0);
          y1 = fmaf(sm[oW3 + jb + j + 1], sm[oH2 + (jb + j + 1) * PS + sl], y1);
        }
        float y = y0 + y1;
        y += __shfl_xor_sync(0xffffffffu, y, 16);
        y += sm[oW3 + H2];                                     // b3 (the ones row of H2 is not needed here)
        const bool valid = (lane < 16) && (tile * TS + sl) < a.count;
        const float R = vb[2 * TS + sl];
        float dz3 = 0.f;
        if (net == 1) {
          const float V = y;                                   // critic: no final activation (ppo_nn.h:118-122)
          const float e = V - R;
          if (valid) {
            loss2_sum = fmaf(e, e, loss2_sum);                 // :288  0.5 * MSE
            dz3 = e * inv_denom;
            const int64_t p = p0 + sl;
            a.tr.vcur[a.idx ? (int64_t)a.idx[p] : p] = V;
          }
        } else if (valid) {
          const float mu = act_tanh(y);                        // ppo_nn.h:102
          const float act = vb[sl];
          const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);  // :140
          const float ratio = expf(lp - vb[TS + sl]);          // :263
          const float A = R - vb[3 * TS + sl];                 // :274 (detached)
          const float surr1 = ratio * A;                       // :284
          const float rc = fminf(fmaxf(ratio, a.hp.clip_lo), a.hp.clip_hi);
          const float surr2 = rc * A;                          // :285
          loss1_sum -= fminf(surr1, surr2);                    // :287
          const bool inside = (ratio >= a.hp.clip_lo) && (ratio <= a.hp.clip_hi);
          // autograd of min(): the smaller branch takes the gradient, ties split it; clamp passes it inside only
          float gr;
          if (surr1 < surr2) gr = A;
          else if (surr1 > surr2) gr = inside ? A : 0.f;
          else gr = 0.5f * A + (inside ? 0.5f * A : 0.f);
          clip_sum += inside ? 0.f : 1.f;
          const float dlogp = -gr * ratio * inv_denom;         // d(mean(-min))/dlogp
          const float dmu = dlogp * (act - mu) * a.hp.inv_var; // :20-31
          dz3 = dmu * (1.0f - mu * mu);
        }
        if (lane < 16) sm[oVec + 4 * TS + sl] = dz3;
      }
      __syncwarp();
      // ---- D2 = dz3 (x) W3 .* (1 - H2^2) ----
      {
        const float4 dz = ld4(sm + oVec + 4 * TS + s0);
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int j = ft + 8 * i;
          const float w3 = sm[oW3 + j];
          const float4 h = ld4(sm + oH2 + j * PS + s0);
          st4(sm + oD2 + j * PS + s0, dz.x * w3 * (1.f - h.x * h.x), dz.y * w3 * (1.f - h.y * h.y),
              dz.z * w3 * (1.f - h.z * h.z), dz.w * w3 * (1.f - h.w * h.w));
        }
      }
      __syncthreads();
      // ---- row sums: db2[j] = sum_s D2[j][s]; dW3[j] = sum_s dz3[s] H2[j][s]; db3 = sum_s dz3[s] ----
      {
        const float4 dz = ld4(sm + oVec + 4 * TS + 4 * lane);
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int j = warp + 8 * i;
          gRow[i] += sum4(ld4(sm + oD2 + j * PS + 4 * lane));
          fma_dot4(gRow[4 + i], dz, ld4(sm + oH2 + j * PS + 4 * lane));
        }
        if (warp == 0) gRow[8] += sum4(dz);
      }
      // ---- dW2[j][k] += sum_s D2[j][s] H1[k][s]  (4x4 register tile, half of the samples) ----
      {
        const int sb = w2_half * (TS / 2);
#pragma unroll 2
        for (int s = 0; s < TS / 2; s += 4) {
          float4 d[4], h[4];
#pragma unroll
          for (int i = 0; i < 4; i++) d[i] = ld4(sm + oD2 + (w2_jt + 8 * i) * PS + sb + s);
#pragma unroll
          for (int i = 0; i < 4; i++) h[i] = ld4(sm + oH1 + (w2_kt + 16 * i) * PS + sb + s);
#pragma unroll
          for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) fma_dot4(gW2[i][j], d[i], h[j]);
        }
      }
      __syncthreads();  // H1 is about to be overwritten by D1
      // ---- D1 = (W2^T D2) .* (1 - H1^2), in place ----
      {
        float acc0[4][4] = {}, acc1[4][4] = {};
#pragma unroll 4
        for (int j = 0; j < H2; j++) {
          const float4 dv = ld4(sm + oD2 + j * PS + s0);
          fma44(acc0, ld4(sm + oW2N + j * H1 + 4 * ft), dv);
          fma44(acc1, ld4(sm + oW2N + j * H1 + 32 + 4 * ft), dv);
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
          float* r0 = sm + oH1 + (ft + 8 * i) * PS + s0;
          float* r1 = sm + oH1 + (32 + ft + 8 * i) * PS + s0;
          const float4 h0 = ld4(r0), h1v = ld4(r1);
          st4(r0, acc0[i][0] * (1.f - h0.x * h0.x), acc0[i][1] * (1.f - h0.y * h0.y),
              acc0[i][2] * (1.f - h0.z * h0.z), acc0[i][3] * (1.f - h0.w * h0.w));
          st4(r1, acc1[i][0] * (1.f - h1v.x * h1v.x), acc1[i][1] * (1.f - h1v.y * h1v.y),
              acc1[i][2] * (1.f - h1v.z * h1v.z), acc1[i][3] * (1.f - h1v.w * h1v.w));
        }
      }
      __syncthreads();
      // ---- dW1aug[k][c] += sum_s D1[k][s] X[c][s]   (c = 6 is the ones row -> db1) ----
      {
        const int sb = w1_sl * 8;
#pragma unroll
        for (int s = 0; s < 8; s += 4) {
          float4 xv[S + 1];
#pragma unroll
          for (int c = 0; c < S + 1; c++) xv[c] = ld4(xb + c * PS + sb + s);
#pragma unroll
          for (int i = 0; i < 4; i++) {
            const float4 d = ld4(sm + oH1 + (w1_kt + 16 * i) * PS + sb + s);
#pragma unroll
            for (int c = 0; c < S + 1; c++) fma_dot4(gW1[i][c], d, xv[c]);
          }
        }
      }
    }  // tiles

    // ---- end of pass: fold the per-thread accumulators and write this CTA's partial row ----
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++)
        sm[oScrW2 + w2_half * (H2 * H1) + (w2_jt + 8 * i) * H1 + (w2_kt + 16 * j)] = gW2[i][j];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int c = 0; c < 8; c++) sm[oScrW1 + w1_sl * (H1 * 8) + (w1_kt + 16 * i) * 8 + c] = gW1[i][c];
    float* gp = part + net * N::SIZE;
#pragma unroll
    for (int i = 0; i < 9; i++) gRow[i] = warp_sum(gRow[i]);
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        gp[N::ob2 + warp + 8 * i] = gRow[i];
        gp[N::oW3 + warp + 8 * i] = gRow[4 + i];
      }
      if (warp == 0) gp[N::ob3] = gRow[8];
    }
    __syncthreads();
    for (int i = tid; i < H2 * H1; i += NT) gp[N::oW2 + i] = sm[oScrW2 + i] + sm[oScrW2 + H2 * H1 + i];
    for (int i = tid; i < H1 * 8; i += NT) {
      const int k = i >> 3, c = i & 7;
      float v = 0.f;
#pragma unroll
      for (int sl = 0; sl < 16; sl++) v += sm[oScrW1 + sl * (H1 * 8) + i];
      if (c < S) gp[N::oW1 + k * S + c] = v;
      else if (c == S) gp[N::ob1 + k] = v;
    }
  }  // nets

  // ---- loss partials of this CTA ----
  __syncthreads();
  loss1_sum = warp_sum(loss1_sum); loss2_sum = warp_sum(loss2_sum); clip_sum = warp_sum(clip_sum);
  if (lane == 0) { sm[oRed + warp] = loss1_sum; sm[oRed + 8 + warp] = loss2_sum; sm[oRed + 16 + warp] = clip_sum; }
  __syncthreads();
  if (tid < 3) {
    float v = 0.f;
    for (int w = 0; w < NT / 32; w++) v += sm[oRed + 8 * tid + w];
    part[N::P + tid] = v;
  }
}

// K4 (+K5): fixed-order sum of the per-CTA rows; on one GPU Adam is applied in the same pass.
// Adam follows libtorch's adam.cpp: scalars are doubles narrowed to f32 inside the tensor kernels.
__device__ __forceinline__ void adam_apply(const UpdateArgs& a, int p, float g, int64_t step) {
  const double bc1 = 1.0 - pow(a.hp.beta1, (double)step);
  const double bc2 = 1.0 - pow(a.hp.beta2, (double)step);
  const float fb1 = (float)a.hp.beta1, f1mb1 = (float)(1.0 - a.hp.beta1);
  const float fb2 = (float)a.hp.beta2, f1mb2 = (float)(1.0 - a.hp.beta2);
  const float sq_bc2 = (float)sqrt(bc2), feps = (float)a.hp.eps, step_size = (float)(-(a.hp.lr / bc1));
  float m = a.adam_m[p], v = a.adam_v[p];
  m = __fadd_rn(__fmul_rn(m, fb1), __fmul_rn(f1mb1, g));                 // exp_avg.mul_(b1).add_(grad, 1-b1)
  v = __fadd_rn(__fmul_rn(v, fb2), __fmul_rn(f1mb2, __fmul_rn(g, g)));   // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), sq_bc2), feps); // (sqrt(v)/sqrt(bc2)).add_(eps)
  a.adam_m[p] = m;
  a.adam_v[p] = v;
  a.params[p] = __fadd_rn(a.params[p], __fmul_rn(step_size, __fdiv_rn(m, denom)));  // p.addcdiv_(m, denom, -step_size)
}

__global__ void __launch_bounds__(128) k_reduce(const UpdateArgs a, int apply_adam) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= N::P + 3) return;
  const bool peers = a.px.world > 1;   // several GPUs: exchange over peer memory inside this kernel (engine.h PeerXchg)
  const int slot = p < N::P ? p : PSTRIDE - 4 + (p - N::P);   // loss1 sum, sum e^2, clipped count ride in the tail
  float g = 0.f;
  for (int c = 0; c < a.grid; c++) g += a.partials[(int64_t)c * PSTRIDE + p];
  a.grads[slot] = g;
  bool ok = true;
  if (peers) g = px_allreduce_elem(a.px, slot, g, &ok);
  if (!ok || !(apply_adam || peers)) return;
  if (p < N::P) adam_apply(a, p, g, *a.adam_step);
  else {
    const int q = p - N::P;
    a.loss_log[a.step_slot * 4 + q] = (q == 1 ? 0.5 : 1.0) * (double)g / (double)a.denom;
  }
}

// Adam after an ncclAllReduce of a.grads (the default multi-GPU path)
__global__ void __launch_bounds__(128) k_adam(const UpdateArgs a) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < N::P) adam_apply(a, p, a.grads[p], *a.adam_step);
  else if (p < N::P + 3) {
    const int q = p - N::P;
    a.loss_log[a.step_slot * 4 + q] = (q == 1 ? 0.5 : 1.0) * (double)a.grads[PSTRIDE - 4 + q] / (double)a.denom;
  }
}

int update_partial_stride() { return PSTRIDE; }

int update_grid_for(int64_t count, int sm_count) {
  int64_t tiles = (count + TS - 1) / TS;
  int64_t g = 2LL * sm_count;  // two resident CTAs per SM
  if (tiles < g) g = tiles;
  if (g < 1) g = 1;
  return (int)g;
}

cudaError_t launch_ppo_grad(const UpdateArgs& a, cudaStream_t st) {
  static b2::PerDeviceOnce once;   // function attributes are per device
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_ppo_grad, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    once.done();
  }
  k_ppo_grad<<<a.grid, NT, SMEM_BYTES, st>>>(a);
  return cudaGetLastError();
}
cudaError_t launch_reduce(const UpdateArgs& a, bool apply_adam, cudaStream_t st) {
  k_reduce<<<(N::P + 3 + 127) / 128, 128, 0, st>>>(a, apply_adam ? 1 : 0);
  return cudaGetLastError();
}
cudaError_t launch_adam(const UpdateArgs& a, cudaStream_t st) {
  k_adam<<<(N::P + 3 + 127) / 128, 128, 0, st>>>(a);
  return cudaGetLastError();
}

}  // namespace b2
