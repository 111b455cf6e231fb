// wide_update.cu -- PPO optimizer step for wide hidden layers (e.g. 2x256, BASELINE configs[4]): the FP32-FMA side
// kernels around the three tcgen05 GEMMs of wide_gemm.cu, the deterministic gradient reduction + Adam, and the
// host sequence that strings them together.  See wide.h for the dataflow.  Replaces ActorCritic::evaluate
// (src/ppo_nn.cpp:126-194), the loss (:263-290), autograd backward (:301-302) and Adam::step (:303).
//
// Memory plan for a chunk of CAP samples (CAP % 128 == 0), g = net (0 actor, 1 critic), pl = bf16 plane (hi, mid):
//   w2p  [g*2+pl][H2][H1]   w2tp [g*2+pl][H1][H2]   bias2 [g][H2]                    (re-split every step)
//   h1p  [g*2+pl][CAP][H1]  h1tp [g*2+pl][H1][CAP]                                   (layer-1 activations, both majors)
//   h2   [g][CAP][H2] f32                                                            (layer-2 activations)
//   d2p  [g*2+pl][CAP][H2]  d2tp [g*2+pl][H2][CAP]                                   (dL/dz2, both majors)
//   d1   [g][CAP][H1] f32                                                            (dL/dz1)
// Both majors exist so that every GEMM reads K-major operands through plain 128-B-swizzled TMA boxes.
// Samples in [count, round_up(count, 128)) are written as zeros by the producers, so the sample-contraction GEMM
// (dW2) needs no masking.
#include <algorithm>
#include <new>

#include "tc_ptx.cuh"
#include "wide.h"
#include "mlp.cuh"

namespace b2 {

namespace {

constexpr int S = OBS;
constexpr int TILE_S = 32;   // samples per tile of the FP32 side kernels

struct WideArgs {
  NetDims d;
  const float* params;
  Traj tr;
  const int32_t* idx;
  int64_t first;      // position of sample 0 of this chunk
  int count, m_pad;   // valid samples, padded to 128
  int64_t cap;        // CAP
  float inv_denom;
  HyperDev hp;
  __nv_bfloat16 *w2p, *w2tp, *h1p, *h1tp, *d2p, *d2tp;
  float *bias2, *h2, *d1;
  float *part_head, *part_dw1;   // [blocks][2][head_row] , [blocks][2][H1 * 8]
  int head_row;
};

__device__ __forceinline__ int64_t sample_pos(const WideArgs& a, int s) {
  const int64_t p = a.first + s;
  return a.idx ? (int64_t)a.idx[p] : p;
}

// weights -> bf16 planes in both majors (runs once per optimizer step: the weights just changed)
__global__ void __launch_bounds__(256) k_wide_prep(const WideArgs a, int64_t* adam_step) {
  const int H1 = a.d.H1, H2 = a.d.H2;
  if (blockIdx.x == 0 && threadIdx.x == 0) *adam_step += 1;   // one prep launch == one optimizer step
  const int n = 2 * H2 * H1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int g = i / (H2 * H1), r = i % (H2 * H1), j = r / H1, k = r % H1;
    const float w = a.params[g * a.d.SIZE + a.d.oW2 + r];
    __nv_bfloat16 hi, mid;
    tc::split_bf16(w, hi, mid);
    const int64_t ps = (int64_t)H2 * H1;
    a.w2p[(g * 2 + 0) * ps + r] = hi;
    a.w2p[(g * 2 + 1) * ps + r] = mid;
    a.w2tp[(g * 2 + 0) * ps + (int64_t)k * H2 + j] = hi;
    a.w2tp[(g * 2 + 1) * ps + (int64_t)k * H2 + j] = mid;
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * H2; i += gridDim.x * blockDim.x)
    a.bias2[i] = a.params[(i / H2) * a.d.SIZE + a.d.ob2 + (i % H2)];
}

// layer 1: H1 = tanh(W1 x + b1) for a tile of 32 samples; thread = feature.  grid (m_pad / 32, 2 nets)
__global__ void __launch_bounds__(256) k_wide_l1(const WideArgs a) {
  __shared__ float xs[TILE_S][S + 2];
  const int H1 = a.d.H1, g = blockIdx.y;
  const int s0 = blockIdx.x * TILE_S;
  for (int i = threadIdx.x; i < TILE_S * S; i += blockDim.x) {
    const int s = i / S, c = i % S;
    xs[s][c] = (s0 + s < a.count) ? a.tr.obs[c * a.tr.cap + sample_pos(a, s0 + s)] : 0.0f;
  }
  __syncthreads();
  const float* net = a.params + g * a.d.SIZE;
  const int64_t plane = a.cap * H1;
  for (int k = threadIdx.x; k < H1; k += blockDim.x) {
    float w[S];
#pragma unroll
    for (int c = 0; c < S; c++) w[c] = net[a.d.oW1 + k * S + c];
    const float b = net[a.d.ob1 + k];
    __align__(16) __nv_bfloat16 th[TILE_S], tm[TILE_S];
#pragma unroll
    for (int s = 0; s < TILE_S; s++) {
      float z = b;
#pragma unroll
      for (int c = 0; c < S; c++) z = fmaf(w[c], xs[s][c], z);
      const float h = (s0 + s < a.count) ? act_tanh(z) : 0.0f;
      tc::split_bf16(h, th[s], tm[s]);
      const int64_t o = (int64_t)(s0 + s) * H1 + k;
      a.h1p[(g * 2 + 0) * plane + o] = th[s];
      a.h1p[(g * 2 + 1) * plane + o] = tm[s];
    }
    // feature-major copy: 32 consecutive samples of one feature = 64 B contiguous
    uint4* oh = reinterpret_cast<uint4*>(a.h1tp + (g * 2 + 0) * plane + (int64_t)k * a.cap + s0);
    uint4* om = reinterpret_cast<uint4*>(a.h1tp + (g * 2 + 1) * plane + (int64_t)k * a.cap + s0);
#pragma unroll
    for (int v = 0; v < 4; v++) {
      oh[v] = *reinterpret_cast<const uint4*>(&th[8 * v]);
      om[v] = *reinterpret_cast<const uint4*>(&tm[8 * v]);
    }
  }
}

// layer 3 + loss head + D2 for both nets.  One warp per sample (lanes own j = i*32 + lane), 8 warps walk a tile of
// 32 samples; D2 goes out sample-major directly and feature-major through a shared-memory transpose.  dW3, db2, db3 and
// the loss sums are accumulated per lane over all tiles of the block and written once, to this block's partial row.
template <int NJ>   // H2 / 32
__global__ void __launch_bounds__(256) k_wide_head(const WideArgs a) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  constexpr int H2 = NJ * 32;
  constexpr int ROWW = 17;                                 // words per transposed row: 32 bf16 + 1 pad word
  uint32_t* tileT = reinterpret_cast<uint32_t*>(sm_raw);   // [4 planes][H2][ROWW]
  float* red = reinterpret_cast<float*>(tileT + 4 * H2 * ROWW);  // [8 warps][2 nets][2 * H2 + 4]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const NetDims d = a.d;
  float w3[2][NJ], b3[2];
#pragma unroll
  for (int g = 0; g < 2; g++) {
#pragma unroll
    for (int i = 0; i < NJ; i++) w3[g][i] = a.params[g * d.SIZE + d.oW3 + i * 32 + lane];
    b3[g] = a.params[g * d.SIZE + d.ob3];
  }
  float gW3[2][NJ], gb2[2][NJ], gb3[2] = {0.f, 0.f};
#pragma unroll
  for (int g = 0; g < 2; g++)
#pragma unroll
    for (int i = 0; i < NJ; i++) { gW3[g][i] = 0.f; gb2[g][i] = 0.f; }
  float loss1 = 0.f, loss2 = 0.f, clipped = 0.f;
  const int64_t plane = a.cap * H2;
  const int n_tiles = a.m_pad / TILE_S;
  __nv_bfloat16* tileH = reinterpret_cast<__nv_bfloat16*>(tileT);

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int s0 = tile * TILE_S;
    for (int q = 0; q < TILE_S / 8; q++) {
      const int sl = warp * (TILE_S / 8) + q;
      const int s = s0 + sl;
      const bool valid = s < a.count;
      float h[2][NJ], y[2] = {0.f, 0.f};
#pragma unroll
      for (int g = 0; g < 2; g++)
#pragma unroll
        for (int i = 0; i < NJ; i++) {
          h[g][i] = valid ? a.h2[((int64_t)g * a.cap + s) * H2 + i * 32 + lane] : 0.0f;
          y[g] = fmaf(w3[g][i], h[g][i], y[g]);
        }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        y[0] += __shfl_xor_sync(0xffffffffu, y[0], o);
        y[1] += __shfl_xor_sync(0xffffffffu, y[1], o);
      }
      float dz[2] = {0.f, 0.f};
      if (valid) {
        const int64_t p = sample_pos(a, s);
        const float R = a.tr.ret[p];
        const float V = y[1] + b3[1];                                  // critic: no final activation (ppo_nn.h:118-122)
        const float e = V - R;
        dz[1] = e * a.inv_denom;                                       // d(0.5 mean e^2)/dV  (:288)
        const float mu = act_tanh(y[0] + b3[0]);                       // ppo_nn.h:102
        const float act = a.tr.action[p];
        const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);  // :140
        const float ratio = expf(lp - a.tr.logp[p]);                   // :263
        const float A = R - V;                                         // :274 (detached)
        const float surr1 = ratio * A;                                 // :284
        const float rc = fminf(fmaxf(ratio, a.hp.clip_lo), a.hp.clip_hi);
        const float surr2 = rc * A;                                    // :285
        const bool inside = (ratio >= a.hp.clip_lo) && (ratio <= a.hp.clip_hi);
        float gr;                                                      // autograd of min(): ties split the gradient
        if (surr1 < surr2) gr = A;
        else if (surr1 > surr2) gr = inside ? A : 0.f;
        else gr = 0.5f * A + (inside ? 0.5f * A : 0.f);
        const float dlogp = -gr * ratio * a.inv_denom;
        const float dmu = dlogp * (act - mu) * a.hp.inv_var;           // :20-31
        dz[0] = dmu * (1.0f - mu * mu);
        if (lane == 0) {
          loss2 = fmaf(e, e, loss2);
          loss1 -= fminf(surr1, surr2);                                // :287
          clipped += inside ? 0.f : 1.f;
          gb3[0] += dz[0];
          gb3[1] += dz[1];
        }
      }
#pragma unroll
      for (int g = 0; g < 2; g++)
#pragma unroll
        for (int i = 0; i < NJ; i++) {
          const float dd = dz[g] * w3[g][i] * (1.0f - h[g][i] * h[g][i]);
          gW3[g][i] = fmaf(dz[g], h[g][i], gW3[g][i]);
          gb2[g][i] += dd;
          __nv_bfloat16 hi, mid;
          tc::split_bf16(dd, hi, mid);
          const int j = i * 32 + lane;
          const int64_t o = (int64_t)s * H2 + j;
          a.d2p[(g * 2 + 0) * plane + o] = hi;
          a.d2p[(g * 2 + 1) * plane + o] = mid;
          tileH[(((g * 2 + 0) * H2 + j) * ROWW) * 2 + sl] = hi;
          tileH[(((g * 2 + 1) * H2 + j) * ROWW) * 2 + sl] = mid;
        }
    }
    __syncthreads();
    // feature-major rows: 16 words (32 samples) per (plane, j); a half-warp writes one 64-B row
    for (int r = threadIdx.x >> 4; r < 4 * H2; r += 16) {
      const int wl = threadIdx.x & 15;
      uint32_t* dst = reinterpret_cast<uint32_t*>(a.d2tp + (int64_t)(r / H2) * plane + (int64_t)(r % H2) * a.cap + s0);
      dst[wl] = tileT[r * ROWW + wl];
    }
    __syncthreads();
  }

  // block reduction of the per-lane column sums -> this block's partial row (fixed order: deterministic)
  const int RW = 2 * H2 + 4;
#pragma unroll
  for (int g = 0; g < 2; g++) {
    float* r = red + (warp * 2 + g) * RW;
#pragma unroll
    for (int i = 0; i < NJ; i++) { r[i * 32 + lane] = gW3[g][i]; r[H2 + i * 32 + lane] = gb2[g][i]; }
    if (lane == 0) {
      r[2 * H2] = gb3[g];
      r[2 * H2 + 1] = g == 0 ? loss1 : loss2;
      r[2 * H2 + 2] = g == 0 ? clipped : 0.f;
      r[2 * H2 + 3] = 0.f;
    }
  }
  __syncthreads();
  float* out = a.part_head + (int64_t)blockIdx.x * 2 * a.head_row;
  for (int i = threadIdx.x; i < 2 * RW; i += blockDim.x) {
    float v = 0.f;
    for (int w = 0; w < 8; w++) v += red[w * 2 * RW + i];
    out[(i / RW) * a.head_row + (i % RW)] = v;
  }
}

// dW1aug[k][c] = sum_s D1[s][k] x[s][c] (c = 6: ones -> db1).  thread = feature k, block = contiguous sample range.
// grid (blocks, 2 nets); partial row [blk][g][k * 8 + c]
__global__ void __launch_bounds__(256) k_wide_dw1(const WideArgs a, int samples_per_block) {
  __shared__ float xs[64][S + 2];
  const int H1 = a.d.H1, g = blockIdx.y;
  const int sb = blockIdx.x * samples_per_block;
  const int se = min(sb + samples_per_block, a.count);
  float acc[S + 1];
#pragma unroll
  for (int c = 0; c <= S; c++) acc[c] = 0.f;
  const int k = threadIdx.x;
  for (int t0 = sb; t0 < se; t0 += 64) {
    const int nt = min(64, se - t0);
    __syncthreads();
    for (int i = threadIdx.x; i < nt * S; i += blockDim.x) {
      const int s = i / S, c = i % S;
      xs[s][c] = a.tr.obs[c * a.tr.cap + sample_pos(a, t0 + s)];
    }
    __syncthreads();
    if (k < H1) {
      const float* dp = a.d1 + ((int64_t)g * a.cap + t0) * H1 + k;
#pragma unroll 4
      for (int s = 0; s < nt; s++) {
        const float dv = dp[(int64_t)s * H1];
#pragma unroll
        for (int c = 0; c < S; c++) acc[c] = fmaf(dv, xs[s][c], acc[c]);
        acc[S] += dv;
      }
    }
  }
  if (k < H1) {
    float* out = a.part_dw1 + ((int64_t)blockIdx.x * 2 + g) * H1 * 8 + k * 8;
#pragma unroll
    for (int c = 0; c <= S; c++) out[c] = acc[c];
    out[S + 1] = 0.f;
  }
}

struct ReduceArgs {
  NetDims d;
  UpdateArgs u;
  const float *part_head, *part_dw1, *part_dw2;
  int head_blocks, head_row, dw1_blocks, splits, pstride;
  int first_chunk, last_chunk, apply_adam;
};

// same arithmetic as ppo_update.cu adam_apply (libtorch adam.cpp order), generic parameter count
__device__ __forceinline__ void wide_adam_apply(const UpdateArgs& a, int p, float g, int64_t step) {
  const double bc1 = 1.0 - pow(a.hp.beta1, (double)step);
  const double bc2 = 1.0 - pow(a.hp.beta2, (double)step);
  const float fb1 = (float)a.hp.beta1, f1mb1 = (float)(1.0 - a.hp.beta1);
  const float fb2 = (float)a.hp.beta2, f1mb2 = (float)(1.0 - a.hp.beta2);
  const float sq_bc2 = (float)sqrt(bc2), feps = (float)a.hp.eps, step_size = (float)(-(a.hp.lr / bc1));
  float m = a.adam_m[p], v = a.adam_v[p];
  m = __fadd_rn(__fmul_rn(m, fb1), __fmul_rn(f1mb1, g));
  v = __fadd_rn(__fmul_rn(v, fb2), __fmul_rn(f1mb2, __fmul_rn(g, g)));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), sq_bc2), feps);
  a.adam_m[p] = m;
  a.adam_v[p] = v;
  a.params[p] = __fadd_rn(a.params[p], __fmul_rn(step_size, __fdiv_rn(m, denom)));
}

__device__ __forceinline__ void wide_loss_log(const UpdateArgs& a, int q, float sum) {
  a.loss_log[a.step_slot * 4 + q] = (q == 1 ? 0.5 : 1.0) * (double)sum / (double)a.denom;
}

// fixed-order sum of every partial that feeds parameter p (or loss scalar q); accumulates over chunks; Adam on the last
__global__ void __launch_bounds__(128) k_wide_reduce(const ReduceArgs r) {
  const NetDims d = r.d;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= d.P + 3) return;
  float s = 0.f;
  int slot;
  if (p < d.P) {
    slot = p;
    const int g = p / d.SIZE, o = p % d.SIZE;
    if (o < d.oW2) {                       // W1 [H1][S] or b1 [H1]
      int k, c;
      if (o < d.ob1) { k = o / S; c = o % S; } else { k = o - d.ob1; c = S; }
      for (int b = 0; b < r.dw1_blocks; b++) s += r.part_dw1[((int64_t)b * 2 + g) * d.H1 * 8 + k * 8 + c];
    } else if (o < d.ob2) {                // W2 [H2][H1]
      const int e = o - d.oW2;
      const float* base = r.part_dw2 + (int64_t)g * r.splits * d.H2 * d.H1 + e;
      for (int sp = 0; sp < r.splits; sp++) s += base[(int64_t)sp * d.H2 * d.H1];
    } else {                               // b2, W3, b3 from the head partial rows
      int col;
      if (o < d.oW3) col = d.H2 + (o - d.ob2);
      else if (o < d.ob3) col = o - d.oW3;
      else col = 2 * d.H2;
      for (int b = 0; b < r.head_blocks; b++) s += r.part_head[((int64_t)b * 2 + g) * r.head_row + col];
    }
  } else {
    const int q = p - d.P;                 // 0 loss1 sum, 1 sum e^2, 2 clipped count
    slot = r.pstride - 4 + q;
    const int g = q == 1 ? 1 : 0, col = 2 * d.H2 + (q == 2 ? 2 : 1);
    for (int b = 0; b < r.head_blocks; b++) s += r.part_head[((int64_t)b * 2 + g) * r.head_row + col];
  }
  if (!r.first_chunk) s += r.u.grads[slot];
  r.u.grads[slot] = s;
  if (r.last_chunk && r.apply_adam) {
    if (p < d.P) wide_adam_apply(r.u, p, s, *r.u.adam_step);
    else wide_loss_log(r.u, p - d.P, s);
  }
}

__global__ void __launch_bounds__(128) k_wide_adam(const UpdateArgs a, int P, int pstride) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < P) wide_adam_apply(a, p, a.grads[p], *a.adam_step);
  else if (p < P + 3) wide_loss_log(a, p - P, a.grads[pstride - 4 + (p - P)]);
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
// Forward-only pass for wide nets (action query / value estimate / bootstrap): FP32 FMA, 16 rows per block.
// Not a throughput path: the host-boundary calls are PCIe-bound and bootstraps touch 1 record in n_timestep.
// obs element (row r, component c) = obs[c * stride_c + r * stride_r]; rows == nullptr => r = 0..n-1.
// ------------------------------------------------------------------------------------------------------------
namespace {
constexpr int FW_ROWS = 16;

__global__ void __launch_bounds__(256) k_wide_forward(const NetDims d, const float* __restrict__ net,
                                                      const float* __restrict__ obs, int64_t stride_c, int64_t stride_r,
                                                      const int32_t* __restrict__ rows, const int* __restrict__ n_dev,
                                                      int64_t n_host, int final_tanh, float* __restrict__ y,
                                                      int y_indexed) {
  extern __shared__ __align__(16) float fsm[];
  float* xs = fsm;                         // [16][8]
  float* h1s = xs + FW_ROWS * 8;           // [H1][16]
  float* ys = h1s + d.H1 * FW_ROWS;        // [8 warps][16]
  const int64_t n = n_dev ? (int64_t)*n_dev : n_host;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t r0 = (int64_t)blockIdx.x * FW_ROWS; r0 < n; r0 += (int64_t)gridDim.x * FW_ROWS) {
    __syncthreads();
    for (int i = threadIdx.x; i < FW_ROWS * S; i += blockDim.x) {
      const int s = i / S, c = i % S;
      float v = 0.f;
      if (r0 + s < n) {
        const int64_t r = rows ? (int64_t)rows[r0 + s] : (r0 + s);
        v = obs[c * stride_c + r * stride_r];
      }
      xs[s * 8 + c] = v;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < FW_ROWS * d.H1; e += blockDim.x) {
      const int s = e % FW_ROWS, k = e / FW_ROWS;
      float z = net[d.ob1 + k];
#pragma unroll
      for (int c = 0; c < S; c++) z = fmaf(net[d.oW1 + k * S + c], xs[s * 8 + c], z);
      h1s[k * FW_ROWS + s] = act_tanh(z);
    }
    __syncthreads();
    float part[FW_ROWS];
#pragma unroll
    for (int s = 0; s < FW_ROWS; s++) part[s] = 0.f;
    for (int j = threadIdx.x; j < d.H2; j += blockDim.x) {
      float acc[FW_ROWS];
      const float b = net[d.ob2 + j];
#pragma unroll
      for (int s = 0; s < FW_ROWS; s++) acc[s] = b;
      const float4* wrow = reinterpret_cast<const float4*>(net + d.oW2 + (int64_t)j * d.H1);
      const bool vec = (reinterpret_cast<uintptr_t>(wrow) & 15) == 0;   // the critic block starts at an odd float offset
      for (int k = 0; k < d.H1; k += 4) {
        float w[4];
        if (vec) { const float4 t = __ldg(wrow + (k >> 2)); w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w; }
        else { for (int u = 0; u < 4; u++) w[u] = net[d.oW2 + (int64_t)j * d.H1 + k + u]; }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const float4* hv = reinterpret_cast<const float4*>(h1s + (k + u) * FW_ROWS);
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const float4 t = hv[q];
            acc[4 * q + 0] = fmaf(w[u], t.x, acc[4 * q + 0]); acc[4 * q + 1] = fmaf(w[u], t.y, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(w[u], t.z, acc[4 * q + 2]); acc[4 * q + 3] = fmaf(w[u], t.w, acc[4 * q + 3]);
          }
        }
      }
      const float w3 = net[d.oW3 + j];
#pragma unroll
      for (int s = 0; s < FW_ROWS; s++) part[s] = fmaf(w3, act_tanh(acc[s]), part[s]);
    }
#pragma unroll
    for (int s = 0; s < FW_ROWS; s++) {
      float v = part[s];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) ys[warp * FW_ROWS + s] = v;
    }
    __syncthreads();
    if (threadIdx.x < FW_ROWS && r0 + threadIdx.x < n) {
      float v = net[d.ob3];
      for (int w = 0; w < 8; w++) v += ys[w * FW_ROWS + threadIdx.x];
      if (final_tanh) v = act_tanh(v);
      const int64_t r = r0 + threadIdx.x;
      y[(y_indexed && rows) ? (int64_t)rows[r] : r] = v;
    }
  }
}
}  // namespace

cudaError_t launch_wide_forward(const NetDims& d, const float* net_params, const float* obs, int64_t obs_stride_c,
                                int64_t obs_stride_k, const int32_t* rows, int64_t n, int final_tanh, float* y,
                                cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  int64_t g = (n + FW_ROWS - 1) / FW_ROWS;
  if (g > 148 * 4) g = 148 * 4;
  const int smem = (FW_ROWS * 8 + d.H1 * FW_ROWS + 8 * FW_ROWS) * 4;
  k_wide_forward<<<(int)g, 256, smem, st>>>(d, net_params, obs, obs_stride_c, obs_stride_k, rows, nullptr, n, final_tanh, y, 0);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------------------
// Host side: buffers, tensor maps, the per-step sequence
// ------------------------------------------------------------------------------------------------------------
namespace {
__global__ void k_compact_flagged(const uint8_t* __restrict__ flags, int64_t B, int32_t* __restrict__ idx, int* counter) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < B; k += (int64_t)gridDim.x * blockDim.x)
    if (flags[k] & (B2_FLAG_TIMEUP | B2_FLAG_TRUNC)) idx[atomicAdd(counter, 1)] = (int32_t)k;
}
}  // namespace

struct WideNet {
  NetDims d;
  int sm_count = 148;
  int64_t cap = 0;                 // samples per chunk, multiple of 128
  __nv_bfloat16 *w2p = nullptr, *w2tp = nullptr, *h1p = nullptr, *h1tp = nullptr, *d2p = nullptr, *d2tp = nullptr;
  float *bias2 = nullptr, *h2 = nullptr, *d1 = nullptr;
  float *part_head = nullptr, *part_dw1 = nullptr, *part_dw2 = nullptr;
  int head_blocks = 0, head_row = 0, dw1_blocks = 0, splits_max = 0;
  CUtensorMap tm_h1, tm_w2, tm_d2, tm_w2t, tm_d2t, tm_h1t;
  int32_t* boot_idx = nullptr;
  int64_t boot_cap = 0;
  int* boot_count = nullptr;
};

int wide_pstride(int H1, int H2) {
  const NetDims d = NetDims::make(S, H1, H2);
  return ((d.P + 4 + 31) / 32) * 32;
}

int64_t wide_chunk_cap(const WideNet* w) { return w ? w->cap : 0; }

void wide_destroy(WideNet* w) {
  if (!w) return;
  cudaFree(w->w2p); cudaFree(w->w2tp); cudaFree(w->h1p); cudaFree(w->h1tp); cudaFree(w->d2p); cudaFree(w->d2tp);
  cudaFree(w->bias2); cudaFree(w->h2); cudaFree(w->d1);
  cudaFree(w->part_head); cudaFree(w->part_dw1); cudaFree(w->part_dw2);
  cudaFree(w->boot_idx); cudaFree(w->boot_count);
  delete w;
}

int wide_create(WideNet** out, int H1, int H2, int64_t max_chunk, int sm_count) {
  *out = nullptr;
  if (H1 < 64 || H2 < 64 || H1 > 256 || H2 > 256 || (H1 % 64) || (H2 % 64))
    return set_error(B2_ERR_INVALID, "wide nets need hidden widths in {64, 128, 192, 256} (got %dx%d)", H1, H2);
  WideNet* w = new (std::nothrow) WideNet();
  if (!w) return set_error(B2_ERR_INVALID, "out of host memory");
  w->d = NetDims::make(S, H1, H2);
  w->sm_count = sm_count;
  const int64_t CAP_LIMIT = 131072;   // 1.6 GB of activations at 2x256; larger minibatches are walked in chunks
  int64_t cap = max_chunk < CAP_LIMIT ? max_chunk : CAP_LIMIT;
  cap = (cap + 127) / 128 * 128;
  w->cap = cap;
  const size_t bf = sizeof(__nv_bfloat16);
#define WCK(expr)                                                                                            \
  do {                                                                                                       \
    cudaError_t _e = (expr);                                                                                 \
    if (_e != cudaSuccess) {                                                                                 \
      wide_destroy(w);                                                                                       \
      return set_error(B2_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e));                                \
    }                                                                                                        \
  } while (0)
  WCK(cudaMalloc(&w->w2p, bf * 4 * H2 * H1));
  WCK(cudaMalloc(&w->w2tp, bf * 4 * H2 * H1));
  WCK(cudaMalloc(&w->h1p, bf * 4 * cap * H1));
  WCK(cudaMalloc(&w->h1tp, bf * 4 * cap * H1));
  WCK(cudaMalloc(&w->d2p, bf * 4 * cap * H2));
  WCK(cudaMalloc(&w->d2tp, bf * 4 * cap * H2));
  WCK(cudaMalloc(&w->bias2, sizeof(float) * 2 * H2));
  WCK(cudaMalloc(&w->h2, sizeof(float) * 2 * cap * H2));
  WCK(cudaMalloc(&w->d1, sizeof(float) * 2 * cap * H1));
  // stale bits must never be NaN where a zero-weighted product could meet them
  WCK(cudaMemset(w->h1p, 0, bf * 4 * cap * H1));
  WCK(cudaMemset(w->h1tp, 0, bf * 4 * cap * H1));
  WCK(cudaMemset(w->d2p, 0, bf * 4 * cap * H2));
  WCK(cudaMemset(w->d2tp, 0, bf * 4 * cap * H2));
  w->head_blocks = 2 * sm_count;
  w->head_row = 2 * H2 + 4;
  w->dw1_blocks = sm_count;
  w->splits_max = sm_count;
  WCK(cudaMalloc(&w->part_head, sizeof(float) * (size_t)w->head_blocks * 2 * w->head_row));
  WCK(cudaMalloc(&w->part_dw1, sizeof(float) * (size_t)w->dw1_blocks * 2 * H1 * 8));
  WCK(cudaMalloc(&w->part_dw2, sizeof(float) * (size_t)2 * w->splits_max * H2 * H1));
  WCK(cudaMalloc(&w->boot_count, sizeof(int)));
#undef WCK
  int rc = 0;
  const uint64_t ucap = (uint64_t)cap;
  // sample-major activations: K = features, rows = samples ; weights: K = in-features, rows = out-features
  if (!rc) rc = make_operand_map(&w->tm_h1, w->h1p, H1, ucap, 4, H1, ucap * H1, 128);
  if (!rc) rc = make_operand_map(&w->tm_w2, w->w2p, H1, H2, 4, H1, (uint64_t)H2 * H1, (uint32_t)H2);
  if (!rc) rc = make_operand_map(&w->tm_d2, w->d2p, H2, ucap, 4, H2, ucap * H2, 128);
  if (!rc) rc = make_operand_map(&w->tm_w2t, w->w2tp, H2, H1, 4, H2, (uint64_t)H2 * H1, (uint32_t)H1);
  // feature-major activations: K = samples
  if (!rc) rc = make_operand_map(&w->tm_d2t, w->d2tp, ucap, H2, 4, ucap, ucap * H2, 128);
  if (!rc) rc = make_operand_map(&w->tm_h1t, w->h1tp, ucap, H1, 4, ucap, ucap * H1, (uint32_t)H1);
  if (rc) { wide_destroy(w); return rc; }
  *out = w;
  return B2_OK;
}

template <int NJ>
static cudaError_t launch_head(const WideArgs& wa, int grid, cudaStream_t st) {
  constexpr int H2 = NJ * 32;
  const int smem = 4 * H2 * 17 * 4 + 8 * 2 * (2 * H2 + 4) * 4;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(k_wide_head<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    attr = true;
  }
  k_wide_head<NJ><<<grid, 256, smem, st>>>(wa);
  return cudaGetLastError();
}

int wide_step(WideNet* w, const UpdateArgs& a, bool apply_adam, cudaStream_t st, int64_t* launches) {
  const NetDims& d = w->d;
  const int H1 = d.H1, H2 = d.H2;
  WideArgs wa{};
  wa.d = d; wa.params = a.params; wa.tr = a.tr; wa.idx = a.idx; wa.cap = w->cap;
  wa.inv_denom = 1.0f / (float)a.denom; wa.hp = a.hp;
  wa.w2p = w->w2p; wa.w2tp = w->w2tp; wa.h1p = w->h1p; wa.h1tp = w->h1tp; wa.d2p = w->d2p; wa.d2tp = w->d2tp;
  wa.bias2 = w->bias2; wa.h2 = w->h2; wa.d1 = w->d1;
  wa.part_head = w->part_head; wa.part_dw1 = w->part_dw1; wa.head_row = w->head_row;
  cudaError_t ce;
#define LCK(what)                                                                              \
  do {                                                                                         \
    if ((ce = cudaGetLastError()) != cudaSuccess)                                              \
      return set_error(B2_ERR_CUDA, "wide %s: %s", what, cudaGetErrorString(ce));              \
    (*launches)++;                                                                             \
  } while (0)
  k_wide_prep<<<128, 256, 0, st>>>(wa, a.adam_step);
  LCK("prep");
  const int64_t n_chunks = (a.count + w->cap - 1) / w->cap;
  for (int64_t c = 0; c < n_chunks; c++) {
    wa.first = a.first + c * w->cap;
    wa.count = (int)std::min<int64_t>(w->cap, a.count - c * w->cap);
    wa.m_pad = (wa.count + 127) / 128 * 128;
    const int m_tiles = wa.m_pad / 128;
    // layer 1
    k_wide_l1<<<dim3(wa.m_pad / TILE_S, 2), 256, 0, st>>>(wa);
    LCK("l1");
    // H2 = tanh(H1 W2^T + b2)
    GemmArgs g{};
    g.G = 2; g.M = wa.count; g.N = H2; g.K = H1; g.m_tiles = m_tiles; g.splits = 1; g.kb_per_split = H1 / 64;
    g.stages = gemm_x3_stages(H2);
    g.out = w->h2; g.out_g_stride = w->cap * H2; g.bias = w->bias2;
    if ((ce = launch_gemm_x3(EPI_TANH, w->tm_h1, w->tm_w2, g, w->sm_count, st)) != cudaSuccess)
      return set_error(B2_ERR_CUDA, "wide gemm fwd: %s", cudaGetErrorString(ce));
    (*launches)++;
    // loss head, D2
    {
      const int n_tiles = wa.m_pad / TILE_S;
      const int grid = n_tiles < w->head_blocks ? n_tiles : w->head_blocks;
      if (grid < w->head_blocks)   // rows of blocks that do not run this time must read as zero
        if ((ce = cudaMemsetAsync(w->part_head, 0, sizeof(float) * (size_t)w->head_blocks * 2 * w->head_row, st)) != cudaSuccess)
          return set_error(B2_ERR_CUDA, "wide memset: %s", cudaGetErrorString(ce));
      switch (H2 / 32) {
        case 2: ce = launch_head<2>(wa, grid, st); break;
        case 4: ce = launch_head<4>(wa, grid, st); break;
        case 6: ce = launch_head<6>(wa, grid, st); break;
        default: ce = launch_head<8>(wa, grid, st); break;
      }
      if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "wide head: %s", cudaGetErrorString(ce));
      (*launches)++;
    }
    // D1 = (D2 W2) .* (1 - H1^2)
    g = GemmArgs{};
    g.G = 2; g.M = wa.count; g.N = H1; g.K = H2; g.m_tiles = m_tiles; g.splits = 1; g.kb_per_split = H2 / 64;
    g.stages = gemm_x3_stages(H1);
    g.out = w->d1; g.out_g_stride = w->cap * H1; g.hpl = w->h1p; g.h_plane_stride = w->cap * H1;
    if ((ce = launch_gemm_x3(EPI_DTANH, w->tm_d2, w->tm_w2t, g, w->sm_count, st)) != cudaSuccess)
      return set_error(B2_ERR_CUDA, "wide gemm bwd: %s", cudaGetErrorString(ce));
    (*launches)++;
    // dW2 partials = D2^T H1 over sample splits
    const int kb_total = wa.m_pad / 64;
    const int jt = (H2 + 127) / 128;
    int splits = w->sm_count / (2 * jt);
    if (splits > kb_total) splits = kb_total;
    if (splits > w->splits_max) splits = w->splits_max;
    if (splits < 1) splits = 1;
    const int kbps = (kb_total + splits - 1) / splits;
    splits = (kb_total + kbps - 1) / kbps;
    g = GemmArgs{};
    g.G = 2; g.M = H2; g.N = H1; g.K = wa.m_pad; g.m_tiles = jt; g.splits = splits; g.kb_per_split = kbps;
    g.stages = gemm_x3_stages(H1);
    g.out = w->part_dw2; g.out_g_stride = (int64_t)splits * H2 * H1;
    if ((ce = launch_gemm_x3(EPI_RAW, w->tm_d2t, w->tm_h1t, g, w->sm_count, st)) != cudaSuccess)
      return set_error(B2_ERR_CUDA, "wide gemm dW2: %s", cudaGetErrorString(ce));
    (*launches)++;
    // dW1, db1
    const int spb = (wa.count + w->dw1_blocks - 1) / w->dw1_blocks;
    k_wide_dw1<<<dim3(w->dw1_blocks, 2), 256, 0, st>>>(wa, spb < 1 ? 1 : spb);
    LCK("dw1");
    // reduce (+ Adam on the last chunk)
    ReduceArgs r{};
    r.d = d; r.u = a; r.part_head = w->part_head; r.part_dw1 = w->part_dw1; r.part_dw2 = w->part_dw2;
    r.head_blocks = w->head_blocks; r.head_row = w->head_row; r.dw1_blocks = w->dw1_blocks; r.splits = splits;
    r.pstride = wide_pstride(H1, H2);
    r.first_chunk = c == 0; r.last_chunk = c == n_chunks - 1; r.apply_adam = apply_adam ? 1 : 0;
    k_wide_reduce<<<(d.P + 3 + 127) / 128, 128, 0, st>>>(r);
    LCK("reduce");
  }
#undef LCK
  return B2_OK;
}

cudaError_t launch_wide_adam(WideNet* w, const UpdateArgs& a, cudaStream_t st) {
  k_wide_adam<<<(w->d.P + 3 + 127) / 128, 128, 0, st>>>(a, w->d.P, wide_pstride(w->d.H1, w->d.H2));
  return cudaGetLastError();
}

namespace {
__global__ void k_wide_act_finish(const float* __restrict__ mu, const float* __restrict__ eps_in, int64_t n, uint64_t seed,
                                  uint64_t call, float* __restrict__ action, float* __restrict__ logp, HyperDev hp,
                                  int* status) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float m = mu[i];
    const float eps = eps_in ? eps_in[i] : gauss_noise(seed ^ 0xA5A5A5A5A5A5A5A5ull, (uint64_t)i, call);
    const float a = (float)__dadd_rn((double)m, __dmul_rn(hp.sigma_d, (double)eps));  // src/ppo_nn.cpp:107-108
    action[i] = a;
    logp[i] = gauss_logprob(a, m, hp.inv_var, hp.half_log_norm);                      // :115
    if (!isfinite(a)) atomicOr(status, 1);
  }
}
}  // namespace

cudaError_t launch_wide_act_finish(const float* mu, const float* eps, int64_t n, uint64_t seed, uint64_t call, float* action,
                                   float* logp, const HyperDev& hp, int* status, cudaStream_t st) {
  int64_t g = (n + 255) / 256;
  if (g > 148 * 8) g = 148 * 8;
  k_wide_act_finish<<<(int)g, 256, 0, st>>>(mu, eps, n, seed, call, action, logp, hp, status);
  return cudaGetLastError();
}

// V(obs_k) for the flagged records of the trajectory (src/ppo_nn.cpp:226-228), wide nets
int wide_bootstrap(WideNet* w, const float* params, const Traj& tr, cudaStream_t st, int64_t* launches) {
  const int64_t B = tr.n_envs * tr.horizon;
  if (B > w->boot_cap) {
    cudaFree(w->boot_idx);
    w->boot_idx = nullptr;
    B2_CUDA_CHECK(cudaMalloc(&w->boot_idx, sizeof(int32_t) * B));
    w->boot_cap = B;
  }
  B2_CUDA_CHECK(cudaMemsetAsync(w->boot_count, 0, sizeof(int), st));
  k_compact_flagged<<<w->sm_count * 4, 256, 0, st>>>(tr.flags, B, w->boot_idx, w->boot_count);
  const NetDims& d = w->d;
  const int smem = (FW_ROWS * 8 + d.H1 * FW_ROWS + 8 * FW_ROWS) * 4;
  k_wide_forward<<<w->sm_count * 2, 256, smem, st>>>(d, params + d.SIZE, tr.obs, tr.cap, 1, w->boot_idx, w->boot_count, 0,
                                                     0, tr.vboot, 1);
  cudaError_t ce = cudaGetLastError();
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "wide bootstrap: %s", cudaGetErrorString(ce));
  *launches += 2;
  return B2_OK;
}

}  // namespace b2
