// wide_reduce.cuh -- the last stage of a wide-net optimizer step: fixed-order sums of the partial rows, the multi-GPU
// gradient exchange, Adam (libtorch operation order) and the refresh of the bf16 weight planes.
//   k_wide_reduce (wide_update.cu): wide_reduce_small (coalesced column sums of the dW1 / head partial rows) +
//     wide_reduce_finish2 (W2, two elements per thread) -- one wave of 128-thread blocks; the fixed-order dW2 mode
//     (B2_WIDE_DETERMINISTIC=1) takes its W2 blocks from wide_reduce_body<false>.
//   tail of the dW2 GEMM kernel (wide_gemm.cu, opt-in B2_WIDE_FUSE_REDUCE=1): wide_reduce_body<true> over virtual
//     blocks behind a grid barrier, one launch less (measured slower, kept measurable).
// Both end in wide_reduce_finish: chunk accumulation, exchange, Adam, planes.
#pragma once
#include "tc_ptx.cuh"
#include "wide.h"

namespace b2 {
namespace {

constexpr int S = OBS;

// libtorch adam.cpp: scalars are doubles narrowed to f32 inside the tensor kernels
__device__ __forceinline__ AdamCoef adam_coef(const HyperDev& hp, int64_t step) {
  const double bc1 = 1.0 - pow(hp.beta1, (double)step);
  const double bc2 = 1.0 - pow(hp.beta2, (double)step);
  AdamCoef c;
  c.fb1 = (float)hp.beta1; c.f1mb1 = (float)(1.0 - hp.beta1);
  c.fb2 = (float)hp.beta2; c.f1mb2 = (float)(1.0 - hp.beta2);
  c.sq_bc2 = (float)sqrt(bc2); c.feps = (float)hp.eps; c.step_size = (float)(-(hp.lr / bc1));
  return c;
}

// a parameter changed: keep the bf16 planes of W2 and the aligned copy of (b2, W3) the fused head reads in step with it
__device__ __forceinline__ void wide_store_planes(const NetDims& d, __nv_bfloat16* w2p, float* bias2,
                                                  int p, float val) {
  const int g = p / d.SIZE, o = p % d.SIZE;
  if (o >= d.oW2 && o < d.ob2) {
    const int e = o - d.oW2;
    __nv_bfloat16 hi, mid;
    tc::split_bf16(val, hi, mid);
    const int64_t ps = (int64_t)d.H2 * d.H1;
    w2p[(g * 2 + 0) * ps + e] = hi;
    w2p[(g * 2 + 1) * ps + e] = mid;
  } else if (o >= d.ob2 && o < d.ob3) {                   // b2 | W3 are adjacent in the parameter vector, and in headw[g][2][H2]
    bias2[g * 2 * d.H2 + (o - d.ob2)] = o < d.oW3 ? val * HEADW_B2_SCALE : val;   // b2 pre-scaled for the head's ex2 (wide_fused.cu)
  }
}

// same arithmetic as ppo_update.cu adam_apply (libtorch adam.cpp order), generic parameter count
__device__ __forceinline__ float wide_adam_math(const AdamCoef& c, float g, float& m, float& v, float w) {
  m = __fadd_rn(__fmul_rn(m, c.fb1), __fmul_rn(c.f1mb1, g));                 // exp_avg.mul_(b1).add_(grad, 1-b1)
  v = __fadd_rn(__fmul_rn(v, c.fb2), __fmul_rn(c.f1mb2, __fmul_rn(g, g)));   // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), c.sq_bc2), c.feps); // (sqrt(v)/sqrt(bc2)).add_(eps)
  return __fadd_rn(w, __fmul_rn(c.step_size, __fdiv_rn(m, denom)));          // p.addcdiv_(m, denom, -step_size)
}
__device__ __forceinline__ float wide_adam_apply(const UpdateArgs& a, const AdamCoef& c, int p, float g) {
  float m = a.adam_m[p], v = a.adam_v[p];
  const float np = wide_adam_math(c, g, m, v, a.params[p]);
  a.adam_m[p] = m;
  a.adam_v[p] = v;
  a.params[p] = np;
  return np;
}

__device__ __forceinline__ void wide_loss_log(const UpdateArgs& a, int q, float sum) {
  a.loss_log[a.step_slot * 4 + q] = (q == 1 ? 0.5 : 1.0) * (double)sum / (double)a.denom;
}

// Fixed-order (bit-reproducible) sum of every partial that feeds a parameter; accumulates over chunks; Adam on the last.
// Blocks [0, w2_blocks): one thread per W2 element, split-K partials summed 8 loads at a time.
// Remaining blocks: one WARP per "small" parameter (W1, b1, b2, W3, b3 and the 3 loss scalars), lanes stride over the
// per-block partial rows and a shuffle tree finishes the sum.
// One finished element of the reduce: p = parameter index (or d.P + l for the loss scalars), slot = its place in grads.
struct ReduceElem { bool valid; int p, slot; float s; };

// vb: the virtual block (128 elements of W2, or 4 small parameters) this call works on; sum = false: indices only
template <bool ADDED>   // ADDED: r.dw2_added known at compile time (drops the 40-partial path: 72 -> <= 32 registers, one wave)
__device__ __forceinline__ ReduceElem wide_reduce_body(const ReduceArgs& r, int w2_blocks, int vb, bool sum) {
  const NetDims d = r.d;
  const ReduceElem none{false, 0, 0, 0.f};
  const int w2n = d.H2 * d.H1;
  if (vb < w2_blocks) {
    const int i = vb * blockDim.x + threadIdx.x;
    if (i >= 2 * w2n) return none;
    const int g = i / w2n, e = i % w2n;
    if (!sum) return ReduceElem{true, g * d.SIZE + d.oW2 + e, g * d.SIZE + d.oW2 + e, 0.f};
    if (ADDED) {                               // already summed by the L2; leave zeros behind for the next step's adds
      float* one = r.part_dw2 + (int64_t)g * w2n + e;
      const float s1 = *one;
      *one = 0.f;
      const int p1 = g * d.SIZE + d.oW2 + e;
      return ReduceElem{true, p1, p1, s1};
    }
    const float* base = r.part_dw2 + (int64_t)g * r.splits * w2n + e;
    float s = 0.f;
    int sp = 0;
    if (r.splits <= 40) {                      // the usual case: every partial of this element in flight at once
      float t[40];
#pragma unroll
      for (int u = 0; u < 40; u++) t[u] = (u < r.splits) ? base[(int64_t)u * w2n] : 0.f;
#pragma unroll
      for (int u = 0; u < 40; u++) s += t[u];  // adding the zero padding does not change the sum
      sp = r.splits;
    }
    for (; sp + 16 <= r.splits; sp += 16) {
      float t[16];
#pragma unroll
      for (int u = 0; u < 16; u++) t[u] = base[(int64_t)(sp + u) * w2n];
#pragma unroll
      for (int u = 0; u < 16; u++) s += t[u];
    }
    for (; sp + 4 <= r.splits; sp += 4) {
      float t[4];
#pragma unroll
      for (int u = 0; u < 4; u++) t[u] = base[(int64_t)(sp + u) * w2n];
#pragma unroll
      for (int u = 0; u < 4; u++) s += t[u];
    }
    for (; sp < r.splits; sp++) s += base[(int64_t)sp * w2n];
    const int p = g * d.SIZE + d.oW2 + e;
    return ReduceElem{true, p, p, s};
  }
  const int lane = threadIdx.x & 31;
  const int per_net = d.H1 * (S + 1) + 2 * d.H2 + 1;     // W1, b1, b2, W3, b3
  const int q = (vb - w2_blocks) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (q >= 2 * per_net + 3) return none;
  const float* src;
  int64_t stride;
  int rows, p, slot;
  if (q < 2 * per_net) {
    const int g = q / per_net, o = q % per_net;
    if (o < d.H1 * (S + 1)) {                              // dW1aug[k][c], c == S is b1
      const int k = o / (S + 1), c = o % (S + 1);
      src = r.part_dw1 + (int64_t)g * d.H1 * 8 + k * 8 + c;
      stride = (int64_t)2 * d.H1 * 8; rows = r.dw1_blocks;
      p = g * d.SIZE + (c < S ? d.oW1 + k * S + c : d.ob1 + k);
    } else {
      const int o2 = o - d.H1 * (S + 1);                   // head row: [0,H2) dW3, [H2,2H2) db2, 2H2 db3
      src = r.part_head + (int64_t)g * r.head_row + o2;
      stride = (int64_t)2 * r.head_row; rows = r.head_blocks;
      p = g * d.SIZE + (o2 < d.H2 ? d.oW3 + o2 : (o2 < 2 * d.H2 ? d.ob2 + (o2 - d.H2) : d.ob3));
    }
    slot = p;
  } else {
    const int l = q - 2 * per_net;                         // 0 loss1 sum, 1 sum e^2, 2 clipped count
    const int g = l == 1 ? 1 : 0, col = 2 * d.H2 + (l == 2 ? 2 : 1);
    src = r.part_head + (int64_t)g * r.head_row + col;
    stride = (int64_t)2 * r.head_row; rows = r.head_blocks;
    p = d.P + l; slot = r.pstride - 4 + l;
  }
  if (!sum) return ReduceElem{lane == 0, p, slot, 0.f};
  float s = 0.f;
  for (int b0 = lane; b0 < rows; b0 += 32 * 8) {     // 8 independent loads per lane per round
    float t[8];
#pragma unroll
    for (int u = 0; u < 8; u++) t[u] = (b0 + 32 * u < rows) ? src[(int64_t)(b0 + 32 * u) * stride] : 0.f;
#pragma unroll
    for (int u = 0; u < 8; u++) s += t[u];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return ReduceElem{lane == 0, p, slot, s};
}

// what happens to one finished element: chunked accumulation, the multi-GPU exchange, Adam / loss log on the last chunk
__device__ __forceinline__ void wide_reduce_finish(const ReduceArgs& r, ReduceElem e) {
  if (!e.valid) return;
  if (!r.first_chunk) e.s += r.u.grads[e.slot];             // chunked gradient accumulation
  r.u.grads[e.slot] = e.s;
  if (!r.last_chunk) return;
  const bool peers = r.u.px.world > 1;   // several GPUs: exchange over NVLink peer memory right here (engine.h PeerXchg)
  bool ok = true;
  if (peers) e.s = px_allreduce_elem(r.u.px, e.slot, e.s, &ok);
  if (!ok || !(r.apply_adam || peers)) return;
  if (e.p < r.d.P) wide_store_planes(r.d, r.w2p, r.bias2, e.p, wide_adam_apply(r.u, *r.coef, e.p, e.s));
  else wide_loss_log(r.u, e.p - r.d.P, e.s);
}

// ---- k_wide_reduce's own paths (one wave of 128-thread blocks; the fused GEMM tail keeps wide_reduce_body) ----------

// Two finished W2 elements of one thread: every load of both is in flight before the first dependent instruction, and
// both exchange slots are pushed before either is awaited.
__device__ __forceinline__ void wide_reduce_finish2(const ReduceArgs& r, int p0, int p1, float s0, float s1) {
  if (!r.first_chunk) {
    const float g0 = r.u.grads[p0], g1 = r.u.grads[p1];
    s0 += g0; s1 += g1;
  }
  r.u.grads[p0] = s0;
  r.u.grads[p1] = s1;
  if (!r.last_chunk) return;
  const bool peers = r.u.px.world > 1;
  bool ok = true;
  if (peers && (ok = px_alive(r.u.px))) {
    px_push(r.u.px, p0, s0);
    px_push(r.u.px, p1, s1);
    s0 = px_pull(r.u.px, p0, &ok);
    s1 = px_pull(r.u.px, p1, &ok);
  }
  if (!ok || !(r.apply_adam || peers)) return;
  const AdamCoef c = *r.coef;
  float m0 = r.u.adam_m[p0], m1 = r.u.adam_m[p1], v0 = r.u.adam_v[p0], v1 = r.u.adam_v[p1];
  const float w0 = r.u.params[p0], w1 = r.u.params[p1];
  const float n0 = wide_adam_math(c, s0, m0, v0, w0), n1 = wide_adam_math(c, s1, m1, v1, w1);
  r.u.adam_m[p0] = m0; r.u.adam_m[p1] = m1;
  r.u.adam_v[p0] = v0; r.u.adam_v[p1] = v1;
  r.u.params[p0] = n0; r.u.params[p1] = n1;
  wide_store_planes(r.d, r.w2p, r.bias2, p0, n0);
  wide_store_planes(r.d, r.w2p, r.bias2, p1, n1);
}

constexpr int RED_THREADS = 128, RED_RG = 8, RED_COLS = 16, RED_INFLIGHT = 28;

// Fixed-order sum of one column of `rows` partial rows (`stride` floats apart), coalesced: the block is 8 row groups x
// 16 adjacent columns, so every half-warp reads 64 contiguous bytes of one row; a row group adds rows rg, rg + 8, ..
// (28 loads in flight), the groups are folded in order through shared memory.  The total is returned to row group 0.
__device__ __forceinline__ float wide_colsum(const float* __restrict__ src, bool live, int rows, int64_t stride,
                                             float* red) {
  const int rg = threadIdx.x / RED_COLS, cx = threadIdx.x % RED_COLS;
  float s = 0.f;
  for (int b0 = rg; b0 < rows; b0 += RED_RG * RED_INFLIGHT) {
    float t[RED_INFLIGHT];
#pragma unroll
    for (int u = 0; u < RED_INFLIGHT; u++) {
      const int b = b0 + RED_RG * u;
      t[u] = (live && b < rows) ? src[(int64_t)b * stride] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < RED_INFLIGHT; u++) s += t[u];
  }
  red[rg * RED_COLS + cx] = s;
  __syncthreads();
  float tot = 0.f;
  if (rg == 0) {
#pragma unroll
    for (int j = 0; j < RED_RG; j++) tot += red[j * RED_COLS + cx];
  }
  return tot;
}

__host__ __device__ inline int wide_reduce_dw1_blocks(const NetDims& d) { return (2 * d.H1 * 8 + RED_COLS - 1) / RED_COLS; }
__host__ __device__ inline int wide_reduce_head_blocks(const NetDims& d) { return 2 * ((2 * d.H2 + 3 + RED_COLS - 1) / RED_COLS); }

// block sb of the small-parameter part: [0, dw1 blocks) the dW1aug partial rows of the dW2 kernel's CTAs,
// then the head rows (dW3, db2, db3 and the three loss sums) of k_wide_head's blocks
__device__ __forceinline__ void wide_reduce_small(const ReduceArgs& r, int sb, float* red) {
  const NetDims d = r.d;
  const int cx = threadIdx.x % RED_COLS;
  const bool lead = threadIdx.x < RED_COLS;
  const int dw1b = wide_reduce_dw1_blocks(d);
  if (sb < dw1b) {
    const int x = sb * RED_COLS + cx;                      // flat [g][k][8] index inside a partial row
    const bool live = x < 2 * d.H1 * 8;
    const float s = wide_colsum(r.part_dw1 + x, live, r.dw1_blocks, (int64_t)2 * d.H1 * 8, red);
    const int c = x & 7, k = (x >> 3) % d.H1, g = x / (d.H1 * 8);
    if (!lead || !live || c > S) return;
    const int p = g * d.SIZE + (c < S ? d.oW1 + k * S + c : d.ob1 + k);
    wide_reduce_finish(r, ReduceElem{true, p, p, s});
    return;
  }
  const int ncol = 2 * d.H2 + 3;                           // [0,H2) dW3, [H2,2H2) db2, 2H2 db3, then the loss columns
  const int bpg = (ncol + RED_COLS - 1) / RED_COLS;
  const int hb = sb - dw1b, g = hb / bpg, col = (hb % bpg) * RED_COLS + cx;
  if (g >= 2) return;                                      // (whole block: no barrier is skipped by part of a block)
  const bool live = col < ncol;
  const float s = wide_colsum(r.part_head + (int64_t)g * r.head_row + col, live, r.head_blocks, (int64_t)2 * r.head_row, red);
  if (!lead || !live) return;
  int p, slot;
  if (col <= 2 * d.H2) {
    p = g * d.SIZE + (col < d.H2 ? d.oW3 + col : (col < 2 * d.H2 ? d.ob2 + (col - d.H2) : d.ob3));
    slot = p;
  } else {
    int l;                                                  // 0 loss1 sum (actor), 1 sum e^2 (critic), 2 clipped count (actor)
    if (col == 2 * d.H2 + 1) l = g == 0 ? 0 : 1;
    else if (g == 0) l = 2;
    else return;
    p = d.P + l; slot = r.pstride - 4 + l;
  }
  wide_reduce_finish(r, ReduceElem{true, p, slot, s});
}

}  // namespace
}  // namespace b2
