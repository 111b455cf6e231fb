// wide_update.cu -- PPO optimizer step for wide hidden layers (e.g. 2x256, BASELINE configs[4]): the FP32-FMA side
// kernels around the three tcgen05 GEMMs of wide_gemm.cu, the deterministic gradient reduction + Adam, and the
// host sequence that strings them together.  See wide.h for the dataflow.  Replaces ActorCritic::evaluate
// (src/ppo_nn.cpp:126-194), the loss (:263-290), autograd backward (:301-302) and Adam::step (:303).
//
// Memory plan for a chunk of CAP samples (CAP % 128 == 0), g = net (0 actor, 1 critic), pl = bf16 plane (hi, mid):
//   w2p  [g*2+pl][H2][H1]   bias2 [g][2][H2] = (b2 * 2 log2 e, W3)   (kept current by the Adam step)
//   h1p  [g*2+pl][CAP][H1]                                           (layer-1 activations, sample-major)
//   h2   [g][CAP][H2] f32                                            (layer-2 activations)
//   d2p  [g*2+pl][CAP][H2]                                           (dL/dz2, sample-major)
//   d1   [g][CAP][H1] f32                                            (dL/dz1)
// The layer-2 forward and dX GEMMs read h1p / d2p as K-major operands (K = features); the dW2 GEMM reads the SAME
// planes as MN-major operands (K = samples), so no transposed copy of any activation is ever written.
// Samples in [count, round_up(count, 128)) are written as zeros by the producers, so the sample-contraction GEMM
// (dW2) needs no masking.
#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

#include "tc_ptx.cuh"
#include "wide.h"
#include "mlp.cuh"
#include "wide_reduce.cuh"

namespace b2 {

namespace {


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
  __nv_bfloat16 *w2p, *h1p, *d2p;
  float *bias2, *h2, *d1;
  float *part_head, *part_dw1;   // [blocks][2][head_row] , [blocks][2][H1 * 8]
  int head_row;
  int64_t* bump_step;   // non-null on the first chunk of an optimizer step: adam_step += 1 ...
  AdamCoef* coef_out;   // ... and the Adam coefficients of the new step, for k_wide_reduce / k_wide_adam at its end
};

__device__ __forceinline__ int64_t sample_pos(const WideArgs& a, int s) {
  const int64_t p = a.first + s;
  return a.idx ? (int64_t)a.idx[p] : p;
}

// weights -> bf16 planes.  Runs at the start of an update; afterwards the Adam step itself rewrites the planes of every
// W2 element it changes (wide_store_planes), so the steady state has no conversion pass.  One copy serves both GEMMs:
// Z2 = H1 W2^T reads it K-major, Z1 = D2 W2 reads it as an MN-major operand.
__global__ void __launch_bounds__(256) k_wide_prep(const WideArgs a) {
  const int H1 = a.d.H1, H2 = a.d.H2;
  const int n = 2 * H2 * H1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int g = i / (H2 * H1), r = i % (H2 * H1);
    const float w = a.params[g * a.d.SIZE + a.d.oW2 + r];
    __nv_bfloat16 hi, mid;
    tc::split_bf16(w, hi, mid);
    const int64_t ps = (int64_t)H2 * H1;
    a.w2p[(g * 2 + 0) * ps + r] = hi;
    a.w2p[(g * 2 + 1) * ps + r] = mid;
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 4 * H2; i += gridDim.x * blockDim.x)   // headw[g][2][H2]: b2, W3
    a.bias2[i] = a.params[(i / (2 * H2)) * a.d.SIZE + a.d.ob2 + (i % (2 * H2))] * ((i % (2 * H2)) < H2 ? HEADW_B2_SCALE : 1.0f);
}

// layer 1: H1 = tanh(W1 x + b1) -> (hi, mid) bf16 planes, sample-major.  grid (blocks, 2 nets); a block owns an equal
// share of the m_pad rows of one net (balanced to a row: L1_CTAS resident blocks on every SM, one wave) and walks it in
// tiles of 64 samples.  A thread owns 4 consecutive features -- their 24 weights and 4 biases live in registers for the
// whole block, so the inner loop reads only the 6 observation values (a shared-memory broadcast) -- and strides over the
// samples; a warp's stores are 256 contiguous bytes of one row per plane.
constexpr int L1_TILE = 64;
constexpr int L1_CTAS = 4;
__global__ void __launch_bounds__(256, L1_CTAS) k_wide_l1(const WideArgs a) {
  __shared__ float xs[L1_TILE][S + 2];
  __shared__ __align__(16) float ws[64 * 28];   // per feature group of 4: 24 weights + 4 biases (28: conflict-free LDS.128)
  const int H1 = a.d.H1, g = blockIdx.y;
  const int lo = (int)((int64_t)blockIdx.x * a.m_pad / gridDim.x), hi = (int)((int64_t)(blockIdx.x + 1) * a.m_pad / gridDim.x);
  auto load_tile = [&](int r0) {
    const int n = min(L1_TILE, hi - r0);
    for (int i = threadIdx.x; i < n * S; i += blockDim.x) {
      const int r = i / S, c = i % S;
      xs[r][c] = (r0 + r < a.count) ? a.tr.obs[c * a.tr.cap + sample_pos(a, r0 + r)] : 0.0f;
    }
  };
  pdl_launch_dependents();
  load_tile(lo);   // the observations are constant during an update: fetched while the previous step's reduce still runs
  pdl_wait();                                   // the weights come from the previous step's Adam
  if (blockIdx.x == 0 && g == 0 && threadIdx.x == 0 && a.bump_step) {   // one optimizer step
    const int64_t step = *a.bump_step + 1;
    *a.bump_step = step;
    *a.coef_out = adam_coef(a.hp, step);   // two fp64 pow() on one thread here instead of a barrier in every reduce block
  }
  for (int i = threadIdx.x; i < H1 * S; i += blockDim.x) ws[(i / 24) * 28 + (i % 24)] = a.params[g * a.d.SIZE + a.d.oW1 + i];
  for (int i = threadIdx.x; i < H1; i += blockDim.x) ws[(i / 4) * 28 + 24 + (i % 4)] = a.params[g * a.d.SIZE + a.d.ob1 + i];
  __syncthreads();
  const int64_t plane = a.cap * H1;
  const int groups = H1 / 4;                    // feature groups of 4
  const int kg = threadIdx.x % groups, sl = threadIdx.x / groups, sstep = blockDim.x / groups;
  const bool worker = sl < sstep;               // H1 = 192: 48 groups, 5 sample lanes, 16 idle threads
  float w[24], b[4];
#pragma unroll
  for (int v = 0; v < 6; v++) *reinterpret_cast<float4*>(&w[4 * v]) = *reinterpret_cast<const float4*>(ws + kg * 28 + 4 * v);
  *reinterpret_cast<float4*>(b) = *reinterpret_cast<const float4*>(ws + kg * 28 + 24);
  // pre-scaled by 2 log2 e: tanh z = 1 - 2 / (2^(z K) + 1), and the four features of a thread share ONE reciprocal (tanh4):
  // the kernel is MUFU-bound (XU pipe the busiest unit, profiles/r1l_prof_wide_side.txt), 1.25 instead of 2 per element
#pragma unroll
  for (int v = 0; v < 24; v++) w[v] *= TANH_K;
#pragma unroll
  for (int v = 0; v < 4; v++) b[v] *= TANH_K;
  for (int r0 = lo; r0 < hi; r0 += L1_TILE) {
    if (r0 != lo) {
      __syncthreads();                          // the previous tile's readers are done
      load_tile(r0);
      __syncthreads();
    }
    const int n = min(L1_TILE, hi - r0);
    for (int r = sl; worker && r < n; r += sstep) {
      const bool valid = r0 + r < a.count;
      float x[S], z[4];
#pragma unroll
      for (int c = 0; c < S; c++) x[c] = xs[r][c];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        z[u] = b[u];
#pragma unroll
        for (int c = 0; c < S; c++) z[u] = fmaf(w[u * S + c], x[c], z[u]);
      }
      tanh4(z[0], z[1], z[2], z[3], z[0], z[1], z[2], z[3]);
      uint2 th, tm;
      split2(valid ? z[0] : 0.f, valid ? z[1] : 0.f, th.x, tm.x);
      split2(valid ? z[2] : 0.f, valid ? z[3] : 0.f, th.y, tm.y);
      const int64_t o = (int64_t)(r0 + r) * H1 + kg * 4;
      *reinterpret_cast<uint2*>(a.h1p + (g * 2 + 0) * plane + o) = th;
      *reinterpret_cast<uint2*>(a.h1p + (g * 2 + 1) * plane + o) = tm;
    }
  }
}

// ---- layer 1 of step s AND the reduce + Adam of step s-1 in one kernel (the default single-GPU sequence) ----------------
// Between the dW2 GEMM of one step and the forward GEMM of the next the tensor pipe used to wait for two launches in a row:
// k_wide_reduce (a chain of L2 round trips, 5.6 us) and k_wide_l1 (7.6 us) -- and only W1 / b1 of the reduce's output is
// what layer 1 needs.  Here the producers of a step leave SUMS instead of partial rows (k_gemm_mn2's dW1 warps and
// k_wide_fh's head add theirs into one row each in the L2, as the dW2 GEMM's split-K tiles always did), and this kernel
//   * applies Adam to W1 / b1 itself: a CTA owns 32 features of one net for a slice of the rows (grid = row slices x
//     feature slices x nets), so it needs 224 parameters -- one per thread, straight into the shared-memory weight table
//     the layer-1 loop reads.  Every row slice repeats that arithmetic (37 x at 2x256 on 148 SMs: 9 M of the kernel's
//     ~170 M instructions) instead of waiting for another CTA's result; row slice 0 writes it back.  The state is double
//     buffered (w1_rd -> w1_wr): CTAs of one launch must all read the OLD moments while slice 0 stores the new ones;
//   * spreads the rest of the reduce (W2, b2, W3, b3, the loss sums: one element per thread, grid-stride) over all CTAs
//     in front of the layer-1 loop: sum -> Adam -> parameter, bf16 planes, headw, exactly wide_reduce_finish;
//   * then runs layer 1 as k_wide_l1 does (same arithmetic per element; a warp stores 4 rows x 64 contiguous bytes).
// pending == 0 (first step of an update): layer 1 only, W1 from the canonical vector (seed), which also primes w1_wr.
// rows == 0 (wide_flush after the last step): the reduce + Adam only.
struct L1rArgs {
  WideArgs a;
  ReduceArgs r;            // of the PENDING step (grads, Adam state, loss log slot, coefficients, planes)
  int pending, seed;
  const float* dw1_sum;    // [2][H1 * 8] dW1aug sums of the pending step
  float* dw1_zero;         // the other buffer: zeroed here, the next dW2 kernel adds into it
  float* head_sum;         // [2][head_row] sums of the pending step's head; consumed and zeroed element by element
  const float* w1_rd;      // [2 nets][3: w, m, v][H1][S + 1] (column S = the bias)
  float* w1_wr;
  float *cparams, *cm, *cv, *cgrads;   // the canonical vectors (row slice 0 keeps W1 / b1 current there too)
  unsigned* w1_flag;       // PEERS: [2 nets][feature slices] -- 2 * epoch (| 1: abandoned) once the writer slice has left the summed W1 / b1 gradients of that epoch in cgrads
  long long* tl;           // optional whole-step timeline (tc::tl_stamp), diagnostic runs only
  int tl_k;                // its kernel row: 0 = this step's launch, 3 = the launch of the step after the stamped one
};
constexpr int L1R_FS = 32;        // features per CTA
constexpr int L1R_THREADS = 288;  // 8 layer-1 warps + the reduce warp
constexpr int L1R_RB = 4;         // elements a lane of the reduce warp keeps in flight
constexpr int L1R_TILE = 448;     // observation rows per shared-memory tile: a CTA's whole row slice at 16,384-sample minibatches
                                  // (443 rows on 148 SMs), fetched BEFORE the dependency wait -- with k_wide_l1's 64-row tiles
                                  // the slice took seven exposed gathers

// element e of the pending step's reduce (everything but W1 / b1) -> where its sum waits, its parameter, its gradient slot
__device__ __forceinline__ void l1r_locate(const L1rArgs& q, int e, float*& src, int& p, int& slot) {
  const NetDims& d = q.a.d;
  const int w2n = d.H2 * d.H1, hn = 2 * d.H2 + 1;
  if (e < 2 * w2n) {                                  // W2: the L2 has added the split-K tiles of the dW2 GEMM
    src = q.r.part_dw2 + e;
    p = slot = (e / w2n) * d.SIZE + d.oW2 + (e % w2n);
  } else if (e < 2 * w2n + 2 * hn) {                  // head row of net g2: [0,H2) dW3, [H2,2H2) db2, 2H2 db3
    const int h = e - 2 * w2n, g2 = h / hn, col = h % hn;
    src = q.head_sum + g2 * q.r.head_row + col;
    p = slot = g2 * d.SIZE + (col < d.H2 ? d.oW3 + col : (col < 2 * d.H2 ? d.ob2 + (col - d.H2) : d.ob3));
  } else {                                            // 0 loss1 sum (actor), 1 sum e^2 (critic), 2 clipped count (actor)
    const int l = e - 2 * w2n - 2 * hn;
    src = q.head_sum + (l == 1 ? 1 : 0) * q.r.head_row + 2 * d.H2 + (l == 2 ? 2 : 1);
    p = d.P + l; slot = q.r.pstride - 4 + l;
  }
}

// one finished (and, with several GPUs, exchanged) sum of the reduce: gradient vector, Adam, planes / loss log
__device__ __forceinline__ void l1r_apply(const L1rArgs& q, const AdamCoef& cf, int p, int slot, float sv, float m, float v, float w) {
  const NetDims& d = q.a.d;
  q.cgrads[slot] = sv;
  if (p < d.P) {
    const float nw = wide_adam_math(cf, sv, m, v, w);
    q.cm[p] = m; q.cv[p] = v; q.cparams[p] = nw;
    wide_store_planes(d, q.r.w2p, q.r.bias2, p, nw);
  } else {
    wide_loss_log(q.r.u, p - d.P, sv);
  }
}

// PEERS: several GPUs -- every finished sum is exchanged over NVLink peer memory first (engine.h PeerXchg, as k_wide_reduce
// does): the reduce warp pushes and pulls its elements; of the W1 / b1 threads the writer slice pushes, every slice pulls
// (all 592 CTAs are resident at once, and a thread waits only for stores that are issued before any wait)
template <bool PEERS>
__global__ void __launch_bounds__(L1R_THREADS, L1_CTAS) k_wide_l1r(const L1rArgs q) {
  __shared__ int s_ok;
  __shared__ __align__(16) float xs[L1R_TILE][2 * S];     // every component twice: (x, x) pairs are the FFMA2 operands of the loop
  __shared__ __align__(16) float ws[(L1R_FS / 4) * 28];   // per feature group of 4: 12 weight PAIRS (features 0|1, then 2|3, of
                                                          // component c) + 4 biases, pre-scaled by TANH_K
  const WideArgs& a = q.a;
  const NetDims& d = a.d;
  const int H1 = d.H1, g = blockIdx.z, k0 = blockIdx.y * L1R_FS, tid = threadIdx.x;
  const int n_cta = (int)(gridDim.x * gridDim.y * gridDim.z);
  const int cta = (int)((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x);
  if (tid == 0) tc::tl_stamp(q.tl, q.tl_k, 0);
  pdl_launch_dependents();

  // PEERS: the reduce is spread over EVERY thread of the grid, one element each (as k_wide_reduce's one wave of threads does):
  // all sums are on the wire before anybody waits, one NVLink latency for the lot, and that latency runs beside the W1
  // exchange of the layer-1 threads.  (A reduce warp with 8 elements per lane, as on one GPU, pulled them one after the other:
  // its chain ended 17-25 us after the layer-1 loop, profiles/r2/r2l_wide_timeline_2gpu.txt.)
  const int n_red = 2 * d.H2 * H1 + 2 * (2 * d.H2 + 1) + 3;
  float* osrc = nullptr;
  int op = 0, oslot = 0;
  float om = 0.f, ov = 0.f, ow = 0.f, osv = 0.f;
  bool olive = false, oalive = true;
  if (PEERS && q.pending) {
    const int e = cta * L1R_THREADS + tid;
    olive = e < n_red;
    if (olive) {
      l1r_locate(q, e, osrc, op, oslot);
      if (op < d.P) { om = q.cm[op]; ov = q.cv[op]; ow = q.cparams[op]; }   // older than the kernel in front: before the wait
    }
  }
  auto own_push = [&]() {                                  // after the dependency wait
    oalive = px_alive(q.r.u.px);
    if (olive) { osv = *osrc; *osrc = 0.f; if (oalive) px_push(q.r.u.px, oslot, osv); }
  };
  auto own_finish = [&](const AdamCoef& c) {
    if (olive && oalive) {
      bool ok = true;
      osv = px_pull(q.r.u.px, oslot, &ok);
      if (ok) l1r_apply(q, c, op, oslot, osv, om, ov, ow);
    }
    for (int e = cta * L1R_THREADS + tid + n_cta * L1R_THREADS; e < n_red; e += n_cta * L1R_THREADS) {   // (a grid smaller than the reduce)
      float* src; int p, slot;
      l1r_locate(q, e, src, p, slot);
      float sv = *src; *src = 0.f;
      bool ok = oalive;
      if (ok) { px_push(q.r.u.px, slot, sv); sv = px_pull(q.r.u.px, slot, &ok); }
      if (ok) l1r_apply(q, c, p, slot, sv, p < d.P ? q.cm[p] : 0.f, p < d.P ? q.cv[p] : 0.f, p < d.P ? q.cparams[p] : 0.f);
    }
  };

  if (tid >= 256) {
    // ---- the reduce warp: the pending step's W2, b2, W3, b3 and loss sums, 256 elements per CTA and round, beside the
    // layer-1 warps (its chain of L2 round trips hides under their loop).  Adam state first -- it is older than the kernel
    // in front -- then the dependency wait, then the sums.
    if (cta == 0 && tid == 256 && a.bump_step) {   // one optimizer step: the counter and the coefficients of THIS step's Adam (two
      const int64_t step = *a.bump_step + 1;       // fp64 pow() on one thread, off the layer-1 warps' path; the counter was last
      *a.bump_step = step;                         // written by the launch before the kernels in front: no wait needed)
      *a.coef_out = adam_coef(a.hp, step);
    }
    if (!q.pending) return;
    const int lane = tid - 256;
    const AdamCoef cf = *q.r.coef;
    if (PEERS) {                                           // this warp's threads own an element each like everybody else
      pdl_wait();
      own_push();
      own_finish(cf);
      if (tid == 256) tc::tl_stamp(q.tl, q.tl_k, 4);
      return;
    }
    bool waited = false;
    for (int e0 = cta * 256; e0 < n_red; e0 += n_cta * 256) {
#pragma unroll 1
      for (int b0 = 0; b0 < 256; b0 += 32 * L1R_RB) {
        float* src[L1R_RB];
        int p[L1R_RB], slot[L1R_RB];
        float m[L1R_RB], v[L1R_RB], w[L1R_RB], sv[L1R_RB];
        bool live[L1R_RB];
#pragma unroll
        for (int u = 0; u < L1R_RB; u++) {
          const int e = e0 + b0 + 32 * u + lane;
          live[u] = e < n_red;
          src[u] = nullptr; p[u] = 0; slot[u] = 0; m[u] = v[u] = w[u] = 0.f;
          if (live[u]) {
            l1r_locate(q, e, src[u], p[u], slot[u]);
            if (p[u] < d.P) { m[u] = q.cm[p[u]]; v[u] = q.cv[p[u]]; w[u] = q.cparams[p[u]]; }
          }
        }
        if (!waited) { pdl_wait(); waited = true; }
#pragma unroll
        for (int u = 0; u < L1R_RB; u++) sv[u] = live[u] ? *src[u] : 0.f;
        const bool ok = true;
#pragma unroll
        for (int u = 0; u < L1R_RB; u++) {
          if (!live[u]) continue;
          *src[u] = 0.f;                                  // the next step adds into zeros
          if (!ok) continue;                              // a peer never arrived: the update is abandoned (engine.h px_alive)
          q.cgrads[slot[u]] = sv[u];
          if (p[u] < d.P) {
            const float nw = wide_adam_math(cf, sv[u], m[u], v[u], w[u]);
            q.cm[p[u]] = m[u]; q.cv[p[u]] = v[u]; q.cparams[p[u]] = nw;
            wide_store_planes(d, q.r.w2p, q.r.bias2, p[u], nw);
          } else {
            wide_loss_log(q.r.u, p[u] - d.P, sv[u]);
          }
        }
      }
    }
    if (!waited) pdl_wait();
    if (tid == 256) tc::tl_stamp(q.tl, q.tl_k, 4);        // reduce warp done
    return;
  }

  const int lo = (int)((int64_t)blockIdx.x * a.m_pad / gridDim.x), hi = (int)((int64_t)(blockIdx.x + 1) * a.m_pad / gridDim.x);
  auto load_tile = [&](int r0) {
    const int n = min(L1R_TILE, hi - r0);
    static_assert(L1R_TILE <= 512, "two rows per thread and component");
    float v[2 * S];                                     // every load of a thread in flight before the first store (one L2 round
#pragma unroll                                          // trip instead of a dozen); a warp reads 32 consecutive rows of ONE component
    for (int u = 0; u < 2 * S; u++) {
      const int c = u >> 1, r = tid + 256 * (u & 1);
      v[u] = (r < n && r0 + r < a.count) ? a.tr.obs[c * a.tr.cap + sample_pos(a, r0 + r)] : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < 2 * S; u++) {
      const int c = u >> 1, r = tid + 256 * (u & 1);
      if (r < n) *reinterpret_cast<float2*>(&xs[r][2 * c]) = make_float2(v[u], v[u]);
    }
  };
  if (hi > lo) load_tile(lo);   // the observations are constant during an update
  // ---- this thread's W1 / b1 parameter: feature k0 + tid / 7, column tid % 7.  Its state is older than the kernel in front
  // (the previous k_wide_l1r launch or the canonical vector): fetched before the dependency wait
  const bool w1t = tid < L1R_FS * (S + 1);
  const int kl = tid / (S + 1), wc = tid % (S + 1);
  const int pcan = g * d.SIZE + (wc < S ? d.oW1 + (k0 + kl) * S + wc : d.ob1 + k0 + kl);
  const int64_t wblk = (int64_t)H1 * (S + 1), wi = (int64_t)(k0 + kl) * (S + 1) + wc;
  float pw = 0.f, pm = 0.f, pv = 0.f;
  if (w1t) {
    if (q.seed) { pw = q.cparams[pcan]; pm = q.cm[pcan]; pv = q.cv[pcan]; }
    else { pw = q.w1_rd[(g * 3 + 0) * wblk + wi]; pm = q.w1_rd[(g * 3 + 1) * wblk + wi]; pv = q.w1_rd[(g * 3 + 2) * wblk + wi]; }
  }
  AdamCoef cf{};
  if (q.pending) cf = *q.r.coef;   // of the pending step: written by the launch that ran ITS layer 1
  pdl_wait();                      // the sums come from the kernels of the pending step
  if (tid == 0) tc::tl_stamp(q.tl, q.tl_k, 1);
  const bool writer = blockIdx.x == 0;
  // PEERS: only the writer slice of a (net, feature slice) exchanges its 224 sums with the peers; the other 36 slices wait for
  // ONE flag word of that slice (a thread per CTA polls it) and read the summed gradients it left in the gradient vector.
  // (Every slice pulling for itself was measured first: 132k threads polling the same 28 KB of slots queue on a few L2
  // slices in front of the very NVLink stores they wait for -- 32.1 instead of 27.5 ms per update on two GPUs.)
  float gsum = 0.f;
  bool xok = true;
  if (PEERS && q.pending) {
    const PeerXchg& px = q.r.u.px;
    unsigned* flag = q.w1_flag + g * gridDim.y + blockIdx.y;
    const unsigned want = 2u * px.epoch;                   // | 1: the exchange of this step was abandoned
    if (tid == 0) s_ok = 1;
    own_push();
    asm volatile("bar.sync 1, 256;" ::: "memory");
    // the PUSH of the W1 sums is slice 1's job, not the writer's: the writer's release below would otherwise wait for the
    // acknowledgement of its own NVLink stores (5 us); in the scatter shape slice 1 also runs the owner's publishing pull
    const int pusher = gridDim.x > 1 ? 1 : 0;
    if (w1t && (int)blockIdx.x == pusher && oalive) {
      bool ok = true;
      px_push(px, pcan, q.dw1_sum[(g * H1 + k0 + kl) * 8 + wc]);
      if (pusher != 0 && px.scatter) (void)px_pull(px, pcan, &ok, true);
    }
    if (writer) {
      if (w1t) {
        bool ok = oalive;
        if (ok) gsum = px_pull(px, pcan, &ok, pusher == 0);
        if (ok) q.cgrads[pcan] = gsum; else s_ok = 0;
        __threadfence();
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (tid == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(want | (s_ok ? 0u : 1u)) : "memory");
    } else {
      if (tid == 0) {
        unsigned v;
        long long spins = 0;
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
          if ((v & ~1u) == want) break;
          if (++spins > (1ll << 26)) { atomicOr(px.status, 2); v = want | 1u; break; }
        } while (true);
        if (v & 1u) s_ok = 0;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (w1t && s_ok) gsum = __ldcg(q.cgrads + pcan);
    }
    xok = s_ok != 0;
  }
  if (w1t) {
    if (q.pending) {
      const int si = (g * H1 + k0 + kl) * 8 + wc;
      const float gs = PEERS ? gsum : q.dw1_sum[si];
      const bool ok = xok;
      if (ok) pw = wide_adam_math(cf, gs, pm, pv, pw);
      if (writer) {
        q.dw1_zero[si] = 0.f;
        if (ok) { q.cgrads[pcan] = gs; q.cparams[pcan] = pw; q.cm[pcan] = pm; q.cv[pcan] = pv; }
      }
    }
    if (writer) { q.w1_wr[(g * 3 + 0) * wblk + wi] = pw; q.w1_wr[(g * 3 + 1) * wblk + wi] = pm; q.w1_wr[(g * 3 + 2) * wblk + wi] = pv; }
    ws[(kl >> 2) * 28 + (wc < S ? (((kl & 3) >> 1) * S + wc) * 2 + (kl & 1) : 24 + (kl & 3))] = pw * TANH_K;
  }
  if (PEERS && q.pending) own_finish(cf);                 // this thread's element of the reduce: its peers' values have had the
                                                          // whole W1 exchange to arrive
  if (hi <= lo) return;
  asm volatile("bar.sync 1, 256;" ::: "memory");
  if (tid == 0) tc::tl_stamp(q.tl, q.tl_k, 2);            // W1 / b1 of this step in shared memory
  // ---- layer 1 of this step: 8 feature groups of 4 x 32 sample lanes
  const int64_t plane = a.cap * H1;
  const int kg = tid & 7, sl = tid >> 3;
  float w[24], b[4];
#pragma unroll
  for (int v = 0; v < 6; v++) *reinterpret_cast<float4*>(&w[4 * v]) = *reinterpret_cast<const float4*>(ws + kg * 28 + 4 * v);
  *reinterpret_cast<float4*>(b) = *reinterpret_cast<const float4*>(ws + kg * 28 + 24);
  for (int r0 = lo; r0 < hi; r0 += L1R_TILE) {
    if (r0 != lo) {
      asm volatile("bar.sync 1, 256;" ::: "memory");
      load_tile(r0);
      asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    const int n = min(L1R_TILE, hi - r0);
    const int nv = max(min(n, a.count - r0), 0);           // rows below hold samples, rows [nv, n) are the padding of the last tile
    __nv_bfloat16* oh = a.h1p + (g * 2 + 0) * plane + (int64_t)(r0 + sl) * H1 + k0 + kg * 4;   // walks 32 rows at a time
    __nv_bfloat16* om = oh + plane;
    const float* xr = &xs[sl][0];
    const int64_t ostep = (int64_t)32 * H1;
    int r = sl;
    for (; r < nv; r += 32, oh += ostep, om += ostep, xr += 32 * 2 * S) {
      float z[4];
      float4 xx[3];                                        // (x0, x0, x1, x1), (x2, x2, x3, x3), (x4, x4, x5, x5)
#pragma unroll
      for (int v = 0; v < 3; v++) xx[v] = *reinterpret_cast<const float4*>(xr + 4 * v);
      float2 z01 = make_float2(b[0], b[1]), z23 = make_float2(b[2], b[3]);
#pragma unroll
      for (int c = 0; c < S; c++) {                        // two features per FFMA2, the same fma chain per feature as before
        const float2 xc = (c & 1) ? make_float2(xx[c >> 1].z, xx[c >> 1].w) : make_float2(xx[c >> 1].x, xx[c >> 1].y);
        z01 = __ffma2_rn(make_float2(w[2 * c], w[2 * c + 1]), xc, z01);
        z23 = __ffma2_rn(make_float2(w[2 * (S + c)], w[2 * (S + c) + 1]), xc, z23);
      }
      tanh4(z01.x, z01.y, z23.x, z23.y, z[0], z[1], z[2], z[3]);
      uint2 th, tm;
      split2(z[0], z[1], th.x, tm.x);
      split2(z[2], z[3], th.y, tm.y);
      *reinterpret_cast<uint2*>(oh) = th;
      *reinterpret_cast<uint2*>(om) = tm;
    }
    for (; r < n; r += 32, oh += ostep, om += ostep) {     // padding rows: zeros (the sample-contraction GEMM needs no masking)
      *reinterpret_cast<uint2*>(oh) = make_uint2(0u, 0u);
      *reinterpret_cast<uint2*>(om) = make_uint2(0u, 0u);
    }
  }
  if (tid == 0) tc::tl_stamp(q.tl, q.tl_k, 3);            // layer-1 loop done
}

// layer 2 activation + layer 3 + loss head + D2 for both nets.  The GEMM left Z2 = H1 W2^T; here H2 = tanh(Z2 + b2).
// One warp per sample; lane l owns the NJ consecutive features j = l*NJ .. l*NJ+NJ-1, so the Z2 row is read and the D2
// planes are written as contiguous vectors.  The critic is finished first (its V feeds the actor's advantage, :274), so
// only one net's activations are live at a time; W3 / b2 sit in shared memory.  dW3, db2, db3 and the loss sums are
// accumulated per lane over all samples of the block and written once, to this block's partial row.
// A warp walks ~7 samples; the rows and scalars of its NEXT sample are fetched before the current one is worked on
// (one exposed memory round trip per warp instead of two per sample), which is what the 128-register budget of
// HEAD_CTAS = 2 blocks per SM pays for.
constexpr int HEAD_CTAS = 2;
template <int NJ>   // H2 / 32
__global__ void __launch_bounds__(256, HEAD_CTAS) k_wide_head(const WideArgs a) {
  extern __shared__ __align__(16) float hsm[];
  constexpr int H2 = NJ * 32;
  float* wsm = hsm;                  // [2 nets][2][H2]: W3, b2
  float* red = hsm + 4 * H2;         // [8 warps][2 nets][2 * H2 + 4]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const NetDims d = a.d;
  pdl_launch_dependents();
  pdl_wait();
  for (int i = threadIdx.x; i < 4 * H2; i += blockDim.x) {
    const int g = i / (2 * H2), r = i % (2 * H2);
    wsm[i] = a.params[g * d.SIZE + (r < H2 ? d.oW3 + r : d.ob2 + (r - H2))];
  }
  const float b3a = a.params[d.ob3], b3c = a.params[d.SIZE + d.ob3];
  __syncthreads();
  float gW3[2][NJ], gb2[2][NJ], gb3[2] = {0.f, 0.f};
#pragma unroll
  for (int g = 0; g < 2; g++)
#pragma unroll
    for (int i = 0; i < NJ; i++) { gW3[g][i] = 0.f; gb2[g][i] = 0.f; }
  float loss1 = 0.f, loss2 = 0.f, clipped = 0.f;
  const int64_t plane = a.cap * H2;

  // what a warp needs of one sample: both Z2 rows (this lane's NJ features) and the three trajectory scalars
  struct Fetched { float2 z[2][NJ / 2]; float R, act, oldlp; };
  auto fetch = [&](int s, Fetched& q) {
    q.R = 0.f; q.act = 0.f; q.oldlp = 0.f;
#pragma unroll
    for (int g = 0; g < 2; g++)
#pragma unroll
      for (int i = 0; i < NJ / 2; i++) q.z[g][i] = make_float2(0.f, 0.f);
    if (s >= a.count) return;                    // padding rows of the last tile: all-zero activations and deltas
    const int64_t p = sample_pos(a, s);
    q.R = a.tr.ret[p]; q.act = a.tr.action[p]; q.oldlp = a.tr.logp[p];
#pragma unroll
    for (int g = 0; g < 2; g++) {
      const float2* zp = reinterpret_cast<const float2*>(a.h2 + ((int64_t)g * a.cap + s) * H2 + lane * NJ);
#pragma unroll
      for (int i = 0; i < NJ / 2; i++) q.z[g][i] = zp[i];
    }
  };
  // one net of one sample: activation, dot with W3 (returned warp-reduced), h left in registers
  auto forward = [&](int g, const float2 (&z)[NJ / 2], bool valid, float (&h)[NJ]) {
    const float* w3 = wsm + g * 2 * H2 + lane * NJ;
    const float* b2 = w3 + H2;
    float y = 0.f;
#pragma unroll
    for (int i = 0; i < NJ; i += 2) {
      float2 t = make_float2(0.f, 0.f);
      if (valid) {
        t.x = act_tanh(z[i >> 1].x + b2[i]);
        t.y = act_tanh(z[i >> 1].y + b2[i + 1]);
      }
      h[i] = t.x; h[i + 1] = t.y;
      y = fmaf(w3[i], t.x, y);
      y = fmaf(w3[i + 1], t.y, y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) y += __shfl_xor_sync(0xffffffffu, y, o);
    return y;
  };
  // D2 = dz (x) W3 .* (1 - h^2) -> planes; column sums
  auto backward = [&](int g, int s, float dz, const float (&h)[NJ]) {
    const float* w3 = wsm + g * 2 * H2 + lane * NJ;
    __align__(16) __nv_bfloat16 dh[NJ], dm[NJ];
#pragma unroll
    for (int i = 0; i < NJ; i++) {
      const float dd = dz * w3[i] * (1.0f - h[i] * h[i]);
      gW3[g][i] = fmaf(dz, h[i], gW3[g][i]);
      gb2[g][i] += dd;
      tc::split_bf16(dd, dh[i], dm[i]);
    }
    const int64_t o = (int64_t)s * H2 + lane * NJ;
    uint32_t* oh = reinterpret_cast<uint32_t*>(a.d2p + (g * 2 + 0) * plane + o);
    uint32_t* om = reinterpret_cast<uint32_t*>(a.d2p + (g * 2 + 1) * plane + o);
    if (NJ == 8) {
      *reinterpret_cast<uint4*>(oh) = *reinterpret_cast<const uint4*>(dh);
      *reinterpret_cast<uint4*>(om) = *reinterpret_cast<const uint4*>(dm);
    } else {
#pragma unroll
      for (int i = 0; i < NJ / 2; i++) {
        oh[i] = reinterpret_cast<const uint32_t*>(dh)[i];
        om[i] = reinterpret_cast<const uint32_t*>(dm)[i];
      }
    }
  };

  const int sstride = (int)gridDim.x * 8;
  Fetched cur, nxt;
  if ((int)blockIdx.x * 8 + warp < a.m_pad) fetch(blockIdx.x * 8 + warp, cur);
  for (int s = blockIdx.x * 8 + warp; s < a.m_pad; s += sstride, cur = nxt) {
    if (s + sstride < a.m_pad) fetch(s + sstride, nxt);   // in flight while this sample is worked on
    const bool valid = s < a.count;
    const float R = cur.R, act = cur.act, oldlp = cur.oldlp;
    float h[NJ];
    // ---- critic: no final activation (ppo_nn.h:118-122)
    const float V = forward(1, cur.z[1], valid, h) + b3c;
    const float e = V - R;
    const float dzc = valid ? e * a.inv_denom : 0.f;                 // d(0.5 mean e^2)/dV  (:288)
    backward(1, s, dzc, h);
    // ---- actor
    const float mu = act_tanh(forward(0, cur.z[0], valid, h) + b3a);   // ppo_nn.h:102
    float dza = 0.f;
    if (valid) {
      const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);  // :140
      const float ratio = expf(lp - oldlp);                          // :263
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
      dza = dmu * (1.0f - mu * mu);
      if (lane == 0) {
        loss2 = fmaf(e, e, loss2);
        loss1 -= fminf(surr1, surr2);                                // :287
        clipped += inside ? 0.f : 1.f;
        gb3[0] += dza;
        gb3[1] += dzc;
      }
    }
    backward(0, s, dza, h);
  }

  // block reduction of the per-lane column sums -> this block's partial row (fixed order: deterministic)
  const int RW = 2 * H2 + 4;
#pragma unroll
  for (int g = 0; g < 2; g++) {
    float* r = red + (warp * 2 + g) * RW;
#pragma unroll
    for (int i = 0; i < NJ; i++) { r[lane * NJ + i] = gW3[g][i]; r[H2 + lane * NJ + i] = gb2[g][i]; }
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

// (dW1 / db1: the work of the former k_wide_dw1 kernel rides in extra warps of the dW2 GEMM, wide_gemm.cu dw1_worker)

// Blocks [0, w2_blocks): W2 -- with the L2 having added the split-K partials already (ADDED), two elements per thread;
// the rest: the small parameters and the loss sums, coalesced column sums (wide_reduce.cuh wide_reduce_small).
// 834 blocks for 2 x 256 nets: one wave (the former warp-per-parameter layout needed 2,178 blocks = two waves at 56
// registers, and its strided 4-byte reads cost 14 of the kernel's 19 us).
template <bool ADDED>
__global__ void __launch_bounds__(RED_THREADS) k_wide_reduce(const ReduceArgs r, int w2_blocks) {
  __shared__ float red[RED_RG * RED_COLS];
  pdl_launch_dependents();
  pdl_wait();
  const int vb = (int)blockIdx.x;
  if (vb >= w2_blocks) {
    wide_reduce_small(r, vb - w2_blocks, red);
    return;
  }
  if (ADDED) {
    const int w2n = r.d.H2 * r.d.H1;
    const int i = (vb * RED_THREADS + (int)threadIdx.x) * 2;
    if (i >= 2 * w2n) return;
    float2* two = reinterpret_cast<float2*>(r.part_dw2 + i);   // [net][H2 * H1], already summed; zeroed for the next step's adds
    const float2 s = *two;
    *two = make_float2(0.f, 0.f);
    auto pidx = [&](int j) { return (j / w2n) * r.d.SIZE + r.d.oW2 + (j % w2n); };
    wide_reduce_finish2(r, pidx(i), pidx(i + 1), s.x, s.y);
  } else {
    wide_reduce_finish(r, wide_reduce_body<false>(r, w2_blocks, vb, true));
  }
}

__global__ void __launch_bounds__(128) k_wide_adam(const UpdateArgs a, const NetDims d, int pstride, __nv_bfloat16* w2p,
                                                   float* bias2, const AdamCoef* coef_p) {
  pdl_launch_dependents();
  pdl_wait();                          // (a no-op unless launched with the programmatic-serialization attribute)
  const AdamCoef coef = *coef_p;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  // (after an ncclAllReduce of a.grads: the default multi-GPU path)
  if (p < d.P) wide_store_planes(d, w2p, bias2, p, wide_adam_apply(a, coef, p, a.grads[p]));
  else if (p < d.P + 3) wide_loss_log(a, p - d.P, a.grads[pstride - 4 + (p - d.P)]);
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
  __nv_bfloat16 *w2p = nullptr, *h1p = nullptr, *d2p = nullptr;
  float *bias2 = nullptr, *h2 = nullptr, *d1 = nullptr;
  float* vtmp = nullptr;           // [4][cap] V of the critic pass of the fused forward + head kernel
  bool fused_fh = false;           // k_wide_fh instead of the forward GEMM + k_wide_head (B2_WIDE_FUSED=0: the two kernels)
  int fh_grid = 0;                 // CTAs of the last k_wide_fh launch (= partial rows the reduce sums)
  long long* fh_dbg = nullptr;     // B2_FH_DBG=1: clock64 timeline of the head warps, printed once (diagnostic)
  CUtensorMap tm_d2_st;            // d2p planes as a store target (16 x 32 x 2-plane boxes, 32-B swizzle)
  CUtensorMap tm_out_dw2_16;       // the dW2 planes with 16-column boxes (k_gemm_mn2's epilogue)
  CUtensorMap tm_out_d1_16;        // d1 (fp32) as a store target with 16-column boxes (k_wide_fh's dX passes)
  bool fused_dx = false;           // the dX GEMM inside k_wide_fh (B2_WIDE_FUSE_DX=0: its own launch, k_gemm2_ws<true>)
  AdamCoef* coef = nullptr;
  float *part_head = nullptr, *part_dw1 = nullptr, *part_dw2 = nullptr;
  int head_blocks = 0, head_row = 0, dw1_blocks = 0, splits_max = 0;
  int* bar_counter = nullptr;   // grid barrier of the dW2 kernel's fused reduce tail (device, monotonically increasing)
  int bar_count = 0;            // host mirror: value after every launch so far
  bool fuse_reduce = false;     // B2_WIDE_FUSE_REDUCE=1: reduce + Adam in the tail of the dW2 GEMM kernel behind a grid barrier (needs
                                // dw2_add).  Opt-in: measured 38.3 vs 36.5 ms per configs[4] update -- the barrier and the 2-3 serial
                                // rounds of latency-bound partial sums per CTA cost more than the launch they save
  bool dw2_add = true;   // split-K of dW2 through TMA reduce-add (default); B2_WIDE_DETERMINISTIC=1: partial planes, fixed-order sum
  CUtensorMap tm_h1, tm_w2, tm_d2, tm_w2t, tm_d2_mn, tm_h1_mn;
  CUtensorMap tm_out_h2, tm_out_d1, tm_out_dw2;
  CUtensorMap tm_w2_half;          // K-major weights in (H2/2)-row boxes: the resident halves of the CTA-pair GEMM
  bool pair_fwd = false, pair_bwd = false;
  int32_t* boot_idx = nullptr;
  int64_t boot_cap = 0;
  int* boot_count = nullptr;
  // the merged sequence (k_wide_l1r: reduce + Adam of step s-1 inside the layer-1 kernel of step s; B2_WIDE_MERGE=0: off)
  bool merge_env = true;
  bool merge_peers = false;        // B2_WIDE_MERGE_PEERS=1: the merged sequence on TWO GPUs too (k_wide_l1r<PEERS>).  Opt-in: measured
                                   // 27.7 vs 28.25 ms per configs[4] update on 2 x B200 -- the exchange inside the layer-1 kernel costs
                                   // 15 us per step against 10 in k_wide_reduce, which eats most of the launch it saves -- and at
                                   // world = 8 (reduce-scatter shape) it FAILED bench.py's multi_gpu_check (ranks ended with different
                                   // weights; 29.6 vs 30.6 ms) and was not debugged for lack of GPU time: never taken at world > 2
  float* dw1_sum = nullptr;        // [2 phases][2][H1 * 8]
  float* head_sum = nullptr;       // [2][head_row]
  float* w1s = nullptr;            // [2 phases][2 nets][3][H1][S + 1]
  unsigned* w1_flag = nullptr;     // [2 nets][8]: k_wide_l1r<PEERS>'s writer-slice flags
  int phase = 0;                   // flips with every k_wide_l1r launch
  bool pending = false;            // a step's sums wait for the next k_wide_l1r launch (or wide_flush)
  bool seeded = false;             // w1s[phase ^ 1] holds the current W1 state
  ReduceArgs pend_r{};             // of the pending step
  bool l2_tried = false;           // the access-policy window of wide_step (h1p in the L2's persisting set-aside) has been set up
  long long* tl = nullptr;         // B2_WIDE_TL=1: %globaltimer timeline of one step's kernels [4 kernels][8 slots][min, max]
  int tl_state = 0, tl_calls = 0;  // 0 idle, 1 armed: this step's kernels stamp rows 0-2, 2: the next layer-1 launch stamps row 3
};

int wide_pstride(int H1, int H2) {
  const NetDims d = NetDims::make(S, H1, H2);
  return ((d.P + 4 + 31) / 32) * 32;
}

int64_t wide_chunk_cap(const WideNet* w) { return w ? w->cap : 0; }

void wide_destroy(WideNet* w) {
  if (!w) return;
  dfree(w->w2p); dfree(w->h1p); dfree(w->d2p);
  dfree(w->bias2); dfree(w->h2); dfree(w->d1); dfree(w->coef); dfree(w->vtmp); dfree(w->fh_dbg);
  dfree(w->part_head); dfree(w->part_dw1); dfree(w->part_dw2); dfree(w->bar_counter);
  dfree(w->boot_idx); dfree(w->boot_count);
  dfree(w->dw1_sum); dfree(w->head_sum); dfree(w->w1s); dfree(w->tl); dfree(w->w1_flag);
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
  WCK(dmalloc(&w->w2p, bf * 4 * H2 * H1));
  WCK(dmalloc(&w->h1p, bf * 4 * cap * H1));
  WCK(dmalloc(&w->d2p, bf * 4 * cap * H2));
  WCK(dmalloc(&w->bias2, sizeof(float) * 4 * H2));
  WCK(dmalloc(&w->vtmp, sizeof(float) * 4 * cap));
  if (getenv("B2_FH_DBG")) { WCK(dmalloc(&w->fh_dbg, sizeof(long long) * 16 * 160)); WCK(cudaMemset(w->fh_dbg, 0, sizeof(long long) * 16 * 160)); }
  WCK(dmalloc(&w->coef, 2 * sizeof(AdamCoef)));   // [phase]: the merged sequence reads step s-1's while it writes step s's
  WCK(dmalloc(&w->h2, sizeof(float) * 2 * cap * H2));
  WCK(dmalloc(&w->d1, sizeof(float) * 2 * cap * H1));
  // stale bits must never be NaN where a zero-weighted product could meet them
  WCK(cudaMemset(w->h1p, 0, bf * 4 * cap * H1));
  WCK(cudaMemset(w->d2p, 0, bf * 4 * cap * H2));
  w->head_blocks = HEAD_CTAS * sm_count;
  w->head_row = 2 * H2 + 4;
  w->dw1_blocks = sm_count;
  w->splits_max = sm_count;
  WCK(dmalloc(&w->part_head, sizeof(float) * (size_t)w->head_blocks * 2 * w->head_row));
  WCK(dmalloc(&w->part_dw1, sizeof(float) * (size_t)w->dw1_blocks * 2 * H1 * 8));
  WCK(dmalloc(&w->part_dw2, sizeof(float) * (size_t)2 * w->splits_max * H2 * H1));
  WCK(cudaMemset(w->part_dw2, 0, sizeof(float) * (size_t)2 * w->splits_max * H2 * H1));
  { const char* e = getenv("B2_WIDE_DETERMINISTIC"); w->dw2_add = !(e && e[0] == '1'); }
  w->fuse_reduce = w->dw2_add && getenv("B2_WIDE_FUSE_REDUCE") != nullptr;
  WCK(dmalloc(&w->bar_counter, sizeof(int)));
  WCK(cudaMemset(w->bar_counter, 0, sizeof(int)));
  WCK(dmalloc(&w->boot_count, sizeof(int)));
  { const char* e = getenv("B2_WIDE_MERGE"); w->merge_env = !(e && e[0] == '0'); }
  { const char* e = getenv("B2_WIDE_MERGE_PEERS"); w->merge_peers = e && e[0] == '1'; }
  WCK(dmalloc(&w->dw1_sum, sizeof(float) * 2 * 2 * H1 * 8));
  WCK(dmalloc(&w->head_sum, sizeof(float) * 2 * w->head_row));
  WCK(dmalloc(&w->w1s, sizeof(float) * 2 * 2 * 3 * H1 * (S + 1)));
  WCK(cudaMemset(w->dw1_sum, 0, sizeof(float) * 2 * 2 * H1 * 8));
  WCK(cudaMemset(w->head_sum, 0, sizeof(float) * 2 * w->head_row));
  WCK(cudaMemset(w->w1s, 0, sizeof(float) * 2 * 2 * 3 * H1 * (S + 1)));
  WCK(dmalloc(&w->w1_flag, sizeof(unsigned) * 16));
  WCK(cudaMemset(w->w1_flag, 0, sizeof(unsigned) * 16));
#ifdef B2_TIMELINE
  if (getenv("B2_WIDE_TL")) WCK(dmalloc(&w->tl, sizeof(long long) * 4 * 8 * 2));
#endif
#undef WCK
  int rc = 0;
  const uint64_t ucap = (uint64_t)cap;
  // sample-major activations: K = features, rows = samples ; weights: K = in-features, rows = out-features
  if (!rc) rc = make_operand_map(&w->tm_h1, w->h1p, H1, ucap, 4, H1, ucap * H1, 128);
  if (!rc) rc = make_operand_map(&w->tm_w2, w->w2p, H1, H2, 4, H1, (uint64_t)H2 * H1, (uint32_t)H2);
  if (!rc) rc = make_operand_map(&w->tm_d2, w->d2p, H2, ucap, 4, H2, ucap * H2, 128);
  if (!rc) rc = make_operand_map(&w->tm_w2t, w->w2p, H1, H2, 4, H1, (uint64_t)H2 * H1, 64);   // W2 as [K = j][N = k] boxes
  // the same sample-major planes read contraction-major (K = samples) by the dW2 GEMM: boxes of 64 features x 32 samples
  if (!rc) rc = make_operand_map(&w->tm_d2_mn, w->d2p, H2, ucap, 4, H2, ucap * H2, MN_BOX_ROWS);
  if (!rc) rc = make_operand_map(&w->tm_h1_mn, w->h1p, H1, ucap, 4, H1, ucap * H1, MN_BOX_ROWS);
  w->pair_fwd = gemm2_supported(H2, H1, false);
  w->pair_bwd = gemm2_supported(H1, H2, true);
  if (!rc && w->pair_fwd) rc = make_operand_map(&w->tm_w2_half, w->w2p, H1, H2, 4, H1, (uint64_t)H2 * H1, (uint32_t)(H2 / 2));
  { const char* e = getenv("B2_WIDE_FUSED"); w->fused_fh = w->pair_fwd && fh_supported(H2, H1) && !(e && e[0] == '0'); }
  if (!rc && w->fused_fh) rc = make_plane_store_map(&w->tm_d2_st, w->d2p, H2, ucap, 4);
  { const char* e = getenv("B2_WIDE_FUSE_DX"); w->fused_dx = w->fused_fh && w->pair_bwd && !(e && e[0] == '0'); }
  if (!rc && w->fused_dx) rc = make_output_map16(&w->tm_out_d1_16, w->d1, H1, ucap, 2);
  if (!rc) rc = make_output_map(&w->tm_out_h2, w->h2, H2, ucap, 2);
  if (!rc) rc = make_output_map(&w->tm_out_d1, w->d1, H1, ucap, 2);
  if (!rc) rc = make_output_map(&w->tm_out_dw2, w->part_dw2, H1, H2, (uint64_t)2 * w->splits_max);
  if (!rc) rc = make_output_map16(&w->tm_out_dw2_16, w->part_dw2, H1, H2, (uint64_t)2 * w->splits_max);
  if (rc) { wide_destroy(w); return rc; }
  *out = w;
  return B2_OK;
}

template <int NJ>
static cudaError_t launch_head(const WideArgs& wa, int grid, cudaStream_t st) {
  constexpr int H2 = NJ * 32;
  const int smem = (4 * H2 + 8 * 2 * (2 * H2 + 4)) * 4;
  static b2::PerDeviceOnce once;   // function attributes are per device
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(k_wide_head<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch_pdl(k_wide_head<NJ>, dim3(grid), dim3(256), (size_t)smem, st, wa);
}

// one k_wide_l1r launch: layer 1 of the chunk wa describes (m_pad == 0: none) + whatever step is pending
static cudaError_t launch_l1r(WideNet* w, const WideArgs& wa, const UpdateArgs& u, int skip, cudaStream_t st) {
  const NetDims& d = w->d;
  const int H1 = d.H1, n_fs = H1 / L1R_FS;
  L1rArgs q{};
  q.a = wa;
  q.a.coef_out = w->coef + w->phase;
  q.pending = (w->pending && !(skip & 32)) ? 1 : 0;
  q.seed = w->seeded ? 0 : 1;
  q.r = w->pend_r;
  q.dw1_sum = w->dw1_sum + (size_t)(w->phase ^ 1) * 2 * H1 * 8;
  q.dw1_zero = w->dw1_sum + (size_t)w->phase * 2 * H1 * 8;
  q.head_sum = w->head_sum;
  q.w1_rd = w->w1s + (size_t)(w->phase ^ 1) * 2 * 3 * H1 * (S + 1);
  q.w1_wr = w->w1s + (size_t)w->phase * 2 * 3 * H1 * (S + 1);
  q.cparams = u.params; q.cm = u.adam_m; q.cv = u.adam_v; q.cgrads = u.grads;
  q.w1_flag = w->w1_flag;
  if (w->tl_state == 1) { q.tl = w->tl; q.tl_k = 0; }
  else if (w->tl_state == 2) { q.tl = w->tl; q.tl_k = 3; w->tl_state = 3; }
  const int full = std::max(1, w->sm_count * L1_CTAS / (2 * n_fs));
  const int n_rs = wa.m_pad > 0 ? std::max(1, std::min(full, wa.m_pad / 16)) : full;
  const bool peers = (q.pending ? q.r.u.px.world : u.px.world) > 1;
  cudaError_t ce = peers ? launch_pdl(k_wide_l1r<true>, dim3(n_rs, n_fs, 2), dim3(L1R_THREADS), 0, st, q)
                         : launch_pdl(k_wide_l1r<false>, dim3(n_rs, n_fs, 2), dim3(L1R_THREADS), 0, st, q);
  if (ce != cudaSuccess) return ce;
  w->pending = false;
  w->seeded = true;
  w->phase ^= 1;
  return cudaSuccess;
}

int wide_flush(WideNet* w, cudaStream_t st, int64_t* launches) {
  if (!w || !w->pending) return B2_OK;
  WideArgs wa{};
  wa.d = w->d; wa.params = w->pend_r.u.params; wa.tr = w->pend_r.u.tr; wa.cap = w->cap; wa.hp = w->pend_r.u.hp;
  wa.h1p = w->h1p;
  cudaError_t ce = launch_l1r(w, wa, w->pend_r.u, 0, st);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "wide flush: %s", cudaGetErrorString(ce));
  (*launches)++;
  return B2_OK;
}

int wide_step(WideNet* w, const UpdateArgs& a, bool apply_adam, bool more_follow, cudaStream_t st, int64_t* launches) {
  const NetDims& d = w->d;
  const int H1 = d.H1, H2 = d.H2;
  WideArgs wa{};
  wa.d = d; wa.params = a.params; wa.tr = a.tr; wa.idx = a.idx; wa.cap = w->cap;
  wa.inv_denom = 1.0f / (float)a.denom; wa.hp = a.hp;
  wa.w2p = w->w2p; wa.h1p = w->h1p; wa.d2p = w->d2p;
  wa.bias2 = w->bias2; wa.h2 = w->h2; wa.d1 = w->d1;
  wa.part_head = w->part_head; wa.part_dw1 = w->part_dw1; wa.head_row = w->head_row;
  cudaError_t ce;
#define LCK(what)                                                                              \
  do {                                                                                         \
    if ((ce = cudaGetLastError()) != cudaSuccess)                                              \
      return set_error(B2_ERR_CUDA, "wide %s: %s", what, cudaGetErrorString(ce));              \
    (*launches)++;                                                                             \
  } while (0)
  if (!w->l2_tried) {
    // The step's working set (h1p, d2p, d1: 33.5 MB each at 2x256 / 16,384 samples) does not quite stay in the 126-MB L2
    // (k_wide_fh reads 34 MB from DRAM per launch), so h1p -- written once and read by both GEMM kernels of the step -- is put
    // into the L2's persisting set-aside through an access-policy window on the engine's stream: 22.7 ms per configs[4] update
    // and flat over back-to-back updates, against 22.6 drifting to 23.2 without it (less DRAM traffic, less power, steadier
    // clock); d2p or d1 instead: 24.1.  Only when the array fits the set-aside whole; the device-wide limit
    // (cudaLimitPersistingL2CacheSize) is only ever raised.  B2_WIDE_L2_PERSIST=0: off; =d2 / d1 (diagnostic): that array.
    w->l2_tried = true;
    const char* e = getenv("B2_WIDE_L2_PERSIST");
    if (!e) e = "h1";
    void* base = nullptr; size_t bytes = 0;
    if (e[0] == 'h') { base = w->h1p; bytes = sizeof(__nv_bfloat16) * 4 * (size_t)w->cap * H1; }
    else if (e[0] == 'd' && e[1] == '2') { base = w->d2p; bytes = sizeof(__nv_bfloat16) * 4 * (size_t)w->cap * H2; }
    else if (e[0] == 'd' && e[1] == '1') { base = w->d1; bytes = sizeof(float) * 2 * (size_t)w->cap * H1; }
    int dev = 0, max_persist = 0, max_win = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
    cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    if (base && bytes >= ((size_t)8 << 20) && bytes <= (size_t)max_persist && bytes <= (size_t)max_win) {
      size_t cur = 0;
      cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
      cudaError_t e1 = cur >= bytes ? cudaSuccess : cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes);
      cudaStreamAttrValue v{};
      v.accessPolicyWindow.base_ptr = base;
      v.accessPolicyWindow.num_bytes = bytes;
      v.accessPolicyWindow.hitRatio = 1.0f;
      v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      cudaError_t e2 = e1 == cudaSuccess ? cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) : e1;
      if (e2 != cudaSuccess) (void)cudaGetLastError();   // a caching hint: never an error of the step
    }
  }
  static const bool pair_env0 = [] { const char* e = getenv("B2_WIDE_DW2_PAIR"); return !(e && e[0] == '0'); }();
  const int64_t n_chunks = (a.count + w->cap - 1) / w->cap;
  // the merged sequence: one GPU, Adam applied here, one chunk, sums in the L2 (not the fixed-order mode), the fused forward +
  // head kernel and the CTA-pair dW2 kernel (the two producers that can leave sums)
  const bool merge = w->merge_env && apply_adam && (a.px.world <= 1 || (w->merge_peers && a.px.world == 2)) && n_chunks == 1 && w->dw2_add && !w->fuse_reduce && w->fused_fh &&
                     pair_env0 && gemm_mn2_supported(H2, H1) && getenv("B2_WIDE_FUSE_M") == nullptr;
  if (a.step_slot == 0 || !merge) {
    if (w->pending && !merge) { int rc = wide_flush(w, st, launches); if (rc) return rc; }
    w->pending = false;       // (a step left pending by an update that failed part-way is dropped with it)
    w->seeded = false;        // the canonical vector is the truth at the start of an update
  }
  if (a.step_slot == 0) {   // first optimizer step of this update: the weights may have been set from outside
    if (merge) {
      if ((ce = cudaMemsetAsync(w->dw1_sum, 0, sizeof(float) * 2 * 2 * H1 * 8, st)) != cudaSuccess ||
          (ce = cudaMemsetAsync(w->head_sum, 0, sizeof(float) * 2 * w->head_row, st)) != cudaSuccess)
        return set_error(B2_ERR_CUDA, "wide memset: %s", cudaGetErrorString(ce));
    }
    k_wide_prep<<<128, 256, 0, st>>>(wa);
    LCK("prep");
  }
  // diagnostic only (profiles/wide_marginal.sh): B2_WIDE_SKIP = bit mask of step kernels to leave out (1 l1, 2 fwd GEMM,
  // 4 head, 8 dX GEMM, 16 dW2 GEMM + dw1, 32 reduce) -- results are garbage, the timing shows each kernel's marginal cost
  static const int skip = [] { const char* e = getenv("B2_WIDE_SKIP"); return e ? atoi(e) : 0; }();
  // B2_WIDE_TL=1 (diagnostic): the 200th merged step of the process stamps its three kernels, the next step's layer-1 launch
  // the fourth row; printed once, relative to the first stamp (profiles/r2/r2k_wide_timeline.txt)
  if (w->tl && merge) {
    if (w->tl_state == 0 && ++w->tl_calls == 200) {
      std::vector<long long> init(4 * 8 * 2);
      for (size_t i = 0; i < init.size(); i += 2) { init[i] = -1; init[i + 1] = 0; }   // (min as unsigned, max)
      cudaStreamSynchronize(st);
      cudaMemcpy(w->tl, init.data(), sizeof(long long) * init.size(), cudaMemcpyHostToDevice);
      w->tl_state = 1;
    }
  }
  for (int64_t c = 0; c < n_chunks; c++) {
    wa.first = a.first + c * w->cap;
    wa.count = (int)std::min<int64_t>(w->cap, a.count - c * w->cap);
    wa.m_pad = (wa.count + 127) / 128 * 128;
    wa.bump_step = c == 0 ? a.adam_step : nullptr;
    wa.coef_out = w->coef;
    const int m_tiles = wa.m_pad / 128;
    AdamCoef* step_coef = merge ? w->coef + w->phase : w->coef;   // where this step's Adam coefficients are written
    const int sum_phase = w->phase;                               // (launch_l1r flips it)
    // layer 1 (merged: + the reduce and Adam of the step before)
    const int l1_grid = std::max(1, std::min(w->sm_count * L1_CTAS / 2, wa.m_pad / 16));
    if (!(skip & 1))
    if ((ce = merge ? launch_l1r(w, wa, a, skip, st) : launch_pdl(k_wide_l1, dim3(l1_grid, 2), dim3(256), 0, st, wa)) != cudaSuccess)
      return set_error(B2_ERR_CUDA, "wide l1: %s", cudaGetErrorString(ce));
    (*launches)++;
    // H2 = tanh(H1 W2^T + b2)
    GemmArgs g{};
    g.G = 2; g.M = wa.count; g.N = H2; g.K = H1; g.m_tiles = m_tiles; g.splits = 1; g.kb_per_split = H1 / 64;
    g.stages = gemm_x3_stages(H2);
    int head_rows = w->head_blocks;                       // partial rows of the head the reduce kernel sums
    if (w->fused_fh) {                                    // forward GEMM + head in one kernel (wide_fused.cu)
      FhArgs f{};
      f.d = d; f.N = H2; f.K = H1; f.stages = fh_stages(H2, H1); f.count = wa.count; f.m_pad = wa.m_pad; f.cap = w->cap;
      f.first = wa.first; f.idx = a.idx; f.tr = a.tr; f.params = a.params; f.headw = w->bias2; f.vtmp = w->vtmp;
      f.inv_denom = wa.inv_denom; f.hp = a.hp; f.part_head = w->part_head; f.head_row = w->head_row;
      f.dbg = w->fh_dbg;
      f.head_sum = merge ? w->head_sum : nullptr;
      f.tl = w->tl_state == 1 ? w->tl : nullptr;
      f.dx = (w->fused_dx && !(skip & 8)) ? 1 : 0;
      if (!(skip & 6))
      if ((ce = launch_wide_fh(w->tm_h1, w->tm_w2_half, w->tm_d2_st, w->tm_d2, w->tm_w2t, w->tm_out_d1_16, f, w->sm_count, &w->fh_grid, st)) != cudaSuccess)
        return set_error(B2_ERR_CUDA, "wide fwd+head: %s", cudaGetErrorString(ce));
      (*launches)++;
      head_rows = w->fh_grid;
      static int dbg_calls = 0;
      if (w->fh_dbg && ++dbg_calls == 50) {               // one warm launch in the middle of an update
        cudaStreamSynchronize(st);
        std::vector<long long> h(16 * 160);
        cudaMemcpy(h.data(), w->fh_dbg, sizeof(long long) * h.size(), cudaMemcpyDeviceToHost);
        for (int c : {0, 1, 64, 127}) {
          if (c >= w->fh_grid) continue;
          const long long* r = &h[(size_t)c * 16];
          fprintf(stderr, "fh cta %3d:", c);
          for (int it = 0; it < 2; it++)
            fprintf(stderr, "  tile %d wait %6lld acc_ready %6lld pass1_done %6lld pass2_done %6lld |", it, r[it * 4] - r[12], r[it * 4 + 1] - r[12],
                    r[it * 4 + 2] - r[12], r[it * 4 + 3] - r[12]);
          fprintf(stderr, "  head_done %6lld (cycles after entry)\n", r[13] - r[12]);
        }
      }
    } else {
    if (skip & 2) ce = cudaSuccess;
    else if (w->pair_fwd) {   // CTA pairs with the weight halves resident (wide_gemm2.cu)
      const Gemm2Args g2{2, wa.count, H2, H1, gemm2_stages(H2, H1)};
      ce = launch_gemm2(false, w->tm_h1, w->tm_w2_half, w->tm_out_h2, g2, w->sm_count, st);
    } else {
      ce = launch_gemm_x3(w->tm_h1, w->tm_w2, w->tm_out_h2, g, w->sm_count, st);
    }
    if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "wide gemm fwd: %s", cudaGetErrorString(ce));
    (*launches)++;
    // loss head, D2
    {
      const int n_tiles = wa.m_pad / 8;
      const int grid = n_tiles < w->head_blocks ? n_tiles : w->head_blocks;
      if (grid < w->head_blocks)   // rows of blocks that do not run this time must read as zero
        if ((ce = cudaMemsetAsync(w->part_head, 0, sizeof(float) * (size_t)w->head_blocks * 2 * w->head_row, st)) != cudaSuccess)
          return set_error(B2_ERR_CUDA, "wide memset: %s", cudaGetErrorString(ce));
      if (skip & 4) ce = cudaSuccess;
      else switch (H2 / 32) {
        case 2: ce = launch_head<2>(wa, grid, st); break;
        case 4: ce = launch_head<4>(wa, grid, st); break;
        case 6: ce = launch_head<6>(wa, grid, st); break;
        default: ce = launch_head<8>(wa, grid, st); break;
      }
      if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "wide head: %s", cudaGetErrorString(ce));
      (*launches)++;
    }
    }
    // D1 = (D2 W2) .* (1 - H1^2)
    g = GemmArgs{};
    g.G = 2; g.M = wa.count; g.N = H1; g.K = H2; g.m_tiles = m_tiles; g.splits = 1; g.kb_per_split = H2 / 64;
    g.stages = gemm_x3_stages(H1); g.b_mn = 1;   // B = W2 as stored ([j][k]), read as an MN-major operand
    const bool dx_in_fh = w->fused_fh && w->fused_dx && !(skip & 6);     // passes 2, 3 of k_wide_fh
    if ((skip & 8) || dx_in_fh) ce = cudaSuccess;
    else if (w->pair_bwd) {
      const Gemm2Args g2{2, wa.count, H1, H2, gemm2_stages(H1, H2)};
      ce = launch_gemm2(true, w->tm_d2, w->tm_w2t, w->tm_out_d1, g2, w->sm_count, st);
    } else {
      ce = launch_gemm_x3(w->tm_d2, w->tm_w2t, w->tm_out_d1, g, w->sm_count, st);
    }
    if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "wide gemm bwd: %s", cudaGetErrorString(ce));
    if (!dx_in_fh) (*launches)++;
    // dW2 partials = D2^T H1 over sample splits
    const int kb_total = wa.m_pad / MN_BOX_ROWS;
    const int jt = (H2 + 127) / 128;
    // H2 = 256: one CTA owns both 128-row halves of dW2 (two TMEM accumulators), so the H1 block it streams feeds both
    // -- the kernel is L2-bandwidth bound (100 MB of operand reads per step with separate halves, 67 MB fused)
    const bool fuse_m = jt == 2 && gemm_x3_stages(H1, true, true) >= 2 && getenv("B2_WIDE_FUSE_M") != nullptr;   // opt-in: measured equal to two CTAs per net half (36.3 ms), see DESIGN.md s3b
    int splits = w->sm_count / (2 * (fuse_m ? 1 : jt));
    if (splits > kb_total) splits = kb_total;
    if (splits > w->splits_max) splits = w->splits_max;
    if (splits < 1) splits = 1;
    const int kbps = (kb_total + splits - 1) / splits;
    splits = (kb_total + kbps - 1) / kbps;
    g = GemmArgs{};
    g.G = 2; g.M = H2; g.N = H1; g.K = wa.m_pad; g.m_tiles = jt; g.splits = splits; g.kb_per_split = kbps;
    g.stages = gemm_x3_stages(H1, true, fuse_m); g.mn_major = 1; g.reduce_add = w->dw2_add ? 1 : 0; g.fuse_m = fuse_m ? 1 : 0;
    // k_wide_dw1's work rides in eight extra warps of this kernel (FP32 pipe beside the tensor pipe)
    g.dw1.params = (skip & 64) ? nullptr : a.params;   // (diagnostic bit 64: the GEMM without the dw1 warps' work)
    g.dw1.obs = a.tr.obs; g.dw1.obs_cap = a.tr.cap; g.dw1.idx = a.idx; g.dw1.first = wa.first;
    g.dw1.count = wa.count; g.dw1.H1 = H1; g.dw1.SIZE = d.SIZE; g.dw1.oW1 = d.oW1; g.dw1.ob1 = d.ob1;
    g.dw1.z1 = w->d1; g.dw1.cap = w->cap; g.dw1.part = w->part_dw1;
    g.dw1.sum = merge ? w->dw1_sum + (size_t)sum_phase * 2 * H1 * 8 : nullptr;
    g.tl = w->tl_state == 1 ? w->tl : nullptr;
    int dw2_grid = std::min(2 * (fuse_m ? 1 : jt) * splits, w->sm_count);
    // H2 = 256: CTA pairs (tcgen05 cta_group::2) own all 256 rows of dW2, every operand byte is fetched once, and the dW1
    // warps take Z1 through a cp.async ring (wide_gemm_mn2.cu, wide_dw1.cuh); B2_WIDE_DW2_PAIR=0: k_gemm_x3<MN>
    static const bool pair_env = [] { const char* e = getenv("B2_WIDE_DW2_PAIR"); return !(e && e[0] == '0'); }();
    const bool dw2_pair = pair_env && !fuse_m && !w->fuse_reduce && gemm_mn2_supported(H2, H1);
    if (dw2_pair) { g.stages = gemm_mn2_stages(H1); dw2_grid = 2 * std::min(w->sm_count / 2, 2 * splits); }
    // reduce (+ Adam on the last chunk): in the tail of the dW2 kernel behind a grid barrier, or as a launch of its own
    ReduceArgs r{};
    r.d = d; r.u = a; r.part_head = w->part_head; r.part_dw1 = w->part_dw1; r.part_dw2 = w->part_dw2;
    r.head_blocks = head_rows; r.head_row = w->head_row; r.dw1_blocks = dw2_grid; r.splits = splits;
    r.dw2_added = w->dw2_add ? 1 : 0;
    r.pstride = wide_pstride(H1, H2);
    r.w2p = w->w2p; r.bias2 = w->bias2; r.coef = step_coef;
    r.first_chunk = c == 0; r.last_chunk = c == n_chunks - 1; r.apply_adam = apply_adam ? 1 : 0;
    const int small = 2 * (H1 * (S + 1) + 2 * H2 + 1) + 3;
    const bool fused_tail = w->fuse_reduce && !(skip & (16 | 32));
    if (fused_tail) {
      const int bd = 320 + 512;                         // threads of the MN-major GEMM kernel (GEMM roles + dw1 warps)
      g.red = r;
      g.red_w2_blocks = (2 * H2 * H1 + bd - 1) / bd;
      g.red_vblocks = g.red_w2_blocks + (small + bd / 32 - 1) / (bd / 32);
      g.bar_counter = w->bar_counter;
      w->bar_count += dw2_grid;
      g.bar_target = w->bar_count;
    }
    if (!(skip & 16))
    if ((ce = dw2_pair ? launch_gemm_mn2(w->tm_d2_mn, w->tm_h1_mn, w->tm_out_dw2_16, g, w->sm_count, &dw2_grid, st)
                       : launch_gemm_x3(w->tm_d2_mn, w->tm_h1_mn, w->tm_out_dw2, g, w->sm_count, st)) != cudaSuccess)
      return set_error(B2_ERR_CUDA, "wide gemm dW2: %s", cudaGetErrorString(ce));
    (*launches)++;
    if (fused_tail) continue;
    if (merge) {              // this step's reduce + Adam: in the next step's layer-1 kernel, or in a launch without rows right now
      w->pend_r = r;
      w->pending = !(skip & 32);
      if (w->tl_state == 1) w->tl_state = 2;
      else if (w->tl_state == 3) {
        w->tl_state = 4;
        cudaStreamSynchronize(st);
        std::vector<unsigned long long> h(4 * 8 * 2);
        cudaMemcpy(h.data(), w->tl, sizeof(long long) * h.size(), cudaMemcpyDeviceToHost);
        const unsigned long long t0 = h[0];
        static const char* kn[4] = {"l1r", "fh", "mn2", "l1r+1"};
        static const char* sn[4][8] = {
            {"entry", "dep_wait_done", "w1_ready", "l1_done", "reduce_done", "", "", ""},
            {"entry", "dep_wait_done", "acc_critic", "acc_actor", "head_critic_done", "head_actor_done", "exit", "acc_last_dx"},
            {"entry", "dep_wait_done", "ops_first", "ops_last", "acc_ready", "epilogue_done", "dw1_done", "exit"},
            {"entry", "dep_wait_done", "w1_ready", "l1_done", "reduce_done", "", "", ""}};
        for (int k = 0; k < 4; k++)
          for (int sl = 0; sl < 8; sl++) {
            const unsigned long long lo = h[(k * 8 + sl) * 2], hi = h[(k * 8 + sl) * 2 + 1];
            if (!sn[k][sl][0] || hi == 0) continue;
            fprintf(stderr, "tl %-6s %-18s first %8.2f us  last %8.2f us\n", kn[k], sn[k][sl], (double)(lo - t0) * 1e-3, (double)(hi - t0) * 1e-3);
          }
      }
      if (!more_follow) { int rc = wide_flush(w, st, launches); if (rc) return rc; }
      continue;
    }
    const int w2_blocks = w->dw2_add ? (2 * H2 * H1 + 2 * RED_THREADS - 1) / (2 * RED_THREADS) : (2 * H2 * H1 + RED_THREADS - 1) / RED_THREADS;
    const int rgrid = w2_blocks + wide_reduce_dw1_blocks(d) + wide_reduce_head_blocks(d);
    if (!(skip & 32))
    if ((ce = (w->dw2_add ? launch_pdl(k_wide_reduce<true>, dim3(rgrid), dim3(RED_THREADS), 0, st, r, w2_blocks)
                          : launch_pdl(k_wide_reduce<false>, dim3(rgrid), dim3(RED_THREADS), 0, st, r, w2_blocks))) != cudaSuccess)
      return set_error(B2_ERR_CUDA, "wide reduce: %s", cudaGetErrorString(ce));
    (*launches)++;
  }
#undef LCK
  return B2_OK;
}

cudaError_t launch_wide_adam(WideNet* w, const UpdateArgs& a, cudaStream_t st) {
  const int grid = (w->d.P + 3 + 127) / 128;
  k_wide_adam<<<grid, 128, 0, st>>>(a, w->d, wide_pstride(w->d.H1, w->d.H2), w->w2p, w->bias2, w->coef);
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
    dfree(w->boot_idx);
    w->boot_idx = nullptr;
    B2_CUDA_CHECK(dmalloc(&w->boot_idx, sizeof(int32_t) * B));
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
