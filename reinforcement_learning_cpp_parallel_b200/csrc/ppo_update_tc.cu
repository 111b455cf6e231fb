// ppo_update_tc.cu -- K3 on tensor cores: evaluate + clipped-surrogate / value loss + backward of the reference's
// 6-64-32-1 nets for LARGE batches, as warp-level bf16x3 MMAs.
//
// Replaces ActorCritic::evaluate (src/ppo_nn.cpp:126-194), the loss (:263-290) and autograd backward (:301-302); same
// outputs as k_ppo_grad (ppo_update.cu): one row of `partials` per CTA, summed in fixed order by k_reduce.
//
// Why.  k_ppo_grad (FP32 FMA out of shared memory) sits at the ceiling of its formulation: the shared-memory data pipe
// runs at 84 % while the FMA pipe is half idle (profiles/r1l_prof_grad.txt), and 88 % of its flops are the 64 x 32
// layer, which over a batch of millions of rows is a real dense GEMM (north_star: tensor cores "only if a hidden width
// makes it a real dense GEMM" -- [8.4 M x 64] . [64 x 32] is one).  The 1e-4 gradient gate rules out one-pass bf16, so
// every fp32 operand is split x = hi + mid (two bf16, |x - hi - mid| <= 2^-17 |x|) and every product is three MMAs
// (mid.hi, hi.mid, hi.hi) accumulated in fp32: ~2^-16 relative per product, the split the 2 x 256 path already validates.
//
// Shape of the computation.  A WARP owns 16 samples at a time and keeps the whole chain in registers; there is no
// block-level synchronisation after the weights are staged.  With samples along the M axis of mma.m16n8k16 the
// accumulator layout of one product IS the A-fragment layout of the next (row = lane / 4 (+ 8), columns 2 (lane % 4)
// + {0, 1}), so
//     X --k8--> Z1 -tanh-> H1 --> Z2 -tanh-> H2 -> y -> loss head -> dz3 -> D2 --> D1pre -> D1
// never leaves the register file.  The three weight-gradient contractions run over SAMPLES, which must then sit on
// the K axis: each 8 x 8 bf16 block is transposed inside the warp with movmatrix (one instruction per register):
//     dW2[j][k]  += sum_s D2[s][j] H1[s][k]     A = T(D2),  B = T(H1)     (accumulators: 64 registers, all tiles of the warp)
//     dW1a[k][c] += sum_s D1[s][k] Xa[s][c]     A = T(D1),  B = T(Xa)     (Xa = [x, 1, 0]: column 6 is db1)
// and the head row (dW3, db2, db3) is FP32.  Per 16 samples and net: 24 k8 + 156 k16 MMAs, 84 movmatrix, ~130 LDS.32 of
// weight fragments (bank-conflict-free pitches), 48 + 16 tanh -- about 1,300 instructions per thread where the FMA
// kernel issues ~14,000.  Critic pass first (it leaves V in vcur for the actor's A = R - V, src/ppo_nn.cpp:274, Q7).
//
// The engine takes this kernel when the GLOBAL (mini)batch has >= 32,768 samples (engine.cu use_tensor_update); the
// reference-sized batches (2,000 / 2,100 records, tests/golden c1_* / c2_*) stay on the FP32 kernel, whose 80-epoch
// Adam trajectories are held to the libtorch goldens.  B2_UPDATE_TC=0 / 1 forces either (A/B runs, parity tests).
#include "engine.h"
#include "mlp.cuh"

namespace b2 {

namespace {

using N = NetDefault;
constexpr int H1 = N::H1, H2 = N::H2, S = N::S;
static_assert(H1 == 64 && H2 == 32 && S == 6, "k_ppo_grad_tc is specialised for the reference MLP");
constexpr int NT = 256, NW = NT / 32;
constexpr int PSTRIDE = ((N::P + 4 + 31) / 32) * 32;   // == ppo_update.cu

// shared memory (32-bit words).  Weight planes are packed bf16 pairs; pitches make every fragment load conflict-free.
constexpr int PW2 = 36;     // W2  [32 j][64 k]  as 32 words per row + 4
constexpr int PW2T = 20;    // W2T [64 k][32 j]  as 16 words per row + 4
constexpr int oW1h = 0;                    // [64 k][4 words] = 8 bf16 (c 0..5 = W1, 6 = b1, 7 = 0)
constexpr int oW1m = oW1h + H1 * 4;
constexpr int oW2h = oW1m + H1 * 4;
constexpr int oW2m = oW2h + H2 * PW2;
constexpr int oW2Th = oW2m + H2 * PW2;
constexpr int oW2Tm = oW2Th + H1 * PW2T;
constexpr int oW3 = oW2Tm + H1 * PW2T;     // fp32 [32]
constexpr int oB2 = oW3 + H2;              // fp32 [32]
constexpr int oB3 = oB2 + H2;              // fp32 [1] (+3 pad)
constexpr int oRedW2 = oB3 + 4;            // fp32 [32 * 64]  end-of-pass sum over the CTA's warps
constexpr int oRedW1 = oRedW2 + H2 * H1;   // fp32 [64 * 8]
constexpr int oRedRow = oRedW1 + H1 * 8;   // fp32 [2 * 32 + 1 (+3)]  db2, dW3, db3
constexpr int oRedLoss = oRedRow + 2 * H2 + 4;   // fp32 [3 * NW]
constexpr int SMEM_WORDS = oRedLoss + 3 * NW + 8;
constexpr int SMEM_BYTES = SMEM_WORDS * 4;

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {   // lower 16 bits = lo (the lower-index element)
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
// x = hi + mid (+ O(2^-17 |x|)): two packed bf16 pairs from two floats
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& mid) {
  hi = pack_bf16(x0, x1);
  const float h0 = __uint_as_float(hi << 16), h1 = __uint_as_float(hi & 0xffff0000u);
  mid = pack_bf16(x0 - h0, x1 - h1);
}
__device__ __forceinline__ float2 unsplit2(uint32_t hi, uint32_t mid) {
  return make_float2(__uint_as_float(hi << 16) + __uint_as_float(mid << 16),
                     __uint_as_float(hi & 0xffff0000u) + __uint_as_float(mid & 0xffff0000u));
}
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma1688(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(b0));
}
// bf16x3 product: c += A . B with A = ah + am, B = bh + bm (small terms first)
__device__ __forceinline__ void mma3_16(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&am)[4], uint32_t bh0,
                                        uint32_t bh1, uint32_t bm0, uint32_t bm1) {
  mma16816(c, am[0], am[1], am[2], am[3], bh0, bh1);
  mma16816(c, ah[0], ah[1], ah[2], ah[3], bm0, bm1);
  mma16816(c, ah[0], ah[1], ah[2], ah[3], bh0, bh1);
}
// transpose an 8 x 8 bf16 block held one packed register per lane (row = lane / 4, columns 2 (lane % 4) + {0, 1})
__device__ __forceinline__ uint32_t movt(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

__device__ __forceinline__ void stage_weights_tc(uint32_t* smw, const float* __restrict__ g) {
  float* smf = reinterpret_cast<float*>(smw);
  for (int i = threadIdx.x; i < H1 * 4; i += NT) {          // W1aug[k][c pair]
    const int k = i >> 2, c = (i & 3) * 2;
    float a0, a1;
    if (c < S) { a0 = g[N::oW1 + k * S + c]; a1 = g[N::oW1 + k * S + c + 1]; }
    else { a0 = g[N::ob1 + k]; a1 = 0.0f; }
    uint32_t h, m;
    split2(a0, a1, h, m);
    smw[oW1h + i] = h; smw[oW1m + i] = m;
  }
  for (int i = threadIdx.x; i < H2 * (H1 / 2); i += NT) {   // W2[j][k pair]
    const int j = i / (H1 / 2), kp = i % (H1 / 2);
    uint32_t h, m;
    split2(g[N::oW2 + j * H1 + 2 * kp], g[N::oW2 + j * H1 + 2 * kp + 1], h, m);
    smw[oW2h + j * PW2 + kp] = h; smw[oW2m + j * PW2 + kp] = m;
  }
  for (int i = threadIdx.x; i < H1 * (H2 / 2); i += NT) {   // W2T[k][j pair]
    const int k = i / (H2 / 2), jp = i % (H2 / 2);
    uint32_t h, m;
    split2(g[N::oW2 + (2 * jp) * H1 + k], g[N::oW2 + (2 * jp + 1) * H1 + k], h, m);
    smw[oW2Th + k * PW2T + jp] = h; smw[oW2Tm + k * PW2T + jp] = m;
  }
  for (int i = threadIdx.x; i < H2; i += NT) { smf[oW3 + i] = g[N::oW3 + i]; smf[oB2 + i] = g[N::ob2 + i]; }
  if (threadIdx.x == 0) smf[oB3] = g[N::ob3];
}

struct TileIn {          // what one lane needs of a 16-sample warp tile
  float x[4];            // obs[c0][r0], obs[c1][r0], obs[c0][r1], obs[c1][r1]   (c0 = 2 tig, c1 = c0 + 1)
  float act[2], olp[2], ret[2], vc[2];   // rows r0, r1
};

}  // namespace

__global__ void __launch_bounds__(NT, 1) k_ppo_grad_tc(const UpdateArgs a) {
  extern __shared__ __align__(16) uint32_t smw[];
  float* smf = reinterpret_cast<float*>(smw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, tig = lane & 3;
  const int64_t nwt = (a.count + 15) / 16;                 // warp tiles
  const int64_t wstride = (int64_t)gridDim.x * NW;
  const float inv_denom = 1.0f / (float)a.denom;
  float* part = a.partials + (int64_t)blockIdx.x * PSTRIDE;
  const int64_t cap = a.tr.cap;

  if (blockIdx.x == 0 && tid == 0) *a.adam_step += 1;      // one launch == one optimizer step

  float loss1_sum = 0.f, loss2_sum = 0.f, clip_sum = 0.f;  // lanes with tig == 0 only

#pragma unroll 1
  for (int net = 1; net >= 0; net--) {                     // critic first (:274)
    __syncthreads();
    stage_weights_tc(smw, a.params + net * N::SIZE);
    __syncthreads();

    float gW2[2][8][4];      // dW2[j = 16 mt + g (+8)][k = 8 nt + 2 tig + {0,1}]
    float gW1[4][4];         // dW1a[k = 16 mt + g (+8)][c = 2 tig + {0,1}]
    float gB2[4][2], gW3[4][2], gB3 = 0.f;   // partial over this lane's rows: [nt][col]
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int j = 0; j < 8; j++)
#pragma unroll
        for (int q = 0; q < 4; q++) gW2[i][j][q] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
      for (int q = 0; q < 4; q++) gW1[i][q] = 0.f;
      gB2[i][0] = gB2[i][1] = gW3[i][0] = gW3[i][1] = 0.f;
    }

    auto fetch = [&](int64_t wt, TileIn& t) {
      const int64_t base = wt * 16;
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int64_t s = base + g + 8 * h;
        const bool valid = s < a.count;
        const int64_t k = a.first + (valid ? s : 0);       // minibatches are contiguous ranges (no gather: launch_ppo_grad_tc refuses a.idx)
        if (tig < 3) {
          t.x[2 * h] = valid ? __ldg(a.tr.obs + (int64_t)(2 * tig) * cap + k) : 0.f;
          t.x[2 * h + 1] = valid ? __ldg(a.tr.obs + (int64_t)(2 * tig + 1) * cap + k) : 0.f;
        } else {
          t.x[2 * h] = 1.0f;            // the ones column: b1 rides in W1aug[.][6], db1 comes out of dW1a[.][6]
          t.x[2 * h + 1] = 0.0f;
        }
        t.act[h] = valid ? __ldg(a.tr.action + k) : 0.f;
        t.olp[h] = valid ? __ldg(a.tr.logp + k) : 0.f;
        t.ret[h] = valid ? __ldg(a.tr.ret + k) : 0.f;
        t.vc[h] = (valid && net == 0) ? a.tr.vcur[k] : 0.f;   // written by the critic pass of THIS launch: no __ldg
      }
    };

    int64_t wt = (int64_t)blockIdx.x * NW + warp;
    TileIn cur, nxt;
    if (wt < nwt) fetch(wt, cur);
#pragma unroll 1
    for (; wt < nwt; wt += wstride) {
      if (wt + wstride < nwt) fetch(wt + wstride, nxt);
      const int64_t base = wt * 16;

      // ---- X fragments (A of m16n8k8) and their transposes (B of the dW1 contraction) ----
      uint32_t xh[2], xm[2];
      split2(cur.x[0], cur.x[1], xh[0], xm[0]);
      split2(cur.x[2], cur.x[3], xh[1], xm[1]);

      // ---- L1: Z1[s][k] = Xa . W1a^T, H1 = tanh(Z1) ----  (ppo_nn.h:85-90 / :106-111)
      uint32_t h1h[8][2], h1m[8][2];     // [n-tile of k][row half]
#pragma unroll
      for (int nt = 0; nt < 8; nt++) {
        float c[4] = {0.f, 0.f, 0.f, 0.f};
        const uint32_t bh = smw[oW1h + (nt * 8 + g) * 4 + tig], bm = smw[oW1m + (nt * 8 + g) * 4 + tig];
        mma1688(c, xm[0], xm[1], bh);
        mma1688(c, xh[0], xh[1], bm);
        mma1688(c, xh[0], xh[1], bh);
        split2(act_tanh(c[0]), act_tanh(c[1]), h1h[nt][0], h1m[nt][0]);
        split2(act_tanh(c[2]), act_tanh(c[3]), h1h[nt][1], h1m[nt][1]);
      }

      // ---- L2: Z2[s][j] = H1 . W2^T + b2, H2 = tanh(Z2) ----  (:91-96 / :112-117)
      float h2[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; nt++) {
        const float2 b = *reinterpret_cast<const float2*>(smf + oB2 + nt * 8 + 2 * tig);
        float c[4] = {b.x, b.y, b.x, b.y};
#pragma unroll
        for (int ks = 0; ks < 4; ks++) {
          const uint32_t ah[4] = {h1h[2 * ks][0], h1h[2 * ks][1], h1h[2 * ks + 1][0], h1h[2 * ks + 1][1]};
          const uint32_t am[4] = {h1m[2 * ks][0], h1m[2 * ks][1], h1m[2 * ks + 1][0], h1m[2 * ks + 1][1]};
          const int w = (nt * 8 + g) * PW2 + ks * 8 + tig;
          mma3_16(c, ah, am, smw[oW2h + w], smw[oW2h + w + 4], smw[oW2m + w], smw[oW2m + w + 4]);
        }
#pragma unroll
        for (int q = 0; q < 4; q++) h2[nt][q] = act_tanh(c[q]);
      }

      // ---- L3 + loss head (FP32; the four lanes of a quad hold the same two rows) ----
      float w3[4][2];
      float y0 = 0.f, y1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 4; nt++) {
        const float2 w = *reinterpret_cast<const float2*>(smf + oW3 + nt * 8 + 2 * tig);
        w3[nt][0] = w.x; w3[nt][1] = w.y;
        y0 = fmaf(w.x, h2[nt][0], y0); y0 = fmaf(w.y, h2[nt][1], y0);
        y1 = fmaf(w.x, h2[nt][2], y1); y1 = fmaf(w.y, h2[nt][3], y1);
      }
      y0 += __shfl_xor_sync(0xffffffffu, y0, 1); y0 += __shfl_xor_sync(0xffffffffu, y0, 2);
      y1 += __shfl_xor_sync(0xffffffffu, y1, 1); y1 += __shfl_xor_sync(0xffffffffu, y1, 2);
      const float b3 = smf[oB3];
      float dz3[2];
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const float y = (h ? y1 : y0) + b3;
        const int64_t s = base + g + 8 * h;
        const bool valid = s < a.count;
        const bool own = valid && tig == 0;               // one lane of the quad books the sums and stores V
        const float R = cur.ret[h];
        float dz = 0.f;
        if (net == 1) {
          const float V = y;                               // critic: no final activation (ppo_nn.h:118-122)
          const float e = V - R;
          if (valid) dz = e * inv_denom;
          if (own) {
            loss2_sum = fmaf(e, e, loss2_sum);             // :288  0.5 * MSE
            a.tr.vcur[a.first + s] = V;
          }
        } else if (valid) {
          const float mu = act_tanh(y);                    // ppo_nn.h:102
          const float act = cur.act[h];
          const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);  // :140
          const float ratio = expf(lp - cur.olp[h]);       // :263
          const float A = R - cur.vc[h];                   // :274 (detached)
          const float surr1 = ratio * A;                   // :284
          const float rc = fminf(fmaxf(ratio, a.hp.clip_lo), a.hp.clip_hi);
          const float surr2 = rc * A;                      // :285
          const bool inside = (ratio >= a.hp.clip_lo) && (ratio <= a.hp.clip_hi);
          float gr;                                        // autograd of min(): see k_ppo_grad
          if (surr1 < surr2) gr = A;
          else if (surr1 > surr2) gr = inside ? A : 0.f;
          else gr = 0.5f * A + (inside ? 0.5f * A : 0.f);
          if (own) {
            loss1_sum -= fminf(surr1, surr2);              // :287
            clip_sum += inside ? 0.f : 1.f;
          }
          const float dlogp = -gr * ratio * inv_denom;
          const float dmu = dlogp * (act - mu) * a.hp.inv_var;   // :20-31
          dz = dmu * (1.0f - mu * mu);
        }
        dz3[h] = dz;
      }
      gB3 += dz3[0] + dz3[1];

      // ---- D2 = dz3 (x) W3 .* (1 - H2^2); head-row gradients ----
      uint32_t d2h[4][2], d2m[4][2];     // [n-tile of j][row half]
#pragma unroll
      for (int nt = 0; nt < 4; nt++) {
        const float d00 = dz3[0] * w3[nt][0] * (1.f - h2[nt][0] * h2[nt][0]);
        const float d01 = dz3[0] * w3[nt][1] * (1.f - h2[nt][1] * h2[nt][1]);
        const float d10 = dz3[1] * w3[nt][0] * (1.f - h2[nt][2] * h2[nt][2]);
        const float d11 = dz3[1] * w3[nt][1] * (1.f - h2[nt][3] * h2[nt][3]);
        gB2[nt][0] += d00 + d10; gB2[nt][1] += d01 + d11;
        gW3[nt][0] = fmaf(dz3[0], h2[nt][0], fmaf(dz3[1], h2[nt][2], gW3[nt][0]));
        gW3[nt][1] = fmaf(dz3[0], h2[nt][1], fmaf(dz3[1], h2[nt][3], gW3[nt][1]));
        split2(d00, d01, d2h[nt][0], d2m[nt][0]);
        split2(d10, d11, d2h[nt][1], d2m[nt][1]);
      }

      // ---- dW2[j][k] += sum_s D2[s][j] H1[s][k]:  A = T(D2) (M = j, K = s), B = T(H1) (K = s, N = k) ----
      {
        uint32_t ath[2][4], atm[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
          ath[mt][0] = movt(d2h[2 * mt][0]); ath[mt][1] = movt(d2h[2 * mt + 1][0]);
          ath[mt][2] = movt(d2h[2 * mt][1]); ath[mt][3] = movt(d2h[2 * mt + 1][1]);
          atm[mt][0] = movt(d2m[2 * mt][0]); atm[mt][1] = movt(d2m[2 * mt + 1][0]);
          atm[mt][2] = movt(d2m[2 * mt][1]); atm[mt][3] = movt(d2m[2 * mt + 1][1]);
        }
#pragma unroll
        for (int nt = 0; nt < 8; nt++) {
          const uint32_t bh0 = movt(h1h[nt][0]), bh1 = movt(h1h[nt][1]);
          const uint32_t bm0 = movt(h1m[nt][0]), bm1 = movt(h1m[nt][1]);
          mma3_16(gW2[0][nt], ath[0], atm[0], bh0, bh1, bm0, bm1);
          mma3_16(gW2[1][nt], ath[1], atm[1], bh0, bh1, bm0, bm1);
        }
      }

      // ---- D1 = (D2 . W2) .* (1 - H1^2); dW1a[k][c] += sum_s D1[s][k] Xa[s][c]:  A = T(D1), B = T(Xa) ----
      {
        const uint32_t xth0 = movt(xh[0]), xth1 = movt(xh[1]), xtm0 = movt(xm[0]), xtm1 = movt(xm[1]);
        const uint32_t ah0[4] = {d2h[0][0], d2h[0][1], d2h[1][0], d2h[1][1]}, am0[4] = {d2m[0][0], d2m[0][1], d2m[1][0], d2m[1][1]};
        const uint32_t ah1[4] = {d2h[2][0], d2h[2][1], d2h[3][0], d2h[3][1]}, am1[4] = {d2m[2][0], d2m[2][1], d2m[3][0], d2m[3][1]};
#pragma unroll
        for (int mt = 0; mt < 4; mt++) {       // 16 hidden-1 units = n-tiles 2 mt, 2 mt + 1 of D1
          uint32_t dth[4], dtm[4];
#pragma unroll
          for (int u = 0; u < 2; u++) {
            const int nt = 2 * mt + u;
            float c[4] = {0.f, 0.f, 0.f, 0.f};
            const int w = (nt * 8 + g) * PW2T + tig;
            mma3_16(c, ah0, am0, smw[oW2Th + w], smw[oW2Th + w + 4], smw[oW2Tm + w], smw[oW2Tm + w + 4]);
            mma3_16(c, ah1, am1, smw[oW2Th + w + 8], smw[oW2Th + w + 12], smw[oW2Tm + w + 8], smw[oW2Tm + w + 12]);
            const float2 ha = unsplit2(h1h[nt][0], h1m[nt][0]), hb = unsplit2(h1h[nt][1], h1m[nt][1]);
            uint32_t ph0, pm0, ph1, pm1;
            split2(c[0] * (1.f - ha.x * ha.x), c[1] * (1.f - ha.y * ha.y), ph0, pm0);
            split2(c[2] * (1.f - hb.x * hb.x), c[3] * (1.f - hb.y * hb.y), ph1, pm1);
            // A of the k-contraction: rows = hidden unit (u selects the +8 half), K = sample (row half selects +8)
            dth[u] = movt(ph0); dth[2 + u] = movt(ph1);
            dtm[u] = movt(pm0); dtm[2 + u] = movt(pm1);
          }
          mma3_16(gW1[mt], dth, dtm, xth0, xth1, xtm0, xtm1);
        }
      }
      cur = nxt;
    }  // warp tiles

    // ---- end of pass: fold the CTA's eight warps in a fixed order, write this CTA's partial row ----
    // head row: sum over the lanes that share tig (rows g), then lanes g == 0 hold the warp's totals
#pragma unroll
    for (int nt = 0; nt < 4; nt++)
#pragma unroll
      for (int q = 0; q < 2; q++) {
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          gB2[nt][q] += __shfl_xor_sync(0xffffffffu, gB2[nt][q], o);
          gW3[nt][q] += __shfl_xor_sync(0xffffffffu, gW3[nt][q], o);
        }
      }
    if (tig != 0) gB3 = 0.f;                               // every lane of a quad summed the same dz3
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) gB3 += __shfl_xor_sync(0xffffffffu, gB3, o);
    __syncthreads();                                       // every warp is done with the weight planes
    for (int w = 0; w < NW; w++) {
      if (warp == w) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
          for (int nt = 0; nt < 8; nt++)
#pragma unroll
            for (int q = 0; q < 4; q++) {
              const int j = 16 * mt + g + 8 * (q >> 1), k = 8 * nt + 2 * tig + (q & 1);
              float* d = smf + oRedW2 + j * H1 + k;
              *d = (w == 0 ? 0.f : *d) + gW2[mt][nt][q];
            }
#pragma unroll
        for (int mt = 0; mt < 4; mt++)
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const int k = 16 * mt + g + 8 * (q >> 1), c = 2 * tig + (q & 1);
            float* d = smf + oRedW1 + k * 8 + c;
            *d = (w == 0 ? 0.f : *d) + gW1[mt][q];
          }
        if (g == 0) {
#pragma unroll
          for (int nt = 0; nt < 4; nt++)
#pragma unroll
            for (int q = 0; q < 2; q++) {
              const int j = 8 * nt + 2 * tig + q;
              float* d = smf + oRedRow + j;
              *d = (w == 0 ? 0.f : *d) + gB2[nt][q];
              float* e = smf + oRedRow + H2 + j;
              *e = (w == 0 ? 0.f : *e) + gW3[nt][q];
            }
        }
        if (lane == 0) {
          float* d = smf + oRedRow + 2 * H2;
          *d = (w == 0 ? 0.f : *d) + gB3;
        }
      }
      __syncthreads();
    }
    float* gp = part + net * N::SIZE;
    for (int i = tid; i < H2 * H1; i += NT) gp[N::oW2 + i] = smf[oRedW2 + i];
    for (int i = tid; i < H1 * 8; i += NT) {
      const int k = i >> 3, c = i & 7;
      if (c < S) gp[N::oW1 + k * S + c] = smf[oRedW1 + i];
      else if (c == S) gp[N::ob1 + k] = smf[oRedW1 + i];
    }
    for (int i = tid; i < H2; i += NT) { gp[N::ob2 + i] = smf[oRedRow + i]; gp[N::oW3 + i] = smf[oRedRow + H2 + i]; }
    if (tid == 0) gp[N::ob3] = smf[oRedRow + 2 * H2];
  }  // nets

  // ---- loss partials of this CTA ----
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    loss1_sum += __shfl_xor_sync(0xffffffffu, loss1_sum, o);
    loss2_sum += __shfl_xor_sync(0xffffffffu, loss2_sum, o);
    clip_sum += __shfl_xor_sync(0xffffffffu, clip_sum, o);
  }
  __syncthreads();
  if (lane == 0) { smf[oRedLoss + warp] = loss1_sum; smf[oRedLoss + NW + warp] = loss2_sum; smf[oRedLoss + 2 * NW + warp] = clip_sum; }
  __syncthreads();
  if (tid < 3) {
    float v = 0.f;
    for (int w = 0; w < NW; w++) v += smf[oRedLoss + NW * tid + w];
    part[N::P + tid] = v;
  }
}

int update_grid_for_tc(int64_t count, int sm_count) {
  const int64_t nwt = (count + 15) / 16;
  int64_t g = (nwt + NW - 1) / NW;
  if (g > sm_count) g = sm_count;         // one CTA per SM (255 registers per thread)
  if (g < 1) g = 1;
  return (int)g;
}

cudaError_t launch_ppo_grad_tc(const UpdateArgs& a, cudaStream_t st) {
  if (a.idx) return cudaErrorInvalidValue;   // gather indices are a k_ppo_grad feature; the engine's minibatches are ranges
  k_ppo_grad_tc<<<a.grid, NT, SMEM_BYTES, st>>>(a);
  return cudaGetLastError();
}

}  // namespace b2
