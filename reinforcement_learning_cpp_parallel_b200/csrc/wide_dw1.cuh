// wide_dw1.cuh -- the dW1 / db1 partials of a wide-net step, carried by 16 extra warps of the dW2 GEMM kernels
// (k_gemm_x3<MN> in wide_gemm.cu: dw1_worker; k_gemm_mn2 in wide_gemm_mn2.cu: dw1_worker_async): FP32 pipe beside the
// tensor pipe.
#pragma once
#include <type_traits>

#include "tc_ptx.cuh"
#include "wide.h"
#include "mlp.cuh"

namespace b2 {
namespace {

constexpr int DW1_THREADS = 512;              // extra warps 10..25 of the MN-major kernel (fused k_wide_dw1 work)
constexpr int DW1_SMEM = (2 * 64 * 8 + 256 * 8) * 4;   // two 64-sample observation tiles + the fold buffer

// The dW1 / db1 partials of this CTA's slice of samples (see Dw1Args).  512 threads = 2 sample halves x 256 features;
// a thread carries its feature of BOTH nets, so the observation tile is staged once and the Z1 loads of the two nets are
// in flight together.  Loads are coalesced across k and batched: 8 samples x 2 nets in flight (16 x 2 spills at this
// kernel's 72-register cap and is 2 ms per update slower), then 4 / 2 / 1 for the remainder -- an L2 round trip is
// ~2,000 cycles beside the GEMM's operand stream, and these warps are nothing but
// a chain of them (clock64 stamps: with a one-sample-at-a-time tail and the nets one after the other they ran 49k cycles
// against the GEMM's 40k, 5 us per step on the critical path).  H1 is recomputed from x exactly as k_wide_l1 computes
// it; per net and feature the adds keep the sample order; the two halves are folded through shared memory in a fixed order.
__device__ __forceinline__ void dw1_worker(const Dw1Args& w, float* sm, int t) {
  const int spb = (w.count + (int)gridDim.x - 1) / (int)gridDim.x;
  const int sb = min((int)blockIdx.x * spb, w.count), se = min(sb + spb, w.count);
  const int half = t >> 8, k = t & 255;
  const int hlen = (se - sb + 1) / 2;
  const int hb = sb + half * hlen, he = min(hb + hlen, se);
  float* xs = sm + half * 64 * 8;
  float* fold = sm + 2 * 64 * 8;
  const bool live = k < w.H1;
  float acc[2][OBS + 1], w1[2][OBS], b1[2];
#pragma unroll
  for (int g = 0; g < 2; g++) {
#pragma unroll
    for (int c = 0; c <= OBS; c++) acc[g][c] = 0.f;
#pragma unroll
    for (int c = 0; c < OBS; c++) w1[g][c] = live ? w.params[g * w.SIZE + w.oW1 + k * OBS + c] : 0.f;
    b1[g] = live ? w.params[g * w.SIZE + w.ob1 + k] : 0.f;
  }
  for (int off = 0; off < hlen; off += 64) {           // uniform trip count for both halves
    const int t0 = hb + off;
    const int nt = max(min(64, he - t0), 0);
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    for (int i = k; i < nt * OBS; i += 256) {
      const int s = i / OBS, c = i % OBS;
      const int64_t p = w.first + t0 + s;
      xs[s * 8 + c] = w.obs[c * w.obs_cap + (w.idx ? (int64_t)w.idx[p] : p)];
    }
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    if (live && nt > 0) {
      const float* dp0 = w.z1 + (int64_t)t0 * w.H1 + k;
      const float* dp1 = dp0 + w.cap * w.H1;
      auto batch = [&](auto bc, int s) {
        constexpr int B = decltype(bc)::value;
        float dv[2][B];
#pragma unroll
        for (int u = 0; u < B; u++) {
          dv[0][u] = dp0[(int64_t)(s + u) * w.H1];
          dv[1][u] = dp1[(int64_t)(s + u) * w.H1];
        }
#pragma unroll
        for (int u = 0; u < B; u++) {
          const float* x = xs + (s + u) * 8;
#pragma unroll
          for (int g = 0; g < 2; g++) {
            float z = b1[g];
#pragma unroll
            for (int c = 0; c < OBS; c++) z = fmaf(w1[g][c], x[c], z);
            const float h = act_tanh(z);
            const float d = dv[g][u] * (1.0f - h * h);
#pragma unroll
            for (int c = 0; c < OBS; c++) acc[g][c] = fmaf(d, x[c], acc[g][c]);
            acc[g][OBS] += d;
          }
        }
      };
      int s = 0;
      for (; s + 8 <= nt; s += 8) batch(std::integral_constant<int, 8>{}, s);
      if (s + 4 <= nt) { batch(std::integral_constant<int, 4>{}, s); s += 4; }
      if (s + 2 <= nt) { batch(std::integral_constant<int, 2>{}, s); s += 2; }
      if (s < nt) batch(std::integral_constant<int, 1>{}, s);
    }
  }
#pragma unroll
  for (int g = 0; g < 2; g++) {
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    if (half == 1) {
#pragma unroll
      for (int c = 0; c <= OBS; c++) fold[k * 8 + c] = acc[g][c];
    }
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    if (half == 0 && live) {
      float* out = w.part + ((int64_t)blockIdx.x * 2 + g) * w.H1 * 8 + k * 8;
#pragma unroll
      for (int c = 0; c <= OBS; c++) out[c] = acc[g][c] + fold[k * 8 + c];
      out[OBS + 1] = 0.f;
    }
  }
}


// The same work with Z1 arriving through BULK async copies (cp.async.bulk, the TMA's 1-D form) into a shared-memory ring:
// the rows z1[g][s .. s + 7][0 .. H1) of a group of 8 samples are contiguous in memory, so ONE elected thread per sample
// half moves a group with two copies (one per net, 8 KB each at H1 = 256) and an mbarrier counts the bytes; the 256
// threads of the half wait on the barrier, read their own column and hand the slot back through a named barrier.
// Two groups per half are in flight (16 samples x 2 nets per thread, twice what 72 registers allow as register batches)
// at zero cost in registers and LSU issue.  (Per-thread 4-byte cp.async copies were tried first: 26.3 vs 25.3 ms per
// configs[4] update -- an LDGSTS costs ~8 cycles of the LSU per warp and there are 1,760 of them per SM.)
// The warps are a chain of L2 round trips (~2,000 cycles each beside the GEMM's operand stream): halving their number is
// what lets the CTA-pair GEMM (k_gemm_mn2) set the kernel's length instead of these warps.
constexpr int DW1_GS = 8;                     // samples per group
constexpr int DW1_NGR = 2;                    // groups in flight per half
constexpr int DW1_RING_BYTES = 2 * DW1_NGR * 2 * DW1_GS * 256 * 4;   // halves x groups x nets x samples x 256 floats = 64 KB
constexpr int DW1_SMEM_ASYNC = DW1_SMEM + DW1_RING_BYTES + 64;       // + the ring's mbarriers

// (evict-first hint: Z1 is read here for the last time in the step)
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(tc::smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(tc::smem_u32(bar)), "l"(policy) : "memory");
}

__device__ __forceinline__ void dw1_worker_async(const Dw1Args& w, float* sm, int t) {
  const int spb = (w.count + (int)gridDim.x - 1) / (int)gridDim.x;
  const int sb = min((int)blockIdx.x * spb, w.count), se = min(sb + spb, w.count);
  const int half = t >> 8, k = t & 255;
  const int hlen = (se - sb + 1) / 2;
  const int hb = sb + half * hlen, he = min(hb + hlen, se);
  const int n = max(he - hb, 0);                       // samples of this half
  const int H1 = w.H1;
  float* xs = sm + half * 64 * 8;
  float* fold = sm + 2 * 64 * 8;
  float* ring = sm + DW1_SMEM / 4 + half * (DW1_NGR * 2 * DW1_GS * 256);   // [group][net][sample][H1]
  uint64_t* rbar = reinterpret_cast<uint64_t*>(sm + (DW1_SMEM + DW1_RING_BYTES) / 4) + half * DW1_NGR;
  const bool live = k < H1;
  const bool producer = k == 0;
  if (t == 0) {
    uint64_t* b0 = reinterpret_cast<uint64_t*>(sm + (DW1_SMEM + DW1_RING_BYTES) / 4);
    for (int i = 0; i < 2 * DW1_NGR; i++) tc::mbar_init(&b0[i], 1);
    tc::fence_barrier_init();
  }
  float acc[2][OBS + 1], w1[2][OBS], b1[2];
#pragma unroll
  for (int g = 0; g < 2; g++) {
#pragma unroll
    for (int c = 0; c <= OBS; c++) acc[g][c] = 0.f;
#pragma unroll
    for (int c = 0; c < OBS; c++) w1[g][c] = live ? w.params[g * w.SIZE + w.oW1 + k * OBS + c] : 0.f;
    b1[g] = live ? w.params[g * w.SIZE + w.ob1 + k] : 0.f;
  }
  asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");   // the ring's barriers are initialised
  const int total_groups = (n + DW1_GS - 1) / DW1_GS;
  const uint64_t pol_last = tc::l2_policy_evict_first();
  auto issue = [&](int gidx) {                         // producer only: group gidx -> ring slot gidx % NGR
    const int gi = gidx % DW1_NGR;
    const int s0 = gidx * DW1_GS, cnt = min(DW1_GS, n - s0);
    const uint32_t bytes = (uint32_t)(cnt * H1 * 4);
    tc::mbar_arrive_expect_tx(&rbar[gi], 2 * bytes);
#pragma unroll
    for (int g = 0; g < 2; g++)
      bulk_load_1d(ring + ((gi * 2 + g) * DW1_GS) * 256, w.z1 + ((int64_t)g * w.cap + hb + s0) * H1, bytes, &rbar[gi], pol_last);
  };
  if (producer)
    for (int gidx = 0; gidx < DW1_NGR && gidx < total_groups; gidx++) issue(gidx);
  const int trips = (hlen + 63) / 64;                  // uniform trip count for both halves (the x tiles use block barriers)
  for (int tile = 0; tile < trips; tile++) {
    const int t0 = tile * 64;                          // offset inside this half
    const int nt = max(min(64, n - t0), 0);
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    for (int i = k; i < nt * OBS; i += 256) {
      const int s = i / OBS, c = i % OBS;
      const int64_t p = w.first + hb + t0 + s;
      xs[s * 8 + c] = w.obs[c * w.obs_cap + (w.idx ? (int64_t)w.idx[p] : p)];
    }
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    for (int gl = 0; gl < 64 / DW1_GS; gl++) {
      const int gidx = tile * (64 / DW1_GS) + gl;
      if (gidx >= total_groups) break;                 // (uniform within the half)
      const int gi = gidx % DW1_NGR;
      tc::mbar_wait(&rbar[gi], (uint32_t)(gidx / DW1_NGR) & 1u);
      const int cnt = min(DW1_GS, n - gidx * DW1_GS);
      if (live) {
        const float* z0 = ring + ((gi * 2 + 0) * DW1_GS) * 256 + k;
        const float* z1r = ring + ((gi * 2 + 1) * DW1_GS) * 256 + k;
        for (int u = 0; u < cnt; u++) {
          const float dv[2] = {z0[u * H1], z1r[u * H1]};
          const float* x = xs + (gl * DW1_GS + u) * 8;
#pragma unroll
          for (int g = 0; g < 2; g++) {
            float z = b1[g];
#pragma unroll
            for (int c = 0; c < OBS; c++) z = fmaf(w1[g][c], x[c], z);
            const float h = act_tanh(z);
            const float d = dv[g] * (1.0f - h * h);
#pragma unroll
            for (int c = 0; c < OBS; c++) acc[g][c] = fmaf(d, x[c], acc[g][c]);
            acc[g][OBS] += d;
          }
        }
      }
      if (half == 0) asm volatile("bar.sync 2, 256;" ::: "memory");   // every thread of the half has read the slot
      else asm volatile("bar.sync 3, 256;" ::: "memory");
      if (producer && gidx + DW1_NGR < total_groups) issue(gidx + DW1_NGR);
    }
  }
  // this CTA's row [2][H1][8]: to part[cta], or (w.sum) assembled in the drained ring and ADDED into the one row of the step
  // with a single bulk reduce (fp32 adds in the L2: 16 KB per CTA instead of a partial row the reduce kernel sums later)
  float* rowbuf = sm + DW1_SMEM / 4;
#pragma unroll
  for (int g = 0; g < 2; g++) {
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    if (half == 1) {
#pragma unroll
      for (int c = 0; c <= OBS; c++) fold[k * 8 + c] = acc[g][c];
    }
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    if (half == 0 && live) {
      float* out = w.sum ? rowbuf + (g * H1 + k) * 8 : w.part + ((int64_t)blockIdx.x * 2 + g) * H1 * 8 + k * 8;
#pragma unroll
      for (int c = 0; c <= OBS; c++) out[c] = acc[g][c] + fold[k * 8 + c];
      out[OBS + 1] = 0.f;
    }
  }
  if (w.sum) {
    tc::fence_proxy_async();
    asm volatile("bar.sync 1, %0;" ::"n"(DW1_THREADS) : "memory");
    if (t == 0) {
      tc::bulk_reduce_add_f32(w.sum, rowbuf, (uint32_t)(2 * H1 * 8 * 4));
      tc::bulk_commit();
      tc::bulk_wait_read<0>();                         // the CTA's shared memory outlives the read (k_gemm_mn2 ends in a cluster barrier)
    }
  }
}

}  // namespace
}  // namespace b2
