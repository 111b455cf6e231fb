// wide_fused.cu -- k_wide_fh: the layer-2 forward GEMM AND the loss head of a wide-net optimizer step in ONE kernel.
//
//   per net g (critic first, src/ppo_nn.cpp:274):   Z2 = H1 W2^T                      tcgen05, CTA pairs, W2 halves resident
//                                                    H2 = tanh(Z2 + b2), y = W3 H2 + b3   epilogue, straight out of TMEM
//                                                    loss head (:263-290) -> dz
//                                                    D2 = dz (x) W3 .* (1 - H2^2)      -> (hi, mid) bf16 planes d2p (TMA store)
//                                                    dW3, db2, db3, loss sums          -> this CTA's partial row
//
// Replaces k_gemm2_ws<false> + k_wide_head (wide_gemm2.cu, wide_update.cu).  What that pair paid for (profiles/r2/
// r2e_wide_marginal.txt: 30.5 us of a 97-us step for the two together): Z2 went to memory as fp32 (33.5 MB written, 33.5 MB
// read per 16,384-sample step), the head was a launch of its own with two exposed L2 round trips per warp, and the GEMM's
// last epilogue and the head's ramp-up were serial.  Here Z2 never leaves TMEM: the epilogue warps run the head on the
// accumulator while the tensor pipe multiplies the next tile.
//
// Schedule.  A CTA pair (tcgen05 cta_group::2, M = 256) owns the 256-row tiles t = pair, pair + P, ...  It walks them
// TWICE: first with the critic's W2 half resident in shared memory, then with the actor's (the actor's advantage needs V
// of the same sample, A = R - V: V waits in vtmp, written and read by the same thread).  The weights are swapped K block
// by K block: after the last critic tile's MMAs of block kb are issued a tcgen05.commit frees that block (wfree[kb]) and
// the actor's block kb is loaded over it while the later blocks still multiply.
//
// The head per thread (= accumulator row = sample), 32 columns at a time, two passes over the row:
//   1. tcgen05.ld Z2 -> h = tanh(z + b2) -> y += w3 h -> tcgen05.st h back over Z2.  The two warps that share a row (column
//      halves) add their partial dot products through shared memory; every thread then knows y, hence V / mu and dz.
//   2. tcgen05.ld h -> dd = dz w3 (1 - h^2) -> bf16 (hi, mid) tiles in swizzled staging -> TMA store to both planes.
//      Column sums (db2 = sum_s dd, dW3 = sum_s dz h) cross the rows of a warp through a 32 x 32 fp32 transpose in the
//      same staging buffer: lane j adds column j over the 32 rows in a fixed order.  They accumulate in registers over
//      the CTA's tiles; the four row quarters are folded in order at the end of a net's pass.
// Everything is summed in a fixed order: bit-reproducible run to run.
#include <algorithm>

#include "tc_ptx.cuh"
#include "wide.h"
#include "mlp.cuh"

namespace b2 {

namespace {

constexpr int BM = 128, BK = 64;
constexpr int HEAD_WARPS = 16;
constexpr int FH_THREADS = 64 + HEAD_WARPS * 32;   // TMA warp, MMA warp, 16 head warps (4 per sub-partition: the head is latency-bound)
constexpr int A_PLANE = BM * BK * 2;          // 16 KB
constexpr int A_STAGE = 2 * A_PLANE;          // hi + mid
constexpr int STG_WARP = 2048;                // per head warp: one 32 x 16 fp32 tile, or two 32 x 16 bf16 tiles (hi | mid)
constexpr int STG_BYTES = HEAD_WARPS * STG_WARP;
constexpr int YBUF_BYTES = 0;
constexpr int BAR_BYTES = 256;
constexpr int SMEM_LIMIT = 227 * 1024;

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FH_THREADS, 1)
k_wide_fh(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
          const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmA2,
          const __grid_constant__ CUtensorMap tmB2, const __grid_constant__ CUtensorMap tmC, const FhArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = tc::cluster_ctarank();            // 0 = leader
  const int N = a.N, NH = N / 2, KB = a.K / BK, stages = a.stages;
  const int B_KBLOCK = NH * 128;                          // one plane of one K block of this CTA's weight half
  const int B_BYTES = KB * 2 * B_KBLOCK;
  uint8_t* sB = smem;
  uint8_t* sA = smem + B_BYTES;
  uint8_t* sStg = sA + (size_t)stages * A_STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(sStg + STG_BYTES + YBUF_BYTES);
  uint64_t* empty = full + 4;
  uint64_t* tfull = empty + 4;
  uint64_t* tempty = tfull + 2;
  uint64_t* bfull = tempty + 2;                           // [4] one per K block of the resident weights (K <= 256)
  uint64_t* wfree = bfull + 4;                            // [4] weight block kb of the pass that ends is no longer read
  uint64_t* d2done = wfree + 4;                           // [2] (a.dx) every D2 tile of net g written by THIS CTA's head warps has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d2done + 2);
  // Passes.  0, 1: forward GEMM + head of the critic / the actor.  2, 3 (a.dx): the dX GEMM Z1 = D2 W2 of the critic / the
  // actor on the SAME pair -- its A operand is the D2 tile this CTA's head warps have just stored, its weights (W2 as
  // stored, read as an MN-major operand, this CTA's half of the COLUMNS) land over the forward weights, its MMAs run while
  // the tensor pipe would otherwise idle under the actor's head, and its raw epilogue is done by the head warps.
  // One launch, one prologue and one weight fetch less than k_wide_fh + k_gemm2_ws<true>, and no drain between them.
  const int npass = a.dx ? 4 : 2;
  const int N2 = a.K, NH2 = N2 / 2, KB2 = a.N / BK;       // dX: N = H1 output columns, K = H2
  const int B2_KBLOCK = NH2 * 128;                        // (KB2 * 2 * B2_KBLOCK == B_BYTES: the same H1 H2 2 bytes)

  const int n_pairs = gridDim.x / 2, pair = blockIdx.x / 2;
  const int tiles = (a.m_pad + 2 * BM - 1) / (2 * BM);

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    tc::tma_prefetch_desc(&tmD);
    if (a.dx) { tc::tma_prefetch_desc(&tmA2); tc::tma_prefetch_desc(&tmB2); tc::tma_prefetch_desc(&tmC); }
    for (int b = 0; b < 2; b++) tc::mbar_init(&d2done[b], HEAD_WARPS);
    for (int s = 0; s < stages; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; b++) { tc::mbar_init(&tfull[b], 1); tc::mbar_init(&tempty[b], 2 * HEAD_WARPS * 32); }
    for (int kb = 0; kb < 4; kb++) { tc::mbar_init(&bfull[kb], 1); tc::mbar_init(&wfree[kb], 1); }
    tc::fence_barrier_init();
  }
  if (a.dbg && threadIdx.x == 64) a.dbg[blockIdx.x * 16 + 12] = clock64();   // kernel entry (slots 0..11: up to 3 tiles x 4 stamps)
  if (threadIdx.x == 64) tc::tl_stamp(a.tl, 1, 0);
  pdl_launch_dependents();
  if (warp == 1) tc::tmem_alloc_2sm(tmem_slot, 512);
  pdl_wait();                                             // h1p comes from k_wide_l1, the weights from the previous step's Adam
  if (threadIdx.x == 64) tc::tl_stamp(a.tl, 1, 1);
  tc::tc_fence_before();
  tc::cluster_sync();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      uint32_t wfree_seen[4] = {0, 0, 0, 0};              // completions of wfree[kb] consumed so far (parity = count & 1)
      for (int pass = 0; pass < npass; pass++) {
        const bool isdx = pass >= 2;
        const int g = (pass & 1) ? 0 : 1;                 // critic first
        const int KBp = isdx ? KB2 : KB;
        const int KBprev = pass == 0 ? 0 : (pass <= 2 ? KB : KB2);   // K blocks of the pass whose weights are being replaced
        bool d2_ready = !isdx;                            // (the weights do not depend on D2: they are fetched first)
        if (pass >= 2)                                    // different block layout: every block of the old weights must be free
          for (int kb = 0; kb < KBprev; kb++) { tc::mbar_wait(&wfree[kb], wfree_seen[kb] & 1u); wfree_seen[kb]++; }
        bool first_tile = true;
        for (int t = pair; t < tiles; t += n_pairs, first_tile = false) {
          const int row0 = t * 2 * BM + (int)rank * BM;
          // this pass's resident weight half, one barrier per K block.  Forward passes: block kb in front of the first tile's
          // A stage kb (the first MMA waits for a quarter of the weights).  dX passes: all blocks at once, BEFORE the wait for
          // the D2 tiles -- the weights do not depend on them.
          auto load_weights = [&](int kb) {
            if (rank == 0) tc::mbar_arrive_expect_tx(&bfull[kb], (uint32_t)(4 * (isdx ? B2_KBLOCK : B_KBLOCK)));   // 2 planes x 2 CTAs
            for (int pl = 0; pl < 2; pl++) {
              if (isdx) {
                uint8_t* dst = sB + (size_t)(kb * 2 + pl) * B2_KBLOCK;
                for (int fb = 0; fb < NH2 / 64; fb++)
                  tc::tma_load_3d_2sm(dst + fb * (BK * 128), &tmB2, &bfull[kb], (int)rank * NH2 + fb * 64, kb * BK, g * 2 + pl);
              } else {
                tc::tma_load_3d_2sm(sB + (size_t)(kb * 2 + pl) * B_KBLOCK, &tmB, &bfull[kb], kb * BK, (int)rank * NH, g * 2 + pl);
              }
            }
          };
          if (first_tile && isdx)
            for (int kb = 0; kb < KBp; kb++) load_weights(kb);
          for (int kb = 0; kb < KBp; kb++) {
            if (first_tile && !isdx) {
              if (pass == 1) { tc::mbar_wait(&wfree[kb], wfree_seen[kb] & 1u); wfree_seen[kb]++; }   // same layout: block by block
              load_weights(kb);
            }
            if (!d2_ready) {                              // the D2 tiles of net g are complete in global memory
              tc::mbar_wait(&d2done[g], 0);
              tc::fence_proxy_async();
              d2_ready = true;
            }
            tc::mbar_wait(&empty[stage], phase ^ 1);
            if (rank == 0) tc::mbar_arrive_expect_tx(&full[stage], (uint32_t)(2 * A_STAGE));
            uint8_t* s = sA + (size_t)stage * A_STAGE;
            tc::tma_load_3d_2sm(s, isdx ? &tmA2 : &tmA, &full[stage], kb * BK, row0, g * 2 + 0);
            tc::tma_load_3d_2sm(s + A_PLANE, isdx ? &tmA2 : &tmA, &full[stage], kb * BK, row0, g * 2 + 1);
            if (++stage == stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      uint32_t bfull_seen[4] = {0, 0, 0, 0};
      for (int pass = 0; pass < npass; pass++) {
        const bool isdx = pass >= 2;
        const int KBp = isdx ? KB2 : KB;
        const uint32_t idesc = isdx ? (tc::idesc_bf16_f32(2 * BM, N2) | (1u << 16)) : tc::idesc_bf16_f32(2 * BM, N);
        bool first_tile = true;
        for (int t = pair; t < tiles; t += n_pairs, it++, first_tile = false) {
          const bool last_of_pass = pass + 1 < npass && t + n_pairs >= tiles;   // the next pass replaces the weights
          const int buf = it & 1;
          const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
          tc::mbar_wait(&tempty[buf], bphase ^ 1);         // both CTAs' head warps are done with this accumulator
          tc::tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)buf * 256u;
          uint32_t acc = 0;
          for (int kb = 0; kb < KBp; kb++) {
            if (first_tile) { tc::mbar_wait(&bfull[kb], bfull_seen[kb] & 1u); bfull_seen[kb]++; }
            tc::mbar_wait(&full[stage], phase);
            tc::tc_fence_after();
            const uint32_t sa = tc::smem_u32(sA + (size_t)stage * A_STAGE);
            const uint64_t dAh = tc::smem_desc_sw128(sa), dAm = tc::smem_desc_sw128(sa + A_PLANE);
            uint64_t dBh, dBm, bstep;
            if (isdx) {                                   // W2 as stored, [K = j][N = k]: MN-major operand, 64-column boxes BK * 128 B apart
              const uint32_t sb = tc::smem_u32(sB + (size_t)(kb * 2) * B2_KBLOCK);
              dBh = tc::smem_desc_mn_sw128(sb, BK * 128); dBm = tc::smem_desc_mn_sw128(sb + B2_KBLOCK, BK * 128);
              bstep = 2048 >> 4;
            } else {
              const uint32_t sb = tc::smem_u32(sB + (size_t)(kb * 2) * B_KBLOCK);
              dBh = tc::smem_desc_sw128(sb); dBm = tc::smem_desc_sw128(sb + B_KBLOCK);
              bstep = 32 >> 4;
            }
#pragma unroll
            for (int k = 0; k < BK / 16; k++) {
              const uint64_t off = (uint64_t)((k * 32) >> 4), offb = (uint64_t)k * bstep;
              tc::umma_bf16_2sm(d_tmem, dAm + off, dBh + offb, idesc, acc);   // small terms first
              tc::umma_bf16_2sm(d_tmem, dAh + off, dBm + offb, idesc, 1u);
              tc::umma_bf16_2sm(d_tmem, dAh + off, dBh + offb, idesc, 1u);
              acc = 1u;
            }
            tc::umma_commit_2sm(&empty[stage], 3);
            if (last_of_pass) tc::umma_commit_2sm(&wfree[kb], 3);   // both CTAs may overwrite this weight block
            if (++stage == stages) { stage = 0; phase ^= 1; }
          }
          tc::umma_commit_2sm(&tfull[buf], 3);
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ the head, on the accumulator
    // 16 warps: q = TMEM lane quarter (32 rows), cq = column quarter; a warp walks its N / 4 columns 16 at a time
    const int q = warp & 3;
    const int cq = (warp - 2) >> 2;
    const int c_begin = cq * (N / 4), c_end = c_begin + N / 4;
    uint8_t* sb = sStg + (warp - 2) * STG_WARP;           // 2 KB: a 32 x 16 fp32 tile, or the (hi | mid) 32 x 16 bf16 tiles
    const int r = q * 32 + lane;                          // accumulator row inside this CTA's 128
    const int cj = lane & 15, rh = lane >> 4;             // column walk: lane = (row parity, column)
    const NetDims d = a.d;
    int it = 0;
    for (int pass = 0; pass < 2; pass++) {
      const int g = 1 - pass;
      const float* b2p = a.headw + (size_t)(g * 2 + 0) * N;   // b2 * TANH_K
      const float* w3p = a.headw + (size_t)(g * 2 + 1) * N;
      const float b3 = __ldg(a.params + g * d.SIZE + d.ob3);
      float gb2acc[4] = {0.f, 0.f, 0.f, 0.f}, gW3acc[4] = {0.f, 0.f, 0.f, 0.f};
      float gb3 = 0.f, lsum = 0.f, csum = 0.f;            // booked by the cq == 0 warps only
      for (int t = pair; t < tiles; t += n_pairs, it++) {
        const int buf = it & 1;
        const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
        const int s = t * 2 * BM + (int)rank * BM + r;    // sample of this chunk
        const bool valid = s < a.count;
        float R = 0.f, act = 0.f, oldlp = 0.f, V = 0.f;
        if (valid) {                                      // in flight while the tile still multiplies
          const int64_t p0 = a.first + s;
          const int64_t p = a.idx ? (int64_t)a.idx[p0] : p0;
          R = a.tr.ret[p];
          if (g == 0) { act = a.tr.action[p]; oldlp = a.tr.logp[p]; V = a.vtmp[(int64_t)cq * a.cap + s]; }
        }
        if (a.dbg && threadIdx.x == 64) a.dbg[blockIdx.x * 16 + it * 4 + 0] = clock64();
        tc::mbar_wait(&tfull[buf], bphase);
        tc::tc_fence_after();
        if (a.dbg && threadIdx.x == 64) a.dbg[blockIdx.x * 16 + it * 4 + 1] = clock64();
        if (threadIdx.x == 64) tc::tl_stamp(a.tl, 1, 2 + pass);   // accumulator of pass 0 / 1 ready
        const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 256u;
        // ---- pass 1 over the row: activation, dot with W3, h back into TMEM
        float2 yab = make_float2(0.f, 0.f);              // two chains of the row's dot product, one FFMA2 per pair
#pragma unroll
        for (int ci = 0; ci < 4; ci++) {
          const int c = c_begin + ci * 16;
          if (c < c_end) {
            float v[16];
            float4 bb[4], ww[4];
            tc::tmem_ld16(t0 + (uint32_t)c, v);
#pragma unroll
            for (int i = 0; i < 4; i++) {
              bb[i] = __ldg(reinterpret_cast<const float4*>(b2p + c) + i);
              ww[i] = __ldg(reinterpret_cast<const float4*>(w3p + c) + i);
            }
            tc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 4; i++) {
              float h0, h1, h2, h3;
              // (rows of the padding need no masking: their inputs are finite -- zero-filled or stale plane rows -- and their
              //  dz is exactly 0, so they add nothing to any sum and their D2 rows are written as zeros)
              const float2 kk = make_float2(TANH_K, TANH_K);
              const float2 t01 = __ffma2_rn(make_float2(v[4 * i], v[4 * i + 1]), kk, make_float2(bb[i].x, bb[i].y));
              const float2 t23 = __ffma2_rn(make_float2(v[4 * i + 2], v[4 * i + 3]), kk, make_float2(bb[i].z, bb[i].w));
              tanh4(t01.x, t01.y, t23.x, t23.y, h0, h1, h2, h3);
              yab = __ffma2_rn(make_float2(ww[i].x, ww[i].y), make_float2(h0, h1), yab);
              yab = __ffma2_rn(make_float2(ww[i].z, ww[i].w), make_float2(h2, h3), yab);
              v[4 * i] = h0; v[4 * i + 1] = h1; v[4 * i + 2] = h2; v[4 * i + 3] = h3;
            }
            tc::tmem_st16(t0 + (uint32_t)c, v);
          }
        }
        // the four warps of this row quarter add their partial dot products through the staging buffers
        if (lane == 0) tc::bulk_wait_read<0>();           // (the last store of the previous tile has read this buffer)
        __syncwarp();
        reinterpret_cast<float*>(sb)[lane] = yab.x + yab.y;
        tc::tmem_st_wait();
        bar_sync(1 + q, 128);
        float ytot = b3;
        {
          const uint8_t* s0 = sStg + ((warp - 2) & 3) * STG_WARP;   // staging of the cq = 0 warp of this row quarter
          float part[4];
#pragma unroll
          for (int k = 0; k < 4; k++) part[k] = reinterpret_cast<const float*>(s0 + (size_t)k * 4 * STG_WARP)[lane];
          ytot += (part[0] + part[1]) + (part[2] + part[3]);
        }
        bar_sync(1 + q, 128);                             // all four have read: the buffers are staging again
        // ---- loss head (same arithmetic as k_wide_head)
        float dz = 0.f;
        if (g == 1) {                                     // critic: no final activation (ppo_nn.h:118-122)
          const float e = ytot - R;
          if (valid) {
            dz = e * a.inv_denom;                         // d(0.5 mean e^2)/dV  (:288)
            a.vtmp[(int64_t)cq * a.cap + s] = ytot;
            if (cq == 0) { lsum = fmaf(e, e, lsum); gb3 += dz; }
          }
        } else if (valid) {
          const float mu = act_tanh(ytot);                                              // ppo_nn.h:102
          const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);    // :140
          const float ratio = expf(lp - oldlp);                                         // :263
          const float A = R - V;                                                        // :274 (detached)
          const float surr1 = ratio * A;                                                // :284
          const float rc = fminf(fmaxf(ratio, a.hp.clip_lo), a.hp.clip_hi);
          const float surr2 = rc * A;                                                   // :285
          const bool inside = (ratio >= a.hp.clip_lo) && (ratio <= a.hp.clip_hi);
          float gr;                                                                     // autograd of min(): ties split the gradient
          if (surr1 < surr2) gr = A;
          else if (surr1 > surr2) gr = inside ? A : 0.f;
          else gr = 0.5f * A + (inside ? 0.5f * A : 0.f);
          const float dlogp = -gr * ratio * a.inv_denom;
          const float dmu = dlogp * (act - mu) * a.hp.inv_var;                          // :20-31
          dz = dmu * (1.0f - mu * mu);
          if (cq == 0) {
            lsum -= fminf(surr1, surr2);                                                // :287
            csum += inside ? 0.f : 1.f;
            gb3 += dz;
          }
        }
        if (a.dbg && threadIdx.x == 64) a.dbg[blockIdx.x * 16 + it * 4 + 2] = clock64();
        // ---- pass 2 over the row: D2 -> planes, column sums.  One transposed tile (h, fp32, through the staging buffer)
        // serves both sums: lane (rh, cj) walks column cj over the rows of parity rh with the rows' dz by shuffle,
        //   dW3[j] += sum_r dz_r h_rj        db2[j] += w3_j sum_r dz_r (1 - h_rj^2)   (the w3_j factor at the end of the pass)
        const int row0 = t * 2 * BM + (int)rank * BM + q * 32;
        const uint8_t* colb[4];                           // column cj of rows 2 rr + rh, one base per rr & 3 (the swizzle)
#pragma unroll
        for (int k = 0; k < 4; k++) colb[k] = sb + rh * 64 + ((((cj >> 2) ^ k)) << 4) + (cj & 3) * 4;
#pragma unroll
        for (int ci = 0; ci < 4; ci++) {
          const int c = c_begin + ci * 16;
          if (c < c_end) {
            float h[16];
            float4 ww[4];
            tc::tmem_ld16(t0 + (uint32_t)c, h);
#pragma unroll
            for (int i = 0; i < 4; i++) ww[i] = __ldg(reinterpret_cast<const float4*>(w3p + c) + i);
            tc::tmem_ld_wait();
            if (lane == 0) tc::bulk_wait_read<0>();       // the previous chunk's store has read the staging buffer
            __syncwarp();
            // fp32 tile: 32 rows of 64 B, 16-B chunk i of row r at i ^ ((r >> 1) & 3)
#pragma unroll
            for (int i = 0; i < 4; i++)
              *reinterpret_cast<float4*>(sb + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4)) = make_float4(h[4 * i], h[4 * i + 1], h[4 * i + 2], h[4 * i + 3]);
            __syncwarp();
            float2 s12 = make_float2(0.f, 0.f);           // (sum dz h, sum dz (1 - h^2)): one FFMA2 per row step
#pragma unroll
            for (int rr = 0; rr < 16; rr++) {             // this lane's row of the step: 2 rr + rh
              const float hv = *reinterpret_cast<const float*>(colb[rr & 3] + rr * 128);
              const float dr = __shfl_sync(0xffffffffu, dz, 2 * rr + rh);
              s12 = __ffma2_rn(make_float2(dr, dr), make_float2(hv, fmaf(-hv, hv, 1.0f)), s12);
            }
            float s1 = s12.x, s2 = s12.y;
            s1 += __shfl_xor_sync(0xffffffffu, s1, 16);   // even rows + odd rows (the same sum in both halves of the warp)
            s2 += __shfl_xor_sync(0xffffffffu, s2, 16);
            gW3acc[ci] += s1;
            gb2acc[ci] += s2;
            // D2 in place of h, then the (hi | mid) bf16 tiles: 32 rows of 32 B per plane, 16-B chunk j of row r at j ^ ((r >> 2) & 1)
#pragma unroll
            for (int i = 0; i < 4; i++) {                 // dz w3 (1 - h^2), two columns per instruction
              const float2 dz2 = make_float2(dz, dz), one = make_float2(1.0f, 1.0f);
              const float2 h01 = make_float2(h[4 * i], h[4 * i + 1]), h23 = make_float2(h[4 * i + 2], h[4 * i + 3]);
              const float2 d01 = __fmul2_rn(__fmul2_rn(dz2, make_float2(ww[i].x, ww[i].y)), __ffma2_rn(make_float2(-h01.x, -h01.y), h01, one));
              const float2 d23 = __fmul2_rn(__fmul2_rn(dz2, make_float2(ww[i].z, ww[i].w)), __ffma2_rn(make_float2(-h23.x, -h23.y), h23, one));
              h[4 * i] = d01.x; h[4 * i + 1] = d01.y; h[4 * i + 2] = d23.x; h[4 * i + 3] = d23.y;
            }
            __syncwarp();                                 // every lane has read its column of the h tile
#pragma unroll
            for (int j = 0; j < 2; j++) {
              uint4 vh, vm;
              split2(h[8 * j], h[8 * j + 1], vh.x, vm.x);
              split2(h[8 * j + 2], h[8 * j + 3], vh.y, vm.y);
              split2(h[8 * j + 4], h[8 * j + 5], vh.z, vm.z);
              split2(h[8 * j + 6], h[8 * j + 7], vh.w, vm.w);
              const int off = lane * 32 + ((j ^ ((lane >> 2) & 1)) << 4);
              *reinterpret_cast<uint4*>(sb + off) = vh;
              *reinterpret_cast<uint4*>(sb + 1024 + off) = vm;
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {                              // one store: box {16 columns, 32 rows, 2 planes}
              tc::tma_store_3d(&tmD, sb, c, row0, g * 2);
              tc::bulk_commit();
            }
          }
        }
        tc::tc_fence_before();
        tc::mbar_arrive_leader_cta(&tempty[buf]);
        if (a.dbg && threadIdx.x == 64) a.dbg[blockIdx.x * 16 + it * 4 + 3] = clock64();
      }
      // ---- end of this net's pass: fold the four row quarters in order, write this CTA's partial row
      if (a.dx) {                                         // the dX pass of this net reads the D2 tiles back: they must have LANDED
        if (lane == 0) { tc::bulk_wait_all(); tc::fence_proxy_async(); tc::mbar_arrive(&d2done[g]); }
      } else if (lane == 0) {
        tc::bulk_wait_read<0>();                          // staging is about to be reused as plain memory ...
      }
      __syncwarp();
      bar_sync(5, HEAD_WARPS * 32);                       // ... by every head warp, across each other's buffers
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        gb3 += __shfl_xor_sync(0xffffffffu, gb3, o);
        lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
        csum += __shfl_xor_sync(0xffffffffu, csum, o);
      }
      float* fold = reinterpret_cast<float*>(sStg);       // [16 warps][2][4 chunks][16 columns], then [4][4] scalars
      float* fw = fold + (warp - 2) * 128;
      if (rh == 0) {
#pragma unroll
        for (int ci = 0; ci < 4; ci++) { fw[ci * 16 + cj] = gb2acc[ci]; fw[64 + ci * 16 + cj] = gW3acc[ci]; }
      }
      if (cq == 0 && lane == 0) { float* sc = fold + HEAD_WARPS * 128 + ((warp - 2) & 3) * 4; sc[0] = gb3; sc[1] = lsum; sc[2] = csum; }
      bar_sync(5, HEAD_WARPS * 32);
      // the CTA's partial row, or (a.head_sum) its share of the ONE row the merged layer-1 + Adam kernel consumes: red.add in
      // the L2, issued with most of the kernel still ahead (the critic's in the middle, the actor's before the dX passes)
      float* out = a.head_sum ? a.head_sum + (size_t)g * a.head_row : a.part_head + ((size_t)blockIdx.x * 2 + g) * a.head_row;
      const bool addrow = a.head_sum != nullptr;
      auto put = [&](int i, float v) { if (addrow) atomicAdd(out + i, v); else out[i] = v; };
      if (((warp - 2) & 3) == 0) {                        // one warp per column quarter: lanes = (quantity, column)
#pragma unroll
        for (int ci = 0; ci < 4; ci++) {
          const int c = c_begin + ci * 16;
          if (c < c_end) {
            float sv = 0.f;
#pragma unroll
            for (int qq = 0; qq < 4; qq++) sv += fold[(cq * 4 + qq) * 128 + rh * 64 + ci * 16 + cj];
            if (rh == 0) put(N + c + cj, __ldg(w3p + c + cj) * sv);   // [H2, 2 H2) db2 = w3_j sum_s dz_s (1 - h_sj^2)
            else put(c + cj, sv);                                     // [0, H2) dW3
          }
        }
        if (cq == 0 && lane == 0) {
          float t3 = 0.f, tl = 0.f, tcn = 0.f;
#pragma unroll
          for (int qq = 0; qq < 4; qq++) { const float* sc = fold + HEAD_WARPS * 128 + qq * 4; t3 += sc[0]; tl += sc[1]; tcn += sc[2]; }
          put(2 * N, t3);
          put(2 * N + 1, tl);                              // g == 0: loss1 sum, g == 1: sum e^2
          if (g == 0) put(2 * N + 2, tcn);
          else if (!addrow) out[2 * N + 2] = 0.f;
          if (!addrow) out[2 * N + 3] = 0.f;
        }
      }
      bar_sync(5, HEAD_WARPS * 32);                       // fold is staging again
      if (threadIdx.x == 64) tc::tl_stamp(a.tl, 1, 4 + pass);   // head of pass 0 / 1 done
    }
    // ---- passes 2, 3 (a.dx): raw epilogue of the dX GEMM, Z1 tiles -> d1 (fp32), 16 columns at a time through the same staging
    // buffers: TMEM -> registers -> 64-B-swizzled 32 x 16 tile -> TMA store
    if (a.dx) {
      const int c2_begin = cq * (N2 / 4), c2_end = c2_begin + N2 / 4;
      for (int pass = 2; pass < 4; pass++) {
        const int g = (pass & 1) ? 0 : 1;
        for (int t = pair; t < tiles; t += n_pairs, it++) {
          const int buf = it & 1;
          const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
          tc::mbar_wait(&tfull[buf], bphase);
          tc::tc_fence_after();
          if (threadIdx.x == 64 && pass == 3) tc::tl_stamp(a.tl, 1, 7);   // accumulator of the last dX pass ready
          const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 256u;
          const int row0 = t * 2 * BM + (int)rank * BM + q * 32;
          // the LAST accumulator of the kernel: no load is outstanding and tfull says every MMA has read its operands, so the
          // weight / A-stage memory is free -- each chunk gets a staging tile of its own there and none waits for the TMA to
          // have read the chunk before it (this epilogue is the exposed tail of the kernel)
          const bool wide = pass == 3 && t + n_pairs >= tiles && (size_t)B_BYTES + (size_t)stages * A_STAGE >= (size_t)HEAD_WARPS * 4 * STG_WARP;
          if (wide && lane == 0) tc::bulk_wait_read<0>();
#pragma unroll
          for (int ci = 0; ci < 4; ci++) {
            const int c = c2_begin + ci * 16;
            if (c < c2_end) {
              uint8_t* stg = wide ? smem + (size_t)((warp - 2) * 4 + ci) * STG_WARP : sb;
              float v[16];
              tc::tmem_ld16(t0 + (uint32_t)c, v);
              tc::tmem_ld_wait();
              if (!wide) {
                if (lane == 0) tc::bulk_wait_read<0>();
                __syncwarp();
              }
#pragma unroll
              for (int i = 0; i < 4; i++)
                *reinterpret_cast<float4*>(stg + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
              tc::fence_proxy_async();
              __syncwarp();
              if (lane == 0) {
                tc::tma_store_3d(&tmC, stg, c, row0, g);
                tc::bulk_commit();
              }
            }
          }
          tc::tc_fence_before();
          tc::mbar_arrive_leader_cta(&tempty[buf]);
        }
      }
    }
    if (lane == 0) tc::bulk_wait_read<0>();               // every store has READ its staging tile; the writes complete with the grid
                                                          // (the next kernel's griddepcontrol.wait waits for this grid's memory flush)
  }

  if (a.dbg && threadIdx.x == 64) a.dbg[blockIdx.x * 16 + 13] = clock64();   // head done
  if (threadIdx.x == 64) tc::tl_stamp(a.tl, 1, 6);
  tc::tc_fence_before();
  tc::cluster_sync();                                     // no CTA of the pair exits while the other may still signal it
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc_2sm(tmem_base, 512);
  }
}

}  // namespace

int fh_stages(int N, int K) {
  const int b_bytes = (K / BK) * 2 * (N / 2) * 128;
  int s = (SMEM_LIMIT - 1024 - BAR_BYTES - YBUF_BYTES - STG_BYTES - b_bytes) / A_STAGE;
  return s > 4 ? 4 : s;
}

bool fh_supported(int N, int K) {
  if (N % 64 || K % 64 || N > 256 || K > 256 || N < 64 || K < 64) return false;
  return fh_stages(N, K) >= 2;
}

cudaError_t launch_wide_fh(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD, const CUtensorMap& tmA2,
                           const CUtensorMap& tmB2, const CUtensorMap& tmC, const FhArgs& a, int sm_count, int* grid_out, cudaStream_t st) {
  static b2::PerDeviceOnce once;   // function attributes are per device
  if (once.need()) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_wide_fh, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)) != cudaSuccess) return e;
    once.done();
  }
  const int tiles = (a.m_pad + 2 * BM - 1) / (2 * BM);
  int pairs = sm_count / 2;
  if (pairs > tiles) pairs = tiles;
  if (pairs < 1) return cudaErrorInvalidValue;
  const int smem = (a.K / BK) * 2 * (a.N / 2) * 128 + a.stages * A_STAGE + STG_BYTES + YBUF_BYTES + BAR_BYTES + 1024;
  const int smem_req = smem > 120 * 1024 ? smem : 120 * 1024;   // one CTA per SM (512 TMEM columns)
  *grid_out = 2 * pairs;
  return launch_pdl(k_wide_fh, dim3(2 * pairs), dim3(FH_THREADS), (size_t)smem_req, st, tmA, tmB, tmD, tmA2, tmB2, tmC, a);
}

}  // namespace b2
