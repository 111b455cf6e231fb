// wide_gemm2.cu -- k_gemm2_ws: the layer-2 forward / dX GEMMs on CTA PAIRS with the weights resident.
//
//   C[g][M][N] = A[g][M][K] . B[g]      B = W2 of net g, either K-major rows [N][K] (Z2 = H1 W2^T) or as stored
//                                        contraction-major [K][N] (Z1 = D2 W2, BMN = true)
//
// Why: k_gemm_x3 streams A (128 KB) AND the whole weight matrix (256 KB as a bf16 pair) from L2 for every 128-row tile.
// At 131,072-sample minibatches that is 7.4 TB/s of L2->SM traffic and the main loop runs at 0.4 of the MMA rate
// (profiles/r1f_*, gemm_timeline.py).  Here two CTAs of one cluster (one TPC) share a 256-row tile with
// tcgen05.mma.cta_group::2: each CTA keeps HALF of the weight rows (N/2 x K, hi + mid planes = 128 KB at 256 x 256)
// in shared memory for the whole kernel and streams only its own 128 rows of A.  L2 traffic per tile drops 3x, the
// shared-memory port sees 64 B/clk of TMA fill + operand reads instead of 160, and the pair issues M = 256 MMAs.
//
// Per CTA, 320 threads:  warp 0 = TMA producer (weights once, then A stages), warp 1 = TMEM allocation (both CTAs) and,
// in the leader CTA only, the MMA issuer; warps 2-9 = epilogue of this CTA's 128 accumulator rows (as k_gemm_x3).
// Barriers live at the same offset in both CTAs: `full`, `bfull`, `tempty` are used in the leader's copy (TMA bytes of
// both CTAs land there through the cta_group::2 load form; the peer's epilogue arrives remotely), `empty` and `tfull`
// in each CTA's own copy (tcgen05.commit multicasts to both).
// A pair works on ONE net (pairs are split between the nets) so the resident weights are loaded once per launch.
#include <stdio.h>

#include <algorithm>

#include "tc_ptx.cuh"
#include "wide.h"

namespace b2 {

namespace {

constexpr int BM = 128, BK = 64;
constexpr int THREADS = 320;
constexpr int A_PLANE = BM * BK * 2;           // 16 KB
constexpr int A_STAGE = 2 * A_PLANE;           // hi + mid
constexpr int STG_BUF = 4096, STG_BYTES = 8 * STG_BUF;
constexpr int BAR_BYTES = 256;
constexpr int SMEM_LIMIT = 227 * 1024;

template <bool BMN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
k_gemm2_ws(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
           const __grid_constant__ CUtensorMap tmC, const Gemm2Args a) {
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
  uint64_t* full = reinterpret_cast<uint64_t*>(sStg + STG_BYTES);
  uint64_t* empty = full + 4;
  uint64_t* tfull = empty + 4;
  uint64_t* tempty = tfull + 2;
  uint64_t* bfull = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bfull + 4);   // bfull[kb]: one barrier per K block of the weights (K <= 256)

  // schedule: the pairs are split evenly between the G groups; a pair walks the 256-row tiles of its group
  const int n_pairs = gridDim.x / 2, pair = blockIdx.x / 2;
  const int ppg = n_pairs / a.G;                          // pairs per group (host guarantees n_pairs % G == 0, ppg >= 1)
  const int g = pair / ppg, p_local = pair % ppg;
  const int tiles = (a.M + 2 * BM - 1) / (2 * BM);

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    tc::tma_prefetch_desc(&tmC);
    for (int s = 0; s < stages; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; b++) { tc::mbar_init(&tfull[b], 1); tc::mbar_init(&tempty[b], 512); }
    for (int kb = 0; kb < 4; kb++) tc::mbar_init(&bfull[kb], 1);
    tc::fence_barrier_init();
  }
  pdl_launch_dependents();
  if (warp == 1) tc::tmem_alloc_2sm(tmem_slot, 512);
  pdl_wait();
  tc::tc_fence_before();
  tc::cluster_sync();                                     // barriers of both CTAs initialised before any remote use
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ---- resident weights: this CTA's N/2 rows (or columns), hi and mid, one barrier per K block, issued in front of
      // the first tile's A stage of the same K block: the first MMA waits for a quarter of the weights, not all 128 KB
      auto load_weights = [&](int kb) {
        if (rank == 0) tc::mbar_arrive_expect_tx(&bfull[kb], (uint32_t)(4 * B_KBLOCK));   // 2 planes x 2 CTAs
        for (int pl = 0; pl < 2; pl++) {
          uint8_t* dst = sB + (size_t)(kb * 2 + pl) * B_KBLOCK;
          if (BMN) {
            for (int fb = 0; fb < NH / 64; fb++)
              tc::tma_load_3d_2sm(dst + fb * (BK * 128), &tmB, &bfull[kb], (int)rank * NH + fb * 64, kb * BK, g * 2 + pl);
          } else {
            tc::tma_load_3d_2sm(dst, &tmB, &bfull[kb], kb * BK, (int)rank * NH, g * 2 + pl);
          }
        }
      };
      // ---- A: this CTA's 128 rows of every tile of the pair
      int stage = 0; uint32_t phase = 0;
      bool first_tile = true;
      for (int t = p_local; t < tiles; t += ppg, first_tile = false) {
        const int row0 = t * 2 * BM + (int)rank * BM;
        for (int kb = 0; kb < KB; kb++) {
          if (first_tile) load_weights(kb);
          tc::mbar_wait(&empty[stage], phase ^ 1);
          if (rank == 0) tc::mbar_arrive_expect_tx(&full[stage], (uint32_t)(2 * A_STAGE));
          uint8_t* s = sA + (size_t)stage * A_STAGE;
          tc::tma_load_3d_2sm(s, &tmA, &full[stage], kb * BK, row0, g * 2 + 0);
          tc::tma_load_3d_2sm(s + A_PLANE, &tmA, &full[stage], kb * BK, row0, g * 2 + 1);
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      const uint32_t idesc = tc::idesc_bf16_f32(2 * BM, N) | (BMN ? (1u << 16) : 0u);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int t = p_local; t < tiles; t += ppg, it++) {
        const int buf = it & 1;
        const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
        tc::mbar_wait(&tempty[buf], bphase ^ 1);           // both CTAs' epilogues have drained this accumulator
        tc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * 256u;
        uint32_t acc = 0;
        for (int kb = 0; kb < KB; kb++) {
          if (it == 0) tc::mbar_wait(&bfull[kb], 0);         // this K block of the resident weights has landed
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(sA + (size_t)stage * A_STAGE);
          const uint32_t sb = tc::smem_u32(sB + (size_t)(kb * 2) * B_KBLOCK);
          const uint64_t dAh = tc::smem_desc_sw128(sa), dAm = tc::smem_desc_sw128(sa + A_PLANE);
          const uint64_t dBh = BMN ? tc::smem_desc_mn_sw128(sb, BK * 128) : tc::smem_desc_sw128(sb);
          const uint64_t dBm = BMN ? tc::smem_desc_mn_sw128(sb + B_KBLOCK, BK * 128) : tc::smem_desc_sw128(sb + B_KBLOCK);
          const uint64_t bstep = BMN ? (2048 >> 4) : (32 >> 4);
#pragma unroll
          for (int k = 0; k < BK / 16; k++) {
            const uint64_t off = (uint64_t)((k * 32) >> 4), offb = (uint64_t)k * bstep;
            tc::umma_bf16_2sm(d_tmem, dAm + off, dBh + offb, idesc, acc);
            tc::umma_bf16_2sm(d_tmem, dAh + off, dBm + offb, idesc, 1u);
            tc::umma_bf16_2sm(d_tmem, dAh + off, dBh + offb, idesc, 1u);
            acc = 1u;
          }
          tc::umma_commit_2sm(&empty[stage], 3);            // both CTAs may refill this A stage
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit_2sm(&tfull[buf], 3);                // both epilogues may drain
      }
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int n_chunks = N / 32;
    const int c_begin = half ? ((n_chunks + 1) / 2) * 32 : 0;
    const int c_end = half ? N : ((n_chunks + 1) / 2) * 32;
    uint8_t* sb = sStg + (warp - 2) * STG_BUF;
    int it = 0;
    for (int t = p_local; t < tiles; t += ppg, it++) {
      const int buf = it & 1;
      const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
      tc::mbar_wait(&tfull[buf], bphase);
      tc::tc_fence_after();
      const int row0 = t * 2 * BM + (int)rank * BM + q * 32;
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 256u;
      for (int c = c_begin; c < c_end; c += 32) {
        float v[32];
        tc::tmem_ld32(t0 + (uint32_t)c, v);
        tc::tmem_ld_wait();
        if (lane == 0) tc::bulk_wait_read<0>();
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; i++)
          *reinterpret_cast<float4*>(sb + lane * 128 + ((i ^ (lane & 7)) << 4)) =
              make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tc::tma_store_3d(&tmC, sb, c, row0, g);
          tc::bulk_commit();
        }
      }
      tc::tc_fence_before();
      tc::mbar_arrive_leader(&tempty[buf]);
    }
    if (lane == 0) tc::bulk_wait_all();
  }

  tc::tc_fence_before();
  tc::cluster_sync();                                     // no CTA of the pair exits while the other may still signal it
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc_2sm(tmem_base, 512);
  }
}

}  // namespace

int gemm2_stages(int N, int K) {
  const int b_bytes = (K / BK) * 2 * (N / 2) * 128;
  int s = (SMEM_LIMIT - 1024 - BAR_BYTES - STG_BYTES - b_bytes) / A_STAGE;
  return s > 4 ? 4 : s;
}

bool gemm2_supported(int N, int K, bool bmn) {
  if (N % 64 || K % 64 || N > 256 || K > 256) return false;
  if (bmn && (N % 128)) return false;                     // each CTA's column half must be whole 64-wide boxes
  return gemm2_stages(N, K) >= 2;
}

cudaError_t launch_gemm2(bool bmn, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const Gemm2Args& a,
                         int sm_count, cudaStream_t st) {
  static b2::PerDeviceOnce once;   // function attributes are per device
  if (once.need()) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_gemm2_ws<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_gemm2_ws<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)) != cudaSuccess) return e;
    once.done();
  }
  const int tiles = (a.M + 2 * BM - 1) / (2 * BM);
  int ppg = (sm_count / 2) / a.G;                         // pairs per group
  if (ppg > tiles) ppg = tiles;
  if (ppg < 1) return cudaErrorInvalidValue;
  const int grid = 2 * ppg * a.G;
  const int smem = (a.K / BK) * 2 * (a.N / 2) * 128 + a.stages * A_STAGE + STG_BYTES + BAR_BYTES + 1024;
  const int smem_req = smem > 120 * 1024 ? smem : 120 * 1024;   // one CTA per SM (512 TMEM columns)
  if (bmn) return launch_pdl(k_gemm2_ws<true>, dim3(grid), dim3(THREADS), (size_t)smem_req, st, tmA, tmB, tmC, a);
  return launch_pdl(k_gemm2_ws<false>, dim3(grid), dim3(THREADS), (size_t)smem_req, st, tmA, tmB, tmC, a);
}

}  // namespace b2
