// wide_gemm_mn2.cu -- k_gemm_mn2: the sample contraction dW2 = D2^T . H1 of a wide-net step on CTA PAIRS.
//
//   C[g][256][N] += A[g][K][256]^T . B[g][K][N]      K = samples (split over pairs), both operands MN-major: the
//                                                     sample-major (hi, mid) bf16 planes d2p / h1p, read as they lie
//
// Why a pair.  k_gemm_x3<MN> gives every CTA one 128-row half of dW2: per 32-sample K block it streams 16 KB of D2 and ALL
// of H1's block (32 KB), so H1 is fetched twice per step -- 100 MB of L2 -> SM traffic for 67 MB of operands, 12 us at the
// ~8 TB/s the L2 delivers to 148 SMs -- and the shared-memory port carries TMA fill plus operand reads at its whole
// 128 B/clk.  With tcgen05.mma.cta_group::2 (M = 256) the two CTAs of a pair each hold their 128 rows of
// A and HALF of B's columns: 32 KB per K block and CTA instead of 48, every operand byte fetched once.
// The faster GEMM only pays once the dW1 warps beside it stop setting the kernel's length: they run dw1_worker_async
// (wide_dw1.cuh), Z1 through a cp.async ring in the shared memory the smaller stages leave free.
//
// Per CTA, 320 + 512 threads: warp 0 = TMA producer (a lane per 4-KB box), warp 1 = TMEM allocation and, in the leader,
// the MMA issuer; warps 2-9 = epilogue of this CTA's 128 accumulator rows (TMA reduce-add of the split-K tile into the
// net's plane, or a store to the split's own plane); warps 10-25 = the dW1 / db1 work.
// Barriers as in k_gemm2_ws: `full`, `tempty` live in the leader (both CTAs' TMA bytes / epilogue arrivals land there),
// `empty`, `tfull` in each CTA (tcgen05.commit multicasts).
#include <algorithm>

#include "tc_ptx.cuh"
#include "wide.h"
#include "mlp.cuh"
#include "wide_dw1.cuh"

namespace b2 {

namespace {

constexpr int BM = 128;
constexpr int BKS = 32;                       // samples per K block
constexpr int MNB_BYTES = BKS * 128;          // one 64-feature x 32-sample box (128-B swizzle)
constexpr int GEMM_THREADS = 320;
constexpr int STG_BUF = 2048, STG_BYTES = 8 * STG_BUF;   // a 32 x 16 fp32 tile per epilogue warp: the ring of the dW1 warps needs the room
constexpr int BAR_BYTES = 256;
constexpr int SMEM_LIMIT = 227 * 1024;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS + DW1_THREADS, 1)
k_gemm_mn2(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
           const __grid_constant__ CUtensorMap tmC, const GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = tc::cluster_ctarank();            // 0 = leader
  const int N = a.N, NHB = N / 128, stages = a.stages;    // NHB: 64-feature boxes of this CTA's half of B
  const int A_PLANE = 2 * MNB_BYTES;                      // 128 rows of A = two boxes per plane
  const int B_PLANE = NHB * MNB_BYTES;
  const int SB = 2 * A_PLANE + 2 * B_PLANE;               // stage bytes of ONE CTA
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * SB + STG_BYTES);
  uint64_t* empty = full + 8;
  uint64_t* tfull = empty + 8;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int kb_total = a.K / BKS;
  const int n_pairs = gridDim.x / 2, pair = blockIdx.x / 2;
  const int n_items = a.G * a.splits;                     // item = g * splits + split: the pair owns all 256 output rows

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    tc::tma_prefetch_desc(&tmC);
    for (int s = 0; s < stages; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; b++) { tc::mbar_init(&tfull[b], 1); tc::mbar_init(&tempty[b], 512); }
    tc::fence_barrier_init();
  }
  if (threadIdx.x == 0) tc::tl_stamp(a.tl, 2, 0);
  pdl_launch_dependents();
  if (warp == 1) tc::tmem_alloc_2sm(tmem_slot, 512);
  pdl_wait();                                             // operands come from the previous kernels of the step
  if (threadIdx.x == 0) tc::tl_stamp(a.tl, 2, 1);
  tc::tc_fence_before();
  tc::cluster_sync();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    int stage = 0; uint32_t phase = 0;
    const uint64_t pol_last = tc::l2_policy_evict_first();
    for (int item = pair; item < n_items; item += n_pairs) {
      const int sp = item % a.splits, g = item / a.splits;
      const int kb0 = sp * a.kb_per_split, kb1 = min(kb0 + a.kb_per_split, kb_total);
      for (int kb = kb0; kb < kb1; kb++) {
        tc::mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* s = smem + (size_t)stage * SB;
        if (lane == 0 && rank == 0) tc::mbar_arrive_expect_tx(&full[stage], (uint32_t)(2 * SB));   // both CTAs' bytes
        __syncwarp();
        const int per = 2 + NHB;                          // boxes per plane: this CTA's A feature blocks, then its B blocks
        if (lane < 2 * per) {
          const int pl = lane / per, j = lane % per;
          if (j < 2)
            tc::tma_load_3d_2sm_hint(s + pl * A_PLANE + j * MNB_BYTES, &tmA, &full[stage], ((int)rank * 2 + j) * 64, kb * BKS, g * 2 + pl, pol_last);   // D2: last read of the step
          else
            tc::tma_load_3d_2sm(s + 2 * A_PLANE + pl * B_PLANE + (j - 2) * MNB_BYTES, &tmB, &full[stage],
                                ((int)rank * NHB + (j - 2)) * 64, kb * BKS, g * 2 + pl);
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      const uint32_t idesc = tc::idesc_bf16_f32(2 * BM, N) | (1u << 15) | (1u << 16);   // A and B MN-major
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int item = pair; item < n_items; item += n_pairs, it++) {
        const int sp = item % a.splits;
        const int kb0 = sp * a.kb_per_split, kb1 = min(kb0 + a.kb_per_split, kb_total);
        const int buf = it & 1;
        const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
        tc::mbar_wait(&tempty[buf], bphase ^ 1);
        tc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * 256u;
        uint32_t acc = 0;
        for (int kb = kb0; kb < kb1; kb++) {
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          if (kb == kb0) tc::tl_stamp(a.tl, 2, 2);         // first operands have landed
          if (kb == kb1 - 1) tc::tl_stamp(a.tl, 2, 3);     // last operands have landed
          const uint32_t s = tc::smem_u32(smem + (size_t)stage * SB);
          const uint64_t mAh = tc::smem_desc_mn_sw128(s, MNB_BYTES), mAm = tc::smem_desc_mn_sw128(s + A_PLANE, MNB_BYTES);
          const uint64_t mBh = tc::smem_desc_mn_sw128(s + 2 * A_PLANE, MNB_BYTES);
          const uint64_t mBm = tc::smem_desc_mn_sw128(s + 2 * A_PLANE + B_PLANE, MNB_BYTES);
#pragma unroll
          for (int k = 0; k < BKS / 16; k++) {
            const uint64_t off = (uint64_t)((k * 2048) >> 4);   // 16 samples = two 8-row groups of 1024 B
            tc::umma_bf16_2sm(d_tmem, mAm + off, mBh + off, idesc, acc);   // small terms first
            tc::umma_bf16_2sm(d_tmem, mAh + off, mBm + off, idesc, 1u);
            tc::umma_bf16_2sm(d_tmem, mAh + off, mBh + off, idesc, 1u);
            acc = 1u;
          }
          tc::umma_commit_2sm(&empty[stage], 3);
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit_2sm(&tfull[buf], 3);
      }
    }
    __syncwarp();
  } else if (warp >= GEMM_THREADS / 32) {
    if (a.dw1.params)
      dw1_worker_async(a.dw1, reinterpret_cast<float*>(smem + (size_t)stages * SB + STG_BYTES + BAR_BYTES), threadIdx.x - GEMM_THREADS);
    if (threadIdx.x == GEMM_THREADS) tc::tl_stamp(a.tl, 2, 6);   // dW1 warps done
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int n_chunks = N / 32;
    const int c_begin = half ? ((n_chunks + 1) / 2) * 32 : 0;
    const int c_end = half ? N : ((n_chunks + 1) / 2) * 32;
    uint8_t* sb = smem + (size_t)stages * SB + (warp - 2) * STG_BUF;
    // The pair's LAST item leaves the operand stages idle (every load has landed and tfull says every MMA has read them): its
    // epilogue then gives each 16-column chunk a staging tile of its own there, so no chunk waits for the TMA to have read the
    // one before it -- the drain is the exposed tail of the kernel (one item per pair at 2x256 on 148 SMs)
    const bool wide_fits = (size_t)stages * SB >= 8 * (size_t)(((n_chunks + 1) / 2) * 2) * STG_BUF;
    uint8_t* sb_wide = smem + (size_t)(warp - 2) * (size_t)(((n_chunks + 1) / 2) * 2) * STG_BUF;
    int it = 0;
    for (int item = pair; item < n_items; item += n_pairs, it++) {
      const int sp = item % a.splits, g = item / a.splits;
      const int buf = it & 1;
      const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
      const bool wide = wide_fits && item + n_pairs >= n_items;
      tc::mbar_wait(&tfull[buf], bphase);
      tc::tc_fence_after();
      if (threadIdx.x == 64) tc::tl_stamp(a.tl, 2, 4);    // accumulator ready
      const int row0 = (int)rank * BM + q * 32;           // first output row (feature j of D2) of this warp
      const int plane = a.reduce_add ? g : g * a.splits + sp;
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 256u;
      if (wide && lane == 0) tc::bulk_wait_read<0>();     // (the small tile of an earlier item)
      for (int c = c_begin; c < c_end; c += 16) {       // tmC: 16-column boxes, 64-B swizzle (make_output_map16)
        uint8_t* stg = wide ? sb_wide + (size_t)((c - c_begin) / 16) * STG_BUF : sb;
        float v[16];
        tc::tmem_ld16(t0 + (uint32_t)c, v);
        tc::tmem_ld_wait();
        if (!wide) {
          if (lane == 0) tc::bulk_wait_read<0>();
          __syncwarp();
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
          *reinterpret_cast<float4*>(stg + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4)) =
              make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (a.reduce_add) tc::tma_reduce_add_3d(&tmC, stg, c, row0, plane);
          else tc::tma_store_3d(&tmC, stg, c, row0, plane);
          tc::bulk_commit();
        }
      }
      tc::tc_fence_before();
      tc::mbar_arrive_leader_cta(&tempty[buf]);
    }
    if (lane == 0) tc::bulk_wait_all();
    if (threadIdx.x == 64) tc::tl_stamp(a.tl, 2, 5);      // epilogue done
  }

  tc::tc_fence_before();
  tc::cluster_sync();                                     // no CTA of the pair exits while the other may still signal it
  if (threadIdx.x == 0) tc::tl_stamp(a.tl, 2, 7);
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc_2sm(tmem_base, 512);
  }
}

}  // namespace

int gemm_mn2_stages(int N) {
  const int sb = (4 + 2 * (N / 128)) * MNB_BYTES;
  int s = (SMEM_LIMIT - 1024 - BAR_BYTES - STG_BYTES - DW1_SMEM_ASYNC) / sb;
  return s > 6 ? 6 : s;
}

bool gemm_mn2_supported(int M, int N) { return M == 256 && N % 128 == 0 && N >= 128 && N <= 256 && gemm_mn2_stages(N) >= 2; }

// grid = 2 * pairs; a.m_tiles is ignored (the pair owns both 128-row halves), a.fuse_m / a.red are not supported here
cudaError_t launch_gemm_mn2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmArgs& a, int sm_count,
                            int* grid_out, cudaStream_t st) {
  static b2::PerDeviceOnce once;   // function attributes are per device
  if (once.need()) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_gemm_mn2, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)) != cudaSuccess) return e;
    once.done();
  }
  const int n_items = a.G * a.splits;
  int pairs = sm_count / 2;
  if (pairs > n_items) pairs = n_items;
  if (pairs < 1) return cudaErrorInvalidValue;
  const int smem = a.stages * (4 + 2 * (a.N / 128)) * MNB_BYTES + STG_BYTES + BAR_BYTES + DW1_SMEM_ASYNC + 1024;
  const int smem_req = smem > 120 * 1024 ? smem : 120 * 1024;   // one CTA per SM (512 TMEM columns)
  *grid_out = 2 * pairs;
  return launch_pdl(k_gemm_mn2, dim3(2 * pairs), dim3(GEMM_THREADS + DW1_THREADS), (size_t)smem_req, st, tmA, tmB, tmC, a);
}

}  // namespace b2
