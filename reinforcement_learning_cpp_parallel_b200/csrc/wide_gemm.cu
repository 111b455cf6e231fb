// wide_gemm.cu -- k_gemm_x3: persistent, warp-specialised tcgen05 GEMM with bf16x3 operands.
//
//   C[g][M][N] = A[g][M][K] . B[g][N][K]^T        (both operands K-major, fp32 values stored as (hi, mid) bf16 planes)
//
// One CTA per SM, 320 threads:
//   warp 0  : TMA producer   -- one lane streams 64-wide K blocks of A_hi, A_mid (128 rows) and B_hi, B_mid (N rows)
//                               into a ring of shared-memory stages (cp.async.bulk.tensor, 128-B swizzle, mbarrier tx)
//   warp 1  : MMA issuer     -- one lane issues 4 (K steps of 16) x 3 (hi*hi, mid*hi, hi*mid) tcgen05.mma per stage,
//                               M = 128, N = whole width (<= 256), fp32 accumulator in TMEM; tcgen05.commit frees the
//                               stage and, after the last K block, publishes the accumulator
//   warps 2-9: epilogue      -- thread = accumulator row; tcgen05.ld 32 columns at a time, registers -> swizzled
//                               shared tile -> TMA store of the raw fp32 accumulator.  (A direct row-per-thread store
//                               costs the LSU 32 lines per instruction, and bias/tanh/tanh' in this epilogue were
//                               MUFU- and load-latency-bound at 1-2x the MMA time of a tile; both measured with the
//                               clock64 timeline of profiles/gemm_timeline.py.  The elementwise work now rides in the
//                               memory-bound consumers: k_wide_head applies bias + tanh, k_wide_dw1 applies tanh'.)
// TMEM holds two accumulators (2 x 256 columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
// Tiles are dealt statically: item = ((g * m_tiles) + mt) * splits + split, strided by gridDim.x.
// Roofline: tensor (bf16 dense); algorithmic flop = 2 M N K per group, issued flop = 3x that.
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <type_traits>
#include <vector>

#include "tc_ptx.cuh"
#include "wide.h"
#include "mlp.cuh"
#include "wide_reduce.cuh"
#include "wide_dw1.cuh"

namespace b2 {

namespace {

constexpr int BM = 128, BK = 64;
constexpr int GEMM_THREADS = 320;              // TMA warp, MMA warp, 8 epilogue warps
constexpr int A_PLANE_BYTES = BM * BK * 2;          // 16 KB
constexpr int SMEM_LIMIT = 227 * 1024;
constexpr int BAR_BYTES = 256;
constexpr int STG_BUF = 4096;                 // one 32 x 32 fp32 output tile
constexpr int STG_BYTES = 8 * STG_BUF;        // one staging tile per epilogue warp

__host__ __device__ constexpr int stage_bytes(int N) { return 2 * A_PLANE_BYTES + 2 * N * BK * 2; }

// MN = true: both operands are MN-major (the contraction index is the slow one in memory): C[M][N] = A[K][M]^T . B[K][N].
// Used for dW2 = D2^T H1, where K = samples and the activations are stored sample-major.  A K block is 32 samples; an
// operand block of 64 features x 32 samples is one TMA box (32 rows of 128 B, 128-B swizzle) = one UMMA MN-major
// swizzle atom column: 8 sample rows per 1024-B group (SBO), 64-feature blocks MNB_BYTES apart (LBO).
constexpr int BKS = 32;                       // samples per K block in MN mode
constexpr int MNB_BYTES = BKS * 128;          // one 64-feature x 32-sample block

template <bool MN>
__global__ void __launch_bounds__(MN ? GEMM_THREADS + DW1_THREADS : GEMM_THREADS, 1)
k_gemm_x3(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
          const __grid_constant__ CUtensorMap tmC, const GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  if (a.dbg && threadIdx.x == 0) a.dbg[(blockIdx.x * 4 + 1) * 8 + 7] = clock64();   // kernel entry
  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);   // 1024-B alignment for the 128-B swizzle atoms
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N = a.N, stages = a.stages;
  const bool FUSE = MN && a.fuse_m;                    // both 128-row output halves in this CTA (GemmArgs::fuse_m)
  const int AFB = FUSE ? 4 : 2;                        // 64-feature blocks of A per plane and stage (MN mode)
  const int SB = MN ? (2 * AFB + 2 * (N / 64)) * MNB_BYTES : stage_bytes(N);
  const int B_PLANE_BYTES = MN ? (N / 64) * MNB_BYTES : N * BK * 2;
  const int A_PLANE = MN ? AFB * MNB_BYTES : A_PLANE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * SB + STG_BYTES);
  uint64_t* empty = full + 8;
  uint64_t* tfull = empty + 8;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int kb_total = a.K / (MN ? BKS : BK);
  const int mt_n = FUSE ? 1 : a.m_tiles;               // output row tiles an item enumerates
  const int n_items = a.G * mt_n * a.splits;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    tc::tma_prefetch_desc(&tmC);
    for (int s = 0; s < stages; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; b++) { tc::mbar_init(&tfull[b], 1); tc::mbar_init(&tempty[b], 256); }
    tc::fence_barrier_init();
  }
  pdl_launch_dependents();
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  pdl_wait();                                           // operands come from the previous kernels of the step
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (a.dbg && threadIdx.x == 0) a.dbg[(blockIdx.x * 4 + 0) * 8 + 7] = clock64();

  if (warp == 0) {
    // K-major: one lane issues the 4 big boxes of a stage.  MN-major: a stage is 12-16 boxes of 4 KB; issued by one
    // thread they cost ~1,400 cycles per stage (the loop was issue-bound, gemm_timeline.py mn), so every box gets a lane.
    if (lane == 0 || MN) {
      int stage = 0; uint32_t phase = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int sp = item % a.splits;
        const int mt = (item / a.splits) % mt_n;
        const int g = item / (a.splits * mt_n);
        const int kb0 = sp * a.kb_per_split;
        const int kb1 = min(kb0 + a.kb_per_split, kb_total);
        if (a.dbg && lane == 0) a.dbg[(blockIdx.x * 4 + (item / gridDim.x)) * 8 + 0] = clock64();
        for (int kb = kb0; kb < kb1; kb++) {
          tc::mbar_wait(&empty[stage], phase ^ 1);
          if (a.dbg && lane == 0 && kb == kb1 - 1) a.dbg[(blockIdx.x * 4 + (item / gridDim.x)) * 8 + 1] = clock64();
          uint8_t* s = smem + (size_t)stage * SB;
          if (lane == 0) tc::mbar_arrive_expect_tx(&full[stage], (uint32_t)SB);
          if (MN) {
            const int per = AFB + N / 64;              // boxes per plane: A feature blocks, then B feature blocks
            if (lane < 2 * per) {
              const int pl = lane / per, j = lane % per;
              if (j < AFB)
                tc::tma_load_3d(s + pl * A_PLANE + j * MNB_BYTES, &tmA, &full[stage], (mt * 2 + j) * 64, kb * BKS, g * 2 + pl);
              else
                tc::tma_load_3d(s + 2 * A_PLANE + pl * B_PLANE_BYTES + (j - AFB) * MNB_BYTES, &tmB, &full[stage], (j - AFB) * 64,
                                kb * BKS, g * 2 + pl);
            }
            __syncwarp();
            if (++stage == stages) { stage = 0; phase ^= 1; }
            continue;
          }
          tc::tma_load_3d(s, &tmA, &full[stage], kb * BK, mt * BM, g * 2 + 0);
          tc::tma_load_3d(s + A_PLANE_BYTES, &tmA, &full[stage], kb * BK, mt * BM, g * 2 + 1);
          if (a.b_mn) {   // B given contraction-major ([K][N], e.g. W2 itself for Z1 = D2 W2): 64 x 64 boxes, MN-major UMMA operand
            for (int pl = 0; pl < 2; pl++)
              for (int fb = 0; fb < N / 64; fb++)
                tc::tma_load_3d(s + 2 * A_PLANE_BYTES + pl * B_PLANE_BYTES + fb * (BK * 128), &tmB, &full[stage], fb * 64, kb * BK,
                                g * 2 + pl);
          } else {
            tc::tma_load_3d(s + 2 * A_PLANE_BYTES, &tmB, &full[stage], kb * BK, 0, g * 2 + 0);
            tc::tma_load_3d(s + 2 * A_PLANE_BYTES + B_PLANE_BYTES, &tmB, &full[stage], kb * BK, 0, g * 2 + 1);
          }
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = tc::idesc_bf16_f32(BM, N) | (MN ? ((1u << 15) | (1u << 16)) : 0u) | ((!MN && a.b_mn) ? (1u << 16) : 0u);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
        const int sp = item % a.splits;
        const int kb0 = sp * a.kb_per_split;
        const int kb1 = min(kb0 + a.kb_per_split, kb_total);
        const int buf = FUSE ? 0 : (it & 1);
        const uint32_t bphase = FUSE ? ((uint32_t)it & 1u) : ((uint32_t)(it >> 1) & 1u);
        tc::mbar_wait(&tempty[buf], bphase ^ 1);       // epilogue has drained this accumulator
        if (FUSE) tc::mbar_wait(&tempty[1], bphase ^ 1);   // ... and the second half's
        tc::tc_fence_after();
        if (a.dbg) a.dbg[(blockIdx.x * 4 + it) * 8 + 2] = clock64();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * 256u;
        uint32_t acc = 0;
        for (int kb = kb0; kb < kb1; kb++) {
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          if (a.dbg && kb == kb0) a.dbg[(blockIdx.x * 4 + it) * 8 + 3] = clock64();
          if (a.dbg && kb == kb1 - 1) a.dbg[(blockIdx.x * 4 + it) * 8 + 4] = clock64();
          const uint32_t s = tc::smem_u32(smem + (size_t)stage * SB);
          if (MN) {
            const uint64_t mAh = tc::smem_desc_mn_sw128(s, MNB_BYTES), mAm = tc::smem_desc_mn_sw128(s + A_PLANE, MNB_BYTES);
            const uint64_t mBh = tc::smem_desc_mn_sw128(s + 2 * A_PLANE, MNB_BYTES);
            const uint64_t mBm = tc::smem_desc_mn_sw128(s + 2 * A_PLANE + B_PLANE_BYTES, MNB_BYTES);
#pragma unroll
            for (int k = 0; k < BKS / 16; k++) {
              const uint64_t off = (uint64_t)((k * 2048) >> 4);   // 16 samples = two 8-row groups of 1024 B
              tc::umma_bf16(d_tmem, mAm + off, mBh + off, idesc, acc);
              tc::umma_bf16(d_tmem, mAh + off, mBm + off, idesc, 1u);
              tc::umma_bf16(d_tmem, mAh + off, mBh + off, idesc, 1u);
              if (FUSE) {   // rows 128..255: feature blocks 2, 3 of the same A planes, second accumulator, same B block
                const uint64_t up = (uint64_t)((2 * MNB_BYTES) >> 4);
                tc::umma_bf16(d_tmem + 256u, mAm + up + off, mBh + off, idesc, acc);
                tc::umma_bf16(d_tmem + 256u, mAh + up + off, mBm + off, idesc, 1u);
                tc::umma_bf16(d_tmem + 256u, mAh + up + off, mBh + off, idesc, 1u);
              }
              acc = 1u;
            }
            tc::umma_commit(&empty[stage]);
            if (++stage == stages) { stage = 0; phase ^= 1; }
            continue;
          }
          const uint64_t dAh = tc::smem_desc_sw128(s), dAm = tc::smem_desc_sw128(s + A_PLANE_BYTES);
          const uint32_t sBh = s + 2 * A_PLANE_BYTES, sBm = sBh + B_PLANE_BYTES;
          const uint64_t dBh = a.b_mn ? tc::smem_desc_mn_sw128(sBh, BK * 128) : tc::smem_desc_sw128(sBh);
          const uint64_t dBm = a.b_mn ? tc::smem_desc_mn_sw128(sBm, BK * 128) : tc::smem_desc_sw128(sBm);
          const uint64_t bstep = a.b_mn ? (2048 >> 4) : (32 >> 4);   // 16 contraction rows of 128 B, or 16 bf16 along a row
#pragma unroll
          for (int k = 0; k < BK / 16; k++) {
            const uint64_t off = (uint64_t)((k * 32) >> 4);    // 16 bf16 = 32 B along K inside the swizzle atom
            const uint64_t offb = (uint64_t)k * bstep;
            tc::umma_bf16(d_tmem, dAm + off, dBh + offb, idesc, acc);   // small terms first
            tc::umma_bf16(d_tmem, dAh + off, dBm + offb, idesc, 1u);
            tc::umma_bf16(d_tmem, dAh + off, dBh + offb, idesc, 1u);
            acc = 1u;
          }
          tc::umma_commit(&empty[stage]);               // stage reusable once these MMAs have read it
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit(&tfull[buf]);                   // accumulator complete
        if (FUSE) tc::umma_commit(&tfull[1]);
      }
    }
    __syncwarp();
  } else if (MN && warp >= GEMM_THREADS / 32) {
    if (a.dw1.params)
      dw1_worker(a.dw1, reinterpret_cast<float*>(smem + (size_t)stages * SB + STG_BYTES + BAR_BYTES), threadIdx.x - GEMM_THREADS);
  } else {
    const int q = warp & 3;                             // TMEM lane quarter this warp may read
    // two warps share each TMEM lane quarter (and each SM sub-partition): one takes the low half of the columns, the
    // other the high half, so one warp's tcgen05.ld / LDG / store latency hides under the other's tanh (MUFU) work
    const int half = (warp - 2) >> 2;
    const int n_chunks = N / 32;
    const int c_begin = half ? ((n_chunks + 1) / 2) * 32 : 0;
    const int c_end = half ? N : ((n_chunks + 1) / 2) * 32;
    uint8_t* sb = smem + (size_t)stages * SB + (warp - 2) * STG_BUF;
    int it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
      const int sp = item % a.splits;
      const int g = item / (a.splits * mt_n);
      for (int mh = 0; mh < (FUSE ? 2 : 1); mh++) {      // FUSE: this item owns both accumulators (rows 0..127, 128..255)
      const int mt = FUSE ? mh : (item / a.splits) % mt_n;
      const int buf = FUSE ? mh : (it & 1);
      const uint32_t bphase = FUSE ? ((uint32_t)it & 1u) : ((uint32_t)(it >> 1) & 1u);
      tc::mbar_wait(&tfull[buf], bphase);
      tc::tc_fence_after();
      if (a.dbg && threadIdx.x == 64 + 128) a.dbg[(blockIdx.x * 4 + it) * 8 + 5] = clock64();
      const int row0 = mt * BM + q * 32;                 // first accumulator row of this warp
      const int plane = a.reduce_add ? g : g * a.splits + sp;
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 256u;
      for (int c = c_begin; c < c_end; c += 32) {
        float v[32];
        tc::tmem_ld32(t0 + (uint32_t)c, v);
        tc::tmem_ld_wait();
        if (lane == 0) tc::bulk_wait_read<0>();          // the previous store has finished reading the staging tile
        __syncwarp();
        // registers -> shared (row = lane, 128-B swizzle: 16-B chunk i of row r sits at chunk i ^ (r & 7)) -> TMA store
#pragma unroll
        for (int i = 0; i < 8; i++)
          *reinterpret_cast<float4*>(sb + lane * 128 + ((i ^ (lane & 7)) << 4)) =
              make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (a.reduce_add) tc::tma_reduce_add_3d(&tmC, sb, c, row0, plane);
          else tc::tma_store_3d(&tmC, sb, c, row0, plane);
          tc::bulk_commit();
        }
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&tempty[buf]);
      }
      if (a.dbg && threadIdx.x == 64 + 128) a.dbg[(blockIdx.x * 4 + it) * 8 + 6] = clock64();
    }
    if (lane == 0) tc::bulk_wait_all();                  // every output tile has landed before the CTA retires
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
  if (a.dbg && threadIdx.x == 32) a.dbg[(blockIdx.x * 4 + 2) * 8 + 7] = clock64();   // after the TMEM release
  if (MN && a.red.part_dw2) {
    // ---- reduce + Adam in the tail (GemmArgs::red).  Every CTA's dW2 tiles have landed (bulk_wait_all in the epilogue
    // warps) and its dW1 rows are stored; the __syncthreads above orders them before thread 0's fence.  Grid barrier:
    // all CTAs are resident (grid <= SMs, one CTA per SM), so spinning on the counter cannot starve anyone.
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(a.bar_counter, 1);
      long long spins = 0;
      while (true) {
        int v;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(a.bar_counter) : "memory");
        if (v - a.bar_target >= 0) break;
        __nanosleep(40);
        if (++spins > (1ll << 25)) __trap();               // ~2 s: a CUDA error, never a hung GPU
      }
    }
    __syncthreads();
    for (int vb = blockIdx.x; vb < a.red_vblocks; vb += gridDim.x)
      wide_reduce_finish(a.red, wide_reduce_body<true>(a.red, a.red_w2_blocks, vb, true));
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

}  // namespace

int gemm_x3_stages(int N, bool mn_major, bool fuse_m) {
  int s = (SMEM_LIMIT - 1024 - BAR_BYTES - STG_BYTES - (mn_major ? DW1_SMEM : 0)) /
          (mn_major ? ((fuse_m ? 8 : 4) + 2 * (N / 64)) * MNB_BYTES : stage_bytes(N));
  if (s > 6) s = 6;
  return s;
}

int make_operand_map(CUtensorMap* out, const void* base, uint64_t K, uint64_t rows, uint64_t planes,
                     uint64_t row_stride_elems, uint64_t plane_stride_elems, uint32_t box_rows) {
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
      return set_error(B2_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  cuuint64_t dims[3] = {K, rows, planes};
  cuuint64_t strides[2] = {row_stride_elems * 2, plane_stride_elems * 2};  // bytes, dims 1 and 2
  cuuint32_t box[3] = {(cuuint32_t)BK, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(B2_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d): K=%llu rows=%llu planes=%llu box_rows=%u",
                                          (int)r, (unsigned long long)K, (unsigned long long)rows, (unsigned long long)planes, box_rows);
  return B2_OK;
}

// fp32 output map: (N fastest, rows, planes), box 32 x 32 x 1, 128-B swizzle (matches the epilogue's staging tiles)
int make_output_map(CUtensorMap* out, const void* base, uint64_t N, uint64_t rows, uint64_t planes) {
  if (!g_encode) {
    CUtensorMap tmp;
    static __nv_bfloat16* dummy = nullptr;
    if (!dummy) cudaMalloc(&dummy, 64 * 64 * 2);
    int rc = make_operand_map(&tmp, dummy, 64, 64, 1, 64, 64 * 64, 32);   // resolves the driver entry point
    if (rc) return rc;
  }
  cuuint64_t dims[3] = {N, rows, planes};
  cuuint64_t strides[2] = {N * 4, rows * N * 4};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(B2_ERR_CUDA, "cuTensorMapEncodeTiled(output) failed (%d): N=%llu rows=%llu planes=%llu", (int)r,
                                          (unsigned long long)N, (unsigned long long)rows, (unsigned long long)planes);
  return B2_OK;
}

int make_output_map16(CUtensorMap* out, const void* base, uint64_t N, uint64_t rows, uint64_t planes) {
  if (!g_encode) {
    CUtensorMap tmp;
    static __nv_bfloat16* dummy = nullptr;
    if (!dummy) cudaMalloc(&dummy, 64 * 64 * 2);
    int rc = make_operand_map(&tmp, dummy, 64, 64, 1, 64, 64 * 64, 32);   // resolves the driver entry point
    if (rc) return rc;
  }
  cuuint64_t dims[3] = {N, rows, planes};
  cuuint64_t strides[2] = {N * 4, rows * N * 4};
  cuuint32_t box[3] = {16, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(B2_ERR_CUDA, "cuTensorMapEncodeTiled(output16) failed (%d): N=%llu rows=%llu planes=%llu", (int)r,
                                          (unsigned long long)N, (unsigned long long)rows, (unsigned long long)planes);
  return B2_OK;
}

int make_plane_store_map(CUtensorMap* out, const void* base, uint64_t N, uint64_t rows, uint64_t planes) {
  if (!g_encode) {
    CUtensorMap tmp;
    static __nv_bfloat16* dummy = nullptr;
    if (!dummy) cudaMalloc(&dummy, 64 * 64 * 2);
    int rc = make_operand_map(&tmp, dummy, 64, 64, 1, 64, 64 * 64, 32);   // resolves the driver entry point
    if (rc) return rc;
  }
  cuuint64_t dims[3] = {N, rows, planes};
  cuuint64_t strides[2] = {N * 2, rows * N * 2};
  cuuint32_t box[3] = {16, 32, 2};   // 16 columns x 32 rows of BOTH planes of a net (plane coordinate = 2 g): one store per chunk
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(B2_ERR_CUDA, "cuTensorMapEncodeTiled(plane store) failed (%d): N=%llu rows=%llu planes=%llu", (int)r,
                                          (unsigned long long)N, (unsigned long long)rows, (unsigned long long)planes);
  return B2_OK;
}

cudaError_t launch_gemm_x3(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmArgs& a,
                           int sm_count, cudaStream_t st) {
  if (a.fuse_m && !(a.mn_major && a.m_tiles == 2 && a.stages >= 2)) return cudaErrorInvalidValue;
  const int smem = a.stages * (a.mn_major ? ((a.fuse_m ? 8 : 4) + 2 * (a.N / 64)) * MNB_BYTES : stage_bytes(a.N)) + STG_BYTES + BAR_BYTES +
                   1024 + (a.mn_major ? DW1_SMEM : 0);
  static b2::PerDeviceOnce once;   // function attributes are per device
  if (once.need()) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_gemm_x3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_gemm_x3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)) != cudaSuccess) return e;
    once.done();
  }
  const int n_items = a.G * (a.fuse_m ? 1 : a.m_tiles) * a.splits;
  int grid = n_items < sm_count ? n_items : sm_count;
  if (grid < 1) return cudaSuccess;
  // 1 CTA per SM is assumed by the 512-column TMEM allocation: always ask for more than half of the shared memory
  const int smem_req = smem > 120 * 1024 ? smem : 120 * 1024;
  if (a.mn_major) {
    if (a.N % 64) return cudaErrorInvalidValue;
    return launch_pdl(k_gemm_x3<true>, dim3(grid), dim3(GEMM_THREADS + DW1_THREADS), (size_t)smem_req, st, tmA, tmB, tmC, a);
  }
  return launch_pdl(k_gemm_x3<false>, dim3(grid), dim3(GEMM_THREADS), (size_t)smem_req, st, tmA, tmB, tmC, a);
}

// ---- diagnostic GEMM: fp32 device operands -> planes -> k_gemm_x3<EPI_RAW> ----
__global__ void k_split_planes(const float* __restrict__ src, __nv_bfloat16* __restrict__ hi,
                               __nv_bfloat16* __restrict__ mid, int64_t rows, int64_t K, int64_t Kp) {
  const int64_t n = rows * Kp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / Kp, k = i % Kp;
    const float x = k < K ? src[r * K + k] : 0.0f;
    tc::split_bf16(x, hi[i], mid[i]);
  }
}

__global__ void k_sum_splits(const float* __restrict__ part, float* __restrict__ out, int64_t n, int splits) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < splits; p++) s += part[(int64_t)p * n + i];
    out[i] = s;
  }
}

int wide_selftest_gemm(int M, int N, int K, const float* A_dev, const float* B_dev, float* C_dev, int splits,
                       int mn_major, int sm_count, cudaStream_t st) {
  if (M < 1 || N < 32 || N > 256 || (N % 32) || K < 1) return set_error(B2_ERR_INVALID, "selftest gemm: need N %% 32 == 0, 32 <= N <= 256");
  const bool pair = mn_major == 3 || mn_major == 4;   // the CTA-pair weights-resident kernel (wide_gemm2.cu); 4: B is [K][N]
  const bool mixed = mn_major == 2 || mn_major == 4;  // A [M][K] K-major, B [K][N] contraction-major
  if (mixed || pair) mn_major = 0;
  if (pair && (splits > 1 || !gemm2_supported(N, (int)(((int64_t)K + BK - 1) / BK * BK), mixed)))
    return set_error(B2_ERR_INVALID, "selftest gemm (CTA pair): needs splits = 1, N %% 64 == 0 (128 for [K][N] weights), K <= 256");
  if ((mn_major && (M % 64)) || ((mn_major || mixed) && (N % 64)))
    return set_error(B2_ERR_INVALID, "selftest gemm: MN-major operands need M %% 64 == 0 (mode 1) and N %% 64 == 0 (modes 1, 2)");
  const int kblk = mn_major ? BKS : BK;
  const int64_t Kp = ((int64_t)K + kblk - 1) / kblk * kblk;
  __nv_bfloat16 *pa = nullptr, *pb = nullptr;
  float* part = nullptr;
  const int kb_total = (int)(Kp / kblk);
  if (splits < 1) splits = 1;
  if (splits > kb_total) splits = kb_total;
  const int kbps = (kb_total + splits - 1) / splits;
  splits = (kb_total + kbps - 1) / kbps;
  B2_CUDA_CHECK(cudaMalloc(&pa, sizeof(__nv_bfloat16) * 2 * M * Kp));
  B2_CUDA_CHECK(cudaMalloc(&pb, sizeof(__nv_bfloat16) * 2 * N * Kp));
  B2_CUDA_CHECK(cudaMalloc(&part, sizeof(float) * (size_t)splits * M * N));
  CUtensorMap tmA, tmB;
  int rc;
  if (mn_major) {   // A_dev [K][M], B_dev [K][N]; rows K..Kp stay zero
    B2_CUDA_CHECK(cudaMemsetAsync(pa, 0, sizeof(__nv_bfloat16) * 2 * M * Kp, st));
    B2_CUDA_CHECK(cudaMemsetAsync(pb, 0, sizeof(__nv_bfloat16) * 2 * N * Kp, st));
    k_split_planes<<<256, 256, 0, st>>>(A_dev, pa, pa + (int64_t)M * Kp, K, M, M);
    k_split_planes<<<256, 256, 0, st>>>(B_dev, pb, pb + (int64_t)N * Kp, K, N, N);
    rc = make_operand_map(&tmA, pa, M, Kp, 2, M, (uint64_t)M * Kp, MN_BOX_ROWS);
    if (!rc) rc = make_operand_map(&tmB, pb, N, Kp, 2, N, (uint64_t)N * Kp, MN_BOX_ROWS);
  } else if (mixed) {
    B2_CUDA_CHECK(cudaMemsetAsync(pb, 0, sizeof(__nv_bfloat16) * 2 * N * Kp, st));
    k_split_planes<<<256, 256, 0, st>>>(A_dev, pa, pa + (int64_t)M * Kp, M, K, Kp);
    k_split_planes<<<256, 256, 0, st>>>(B_dev, pb, pb + (int64_t)N * Kp, K, N, N);
    rc = make_operand_map(&tmA, pa, Kp, M, 2, Kp, (uint64_t)M * Kp, BM);
    if (!rc) rc = make_operand_map(&tmB, pb, N, Kp, 2, N, (uint64_t)N * Kp, BK);
  } else if (pair) {
    k_split_planes<<<256, 256, 0, st>>>(A_dev, pa, pa + (int64_t)M * Kp, M, K, Kp);
    k_split_planes<<<256, 256, 0, st>>>(B_dev, pb, pb + (int64_t)N * Kp, N, K, Kp);
    rc = make_operand_map(&tmA, pa, Kp, M, 2, Kp, (uint64_t)M * Kp, BM);
    if (!rc) rc = make_operand_map(&tmB, pb, Kp, N, 2, Kp, (uint64_t)N * Kp, (uint32_t)(N / 2));
  } else {
    k_split_planes<<<256, 256, 0, st>>>(A_dev, pa, pa + (int64_t)M * Kp, M, K, Kp);
    k_split_planes<<<256, 256, 0, st>>>(B_dev, pb, pb + (int64_t)N * Kp, N, K, Kp);
    rc = make_operand_map(&tmA, pa, Kp, M, 2, Kp, (uint64_t)M * Kp, BM);
    if (!rc) rc = make_operand_map(&tmB, pb, Kp, N, 2, Kp, (uint64_t)N * Kp, (uint32_t)N);
  }
  if (!rc) {
    GemmArgs g{};
    g.G = 1; g.M = M; g.N = N; g.K = (int)Kp; g.mn_major = mn_major ? 1 : 0; g.b_mn = mixed ? 1 : 0;
    g.m_tiles = (M + BM - 1) / BM; g.splits = splits; g.kb_per_split = kbps; g.stages = gemm_x3_stages(N, mn_major != 0);
    CUtensorMap tmC;
    if ((rc = make_output_map(&tmC, part, N, M, splits))) { cudaFree(pa); cudaFree(pb); cudaFree(part); return rc; }
    if (pair) {
      Gemm2Args g2{1, M, N, (int)Kp, gemm2_stages(N, (int)Kp)};
      cudaError_t e2 = launch_gemm2(mixed, tmA, tmB, tmC, g2, sm_count, st);
      if (e2 == cudaSuccess) {
        k_sum_splits<<<256, 256, 0, st>>>(part, C_dev, (int64_t)M * N, 1);
        e2 = cudaStreamSynchronize(st);
      }
      if (e2 != cudaSuccess) rc = set_error(B2_ERR_CUDA, "gemm2: %s", cudaGetErrorString(e2));
      cudaFree(pa); cudaFree(pb); cudaFree(part);
      return rc;
    }
    long long* dbg = nullptr;
    const char* dbg_env = getenv("B2_GEMM_DBG");
    if (dbg_env) {
      cudaMalloc(&dbg, sizeof(long long) * sm_count * 4 * 8);
      cudaMemsetAsync(dbg, 0, sizeof(long long) * sm_count * 4 * 8, st);
      g.dbg = dbg;
    }
    cudaError_t e = launch_gemm_x3(tmA, tmB, tmC, g, sm_count, st);
    if (dbg_env && e == cudaSuccess) {
      e = launch_gemm_x3(tmA, tmB, tmC, g, sm_count, st);   // second launch: warm L2
      std::vector<long long> h((size_t)sm_count * 32);
      cudaStreamSynchronize(st);
      cudaMemcpy(h.data(), dbg, sizeof(long long) * h.size(), cudaMemcpyDeviceToHost);
      {   // whole-grid view: setup cost (kernel entry -> roles start) and the slowest CTA
        long long worst = 0, setup_max = 0, tot_max = 0;
        for (int c = 0; c < sm_count; c++) {
          const long long t0 = h[(size_t)c * 32 + 7];
          for (int it = 0; it < 4; it++) worst = std::max(worst, h[(size_t)(c * 4 + it) * 8 + 6] - t0);
          setup_max = std::max(setup_max, t0 - h[(size_t)(c * 4 + 1) * 8 + 7]);
          tot_max = std::max(tot_max, h[(size_t)(c * 4 + 2) * 8 + 7] - h[(size_t)(c * 4 + 1) * 8 + 7]);
        }
        fprintf(stderr, "slowest CTA: last epilogue ends %lld cycles after setup; setup (entry -> roles) max %lld cycles; entry -> exit max %lld cycles\n",
                worst, setup_max, tot_max);
      }
      for (int c : {0, 1, 73, 147}) {
        if (c >= sm_count) continue;
        const long long t0 = h[(size_t)c * 32 + 7];
        for (int it = 0; it < 4; it++) {
          const long long* r = &h[(size_t)(c * 4 + it) * 8];
          if (!r[0] && !r[5]) continue;
          fprintf(stderr, "cta %3d item %d: prod_first %6lld prod_last_issue %6lld | mma_start %6lld first_full %6lld last_full %6lld | epi_start %6lld epi_end %6lld (cycles after setup)\n",
                  c, it, r[0] - t0, r[1] - t0, r[2] - t0, r[3] - t0, r[4] - t0, r[5] - t0, r[6] - t0);
        }
      }
      cudaFree(dbg);
    }
    if (e != cudaSuccess) rc = set_error(B2_ERR_CUDA, "gemm_x3: %s", cudaGetErrorString(e));
    else {
      k_sum_splits<<<256, 256, 0, st>>>(part, C_dev, (int64_t)M * N, splits);
      e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) rc = set_error(B2_ERR_CUDA, "gemm_x3 sync: %s", cudaGetErrorString(e));
    }
  }
  cudaFree(pa); cudaFree(pb); cudaFree(part);
  return rc;
}

}  // namespace b2
