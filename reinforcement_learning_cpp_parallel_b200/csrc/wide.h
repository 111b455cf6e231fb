// wide.h -- the tensor-core update path for hidden widths that make the PPO update a real dense GEMM
// (BASELINE.json configs[4]: 2x256 MLPs).  The default 6-64-32-1 nets never come here (ppo_update.cu).
//
// Per optimizer step and per net the reference computes (src/ppo_nn.cpp:126-194 evaluate, :263-290 loss,
// :301-303 backward + Adam):
//     H1 = tanh(X W1^T + b1)        K = 6      -> FP32 FMA      (k_wide_l1)
//     Z2 = H1 W2^T                  K = H1     -> tcgen05 GEMM  (k_gemm_x3<false>)
//     H2 = tanh(Z2 + b2), y = H2 W3^T + b3, loss head, D2 = dz3 (x) W3 .* (1 - H2^2)   (k_wide_head)
//     Z1 = D2 W2                    K = H2     -> tcgen05 GEMM  (k_gemm_x3<false>)
//     dW2 = D2^T H1                 K = samples-> tcgen05 GEMM  (k_gemm_x3<true>, MN-major operands, split-K partials)
//     D1 = Z1 .* (1 - H1^2), dW1 = D1^T X, db1 -> FP32 FMA      (k_wide_dw1, recomputes H1 from X)
//     db2, dW3, db3                            -> FP32 FMA      (k_wide_head)
// The three GEMMs hold 99 % of the 801,792 flop per sample-epoch.  The 1e-4 gradient gate
// rules out single-pass BF16/TF32, so every fp32 operand is split into a (hi, mid) bf16 pair and each product
// is issued as three bf16 MMAs hi*hi + mid*hi + hi*mid with fp32 accumulation in TMEM ("bf16x3", error
// ~2^-16 relative): 1.5x cheaper on the tensor pipe than 3xTF32 and the same 4 B per element in memory.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "engine.h"

namespace b2 {

// Programmatic dependent launch: the kernels of one optimizer step form a strict chain of 6 short launches (15-30 us
// each), so launch latency and grid ramp-up are a visible share of the step.  Every kernel of the chain calls
// pdl_launch_dependents() on entry (the next grid may start being scheduled as SMs free up) and pdl_wait() before its
// first read of anything an earlier kernel wrote (blocks until ALL earlier grids have completed and flushed).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// runtime layout of policy.parameters() for arbitrary hidden widths (same order as common.cuh Net<>)
struct NetDims {
  int S, H1, H2;
  int oW1, ob1, oW2, ob2, oW3, ob3, SIZE, P;
  __host__ __device__ static NetDims make(int S, int H1, int H2) {
    NetDims d;
    d.S = S; d.H1 = H1; d.H2 = H2;
    d.oW1 = 0; d.ob1 = d.oW1 + H1 * S; d.oW2 = d.ob1 + H1; d.ob2 = d.oW2 + H2 * H1;
    d.oW3 = d.ob2 + H2; d.ob3 = d.oW3 + H2; d.SIZE = d.ob3 + 1; d.P = 2 * d.SIZE;
    return d;
  }
};

// k_wide_dw1's work, carried by eight extra warps of the dW2 GEMM kernel (FP32 pipe beside the tensor pipe): per CTA a
// contiguous slice of the chunk's samples, both nets; D1 = Z1 .* (1 - H1^2), dW1aug = D1^T [x, 1] -> part[cta][g][k*8+c]
struct Dw1Args {
  const float* params;      // null: no fused dw1 work
  const float* obs;         // trajectory observations, SoA: obs[c * obs_cap + position]
  int64_t obs_cap;
  const int32_t* idx;
  int64_t first;
  int count, H1, SIZE, oW1, ob1;
  const float* z1;          // [2][cap][H1]
  int64_t cap;
  float* part;              // [grid][2][H1 * 8]
  float* sum;               // k_gemm_mn2 only, non-null: ONE row [2][H1 * 8] every CTA adds its partial into (bulk reduce-add)
                            // instead of writing part[cta] -- the merged layer-1 + Adam kernel reads it (wide_update.cu k_wide_l1r)
};

// Adam coefficients of the current optimizer step (computed once per step by k_wide_l1)
struct AdamCoef {
  float fb1, f1mb1, fb2, f1mb2, sq_bc2, feps, step_size;
};

// arguments of the reduce + Adam stage (wide_reduce.cuh)
struct ReduceArgs {
  NetDims d;
  UpdateArgs u;
  const float *part_head, *part_dw1;
  float* part_dw2;
  int dw2_added;        // the dW2 GEMM added its split-K tiles into ONE plane per net (TMA reduce): read it, then zero it
  int head_blocks, head_row, dw1_blocks, splits, pstride;
  int first_chunk, last_chunk, apply_adam;
  __nv_bfloat16* w2p;          // weight planes the Adam step keeps current
  float* bias2;
  const AdamCoef* coef;        // written by k_wide_l1 at the start of this optimizer step
};


struct GemmArgs {
  int G, M, N, K;                // per group g < G: C[M][N] = A[M][K] . B[N][K]^T ; K % 64 == 0, N % 16 == 0, N <= 256
  int m_tiles, splits, kb_per_split, stages;
  int b_mn;                      // K-major kernel only: B is given contraction-major [K][N] (N % 64 == 0), read as an MN-major operand
  int mn_major;                  // 1: C[M][N] = A[K][M]^T . B[K][N] (K blocks of 32, K % 32 == 0, N % 64 == 0)
                                 // output: raw fp32 accumulator tiles, TMA-stored to plane g * splits + split of tmC
  int fuse_m;                    // MN-major kernel, m_tiles == 2 only: one CTA owns BOTH 128-row halves of the output (two TMEM
                                 // accumulators fed from one copy of the B block), so B is streamed once instead of twice
  int reduce_add;                // split-K: every split ADDS its tile into plane g (TMA reduce, fp32 adds in L2) instead of
                                 // storing to plane g * splits + split; the output must be zero before the launch
  long long* dbg;                // optional per-CTA timeline (clock64), diagnostic runs only
  long long* tl;                 // optional whole-step timeline (tc::tl_stamp), diagnostic runs only
  Dw1Args dw1;                   // MN-major kernel only
  // MN-major kernel only: the reduce + Adam stage of the step runs in the tail of this kernel, behind a grid barrier
  // (every CTA is resident: grid <= SM count, one CTA per SM), instead of in a launch of its own.  red.part_dw2 == nullptr: off
  ReduceArgs red;
  int red_w2_blocks, red_vblocks;   // virtual blocks of blockDim.x elements (W2 part) / blockDim.x / 32 small parameters
  int* bar_counter;                 // device counter, monotonically increasing across launches
  int bar_target;                   // value it reaches when every CTA of THIS launch has arrived
};

constexpr float HEADW_B2_SCALE = 2.8853900817779268f;   // 2 log2(e): tanh x = 1 - 2 / (2^(x * this) + 1)
// k_wide_fh (wide_fused.cu): layer-2 forward GEMM + loss head in one kernel, both nets (critic first)
struct FhArgs {
  NetDims d;
  int N, K, stages;            // N = H2, K = H1
  int count, m_pad;            // valid samples of this chunk, padded to 128
  int64_t cap, first;          // chunk capacity (row count of the planes), position of sample 0
  const int32_t* idx;
  Traj tr;
  const float* params;
  const float* headw;          // [g][2][H2]: b2 * HEADW_B2_SCALE, W3 (16-B aligned copies kept current by the Adam step)
  float* vtmp;                 // [4][cap]: V of the critic pass, one copy per column-quarter warp
  float inv_denom;
  HyperDev hp;
  float* part_head;            // [grid][2][head_row]
  float* head_sum;             // non-null: ONE row [2][head_row] every CTA adds its partial into (red.add) instead of part_head[cta]
  int head_row;
  long long* dbg;              // optional clock64 timeline of the head warps, [grid][16] (B2_FH_DBG=1, diagnostic runs)
  int dx;                      // 1: the dX GEMM Z1 = D2 W2 of both nets runs in this kernel too (passes 2, 3)
  long long* tl;               // optional whole-step timeline (tc::tl_stamp), diagnostic runs only
};
int fh_stages(int N, int K);
bool fh_supported(int N, int K);
// tmA: h1p operand map (128-row boxes), tmB: K-major W2 in (H2/2)-row boxes, tmD: bf16 plane store map of d2p (32 x 32 boxes)
// a.dx: tmA2 = d2p operand map (128-row boxes), tmB2 = W2 as [K = j][N = k] in 64 x 64 boxes, tmC = d1 (fp32) store map with
// 16-column boxes (make_output_map16); ignored otherwise
cudaError_t launch_wide_fh(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD, const CUtensorMap& tmA2,
                           const CUtensorMap& tmB2, const CUtensorMap& tmC, const FhArgs& a, int sm_count, int* grid_out, cudaStream_t st);
// fp32 output map with boxes of 16 columns x 32 rows and the 64-B swizzle (the head warps' 2-KB staging tiles)
int make_output_map16(CUtensorMap* out, const void* base, uint64_t N, uint64_t rows, uint64_t planes);
// bf16 planes [planes][rows][N] written in boxes of 16 columns x 32 rows x 2 planes with the 32-B swizzle (the head's staging tiles)
int make_plane_store_map(CUtensorMap* out, const void* base, uint64_t N, uint64_t rows, uint64_t planes);

// the CTA-pair, weights-resident GEMM of wide_gemm2.cu (layer-2 forward and dX): C[g][M][N] = A[g][M][K] . B[g]
struct Gemm2Args {
  int G, M, N, K;      // K, N <= 256, multiples of 64 (N % 128 == 0 when B is contraction-major)
  int stages;          // A stages in shared memory (gemm2_stages)
};
int gemm2_stages(int N, int K);
bool gemm2_supported(int N, int K, bool bmn);
// tmA: operand map with 128-row boxes; tmB: K-major weights with (N/2)-row boxes, or (bmn) 64 x 64 boxes of [K][N];
// tmC: fp32 output map, plane = g
cudaError_t launch_gemm2(bool bmn, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const Gemm2Args& a,
                         int sm_count, cudaStream_t st);

// the sample contraction on CTA pairs (wide_gemm_mn2.cu): M = 256 output rows per pair, split-K items g * splits + split
int gemm_mn2_stages(int N);
bool gemm_mn2_supported(int M, int N);
cudaError_t launch_gemm_mn2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmArgs& a, int sm_count,
                            int* grid_out, cudaStream_t st);

cudaError_t launch_gemm_x3(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmArgs& a,
                           int sm_count, cudaStream_t st);
// fp32 output tensor [planes][rows][N] the epilogue TMA-stores 32 x 32 tiles into (plane = group, or group*splits + split)
int make_output_map(CUtensorMap* out, const void* base, uint64_t N, uint64_t rows, uint64_t planes);
int gemm_x3_stages(int N, bool mn_major = false, bool fuse_m = false);

// 3-D bf16 tensor map (K fastest, rows, planes) with a (64 x box_rows x 1) box and the 128-B swizzle.
// Returns 0 on success; the driver entry point is resolved through the runtime (no -lcuda link dependency).
int make_operand_map(CUtensorMap* out, const void* base, uint64_t K, uint64_t rows, uint64_t planes,
                     uint64_t row_stride_elems, uint64_t plane_stride_elems, uint32_t box_rows);
enum { MN_BOX_ROWS = 32 };   // box rows of the maps the MN-major GEMM reads (samples per K block)

// ---- the whole wide-net learner: buffers, tensor maps and the per-step kernel sequence ----
struct WideNet;
int wide_create(WideNet** out, int H1, int H2, int64_t max_chunk, int sm_count);
void wide_destroy(WideNet* w);
int64_t wide_chunk_cap(const WideNet* w);
// gradient of one minibatch [first, first+count) (mapped through idx if set) into a.grads (P grads + 3 loss sums in
// the tail, same convention as k_reduce); apply_adam fuses the Adam step (single GPU).  *launches += kernels launched.
// more_follow: another step of the same update comes next -- the reduce + Adam of this step may then ride in the layer-1
// kernel of the next one (k_wide_l1r); wide_flush() applies whatever is still pending (call it before the parameters,
// gradients or the loss log are read; a no-op when nothing is pending).
int wide_step(WideNet* w, const UpdateArgs& a, bool apply_adam, bool more_follow, cudaStream_t st, int64_t* launches);
int wide_flush(WideNet* w, cudaStream_t st, int64_t* launches);
cudaError_t launch_wide_adam(WideNet* w, const UpdateArgs& a, cudaStream_t st);
int wide_bootstrap(WideNet* w, const float* params, const Traj& tr, cudaStream_t st, int64_t* launches);
int wide_pstride(int H1, int H2);
// forward-only helpers (action query / value estimate / bootstrap) for wide nets
cudaError_t launch_wide_forward(const NetDims& d, const float* net_params, const float* obs, int64_t obs_stride_c,
                                int64_t obs_stride_k, const int32_t* rows, int64_t n, int final_tanh, float* y,
                                cudaStream_t st);

// device-resident rollout for wide actors, one env step = observe + actor forward + step (rollout_wide.cu)
cudaError_t launch_rollout_wide(const RolloutArgs& a, const NetDims& d, int precision, bool trace, cudaStream_t st,
                                int64_t* launches);

cudaError_t launch_wide_act_finish(const float* mu, const float* eps, int64_t n, uint64_t seed, uint64_t call, float* action,
                                   float* logp, const HyperDev& hp, int* status, cudaStream_t st);

// diagnostic: C = A . B^T through the bf16x3 tcgen05 GEMM on device buffers (fp32 in, fp32 out)
int wide_selftest_gemm(int M, int N, int K, const float* A_dev, const float* B_dev, float* C_dev, int splits,
                       int mn_major, int sm_count, cudaStream_t st);

}  // namespace b2
