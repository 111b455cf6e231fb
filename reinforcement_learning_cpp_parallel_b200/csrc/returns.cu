// returns.cu -- K2: discounted returns as a reverse SEGMENTED scan over the trajectory buffer.
//
// Reference semantics (src/ppo_nn.cpp:207-236): walking the records backwards,
//   carry = TIMEUP ? V(obs_k) : TERMINAL ? 0 : carry ;  R_k = r_k + gamma * carry      (all f32)
// i.e. GAE(lambda=1) whose segments end at flagged records and whose bootstrap uses the critic at the
// PRE-step observation with the weights the update starts from.  With b2_config.gae_lambda < 1 (engine
// extension, SURVEY Q5 "expose lambda but default 1") the carry of an unflagged record becomes the lambda-return
//   carry = (1 - lambda) V(obs_{k+1}) + lambda R_{k+1}
// which needs V at every record (k_values) and reduces to the reference at lambda = 1.
//
// Every flagged record CUTS the recurrence (its carry does not depend on R_{k+1}), so the records of one env between
// two flags are an independent segment.  Two ways to walk them:
//   k_returns      time-major batches with many envs ([T][N], N >= 1024: the device-resident rollout): one thread per
//                  env walks t = T-1..0, loads coalesced across envs and batched 8 deep.  Already N-way parallel.
//   k_seg_ends +   few envs (the reference's merged memory is ONE column of B records, src/main.cpp:290-303 ->
//   k_seg_returns  PPO::update(MemoryNN)): the segment ends (flagged records and each env's last record) are compacted
//                  into a list and every segment gets a WARP: 32 records per round are loaded coalesced, the previous
//                  segment's end is found with a ballot, and the recurrence is evaluated in the reference's order
//                  (mul, then add, no contraction) with the operands handed from lane to lane -- bit-identical to the
//                  sequential f32 evaluation.  A segment longer than 2048 records (no flags for thousands of steps:
//                  not a cart-pole episode) finishes as an affine-map scan, R_k = a_k + b_k R_{k+1} composed with five
//                  shuffle rounds per 32 records; that re-associates, ~1e-7 relative, inside the 1e-5 gate.
// Bootstrap values: k_bootstrap evaluates the critic at the (rare) flagged records only.
// Roofline: HBM, 13 B/record (r 4 + flag 1 + vboot 4 (flagged only) + R 4).
#include "engine.h"
#include "mlp.cuh"

namespace b2 {

constexpr unsigned FLAG_BOOT = B2_FLAG_TIMEUP | B2_FLAG_TRUNC;
constexpr unsigned FLAG_CUT = FLAG_BOOT | B2_FLAG_TERMINAL;

// V(obs_k) with the weights the update starts from: ALL records -> tr.vcur (lambda < 1), or the flagged ones -> tr.vboot
template <class N, bool ALL>
__global__ void __launch_bounds__(128, 4) k_bootstrap(const float* __restrict__ params, Traj tr, int64_t B) {
  extern __shared__ __align__(16) float sm[];
  float* hs = sm + NetSmem<N>::FLOATS;
  stage_net<N>(sm, params + N::SIZE);
  __syncthreads();
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < B; k += (int64_t)gridDim.x * blockDim.x) {
    const bool boot = tr.flags[k] & FLAG_BOOT;
    if (ALL || boot) {
      float x[N::S];
#pragma unroll
      for (int c = 0; c < N::S; c++) x[c] = tr.obs[c * tr.cap + k];
      const float v = mlp_row<N>(sm, hs, x);  // src/ppo_nn.cpp:227
      if (ALL) tr.vcur[k] = v;
      if (boot) tr.vboot[k] = v;
    }
  }
}

// one step of the recurrence, in the reference's operation order (:226-231); LAM adds the lambda-return mix
template <bool LAM>
__device__ __forceinline__ float returns_step(float r, unsigned f, float vboot, float vnext, float carry, float gamma,
                                              float lam) {
  if (f & FLAG_BOOT) carry = vboot;                      // :226-228
  else if (f & B2_FLAG_TERMINAL) carry = 0.0f;           // :229-230
  else if (LAM) carry = __fadd_rn(__fmul_rn(1.0f - lam, vnext), __fmul_rn(lam, carry));
  return __fadd_rn(r, __fmul_rn(gamma, carry));          // :231
}

template <bool LAM>
__global__ void __launch_bounds__(128) k_returns(Traj tr, float gamma, float lam) {
  const int64_t NE = tr.n_envs;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= NE) return;
  float carry = 0.0f;  // src/ppo_nn.cpp:215
  float vnext = 0.0f;  // V(obs_{t+1}); nothing follows the last record
  int64_t t = tr.horizon - 1;
  constexpr int U = 8;
  for (; t >= U - 1; t -= U) {
    float r[U], v[U];
    uint8_t f[U];
#pragma unroll
    for (int i = 0; i < U; i++) {
      const int64_t k = (t - i) * NE + e;
      r[i] = tr.reward[k];
      f[i] = tr.flags[k];
      if (LAM) v[i] = tr.vcur[k];
    }
#pragma unroll
    for (int i = 0; i < U; i++) {
      const int64_t k = (t - i) * NE + e;
      carry = returns_step<LAM>(r[i], f[i], (f[i] & FLAG_BOOT) ? tr.vboot[k] : 0.0f, vnext, carry, gamma, lam);
      if (LAM) vnext = v[i];
      tr.ret[k] = carry;
    }
  }
  for (; t >= 0; t--) {
    const int64_t k = t * NE + e;
    const uint8_t f = tr.flags[k];
    carry = returns_step<LAM>(tr.reward[k], f, (f & FLAG_BOOT) ? tr.vboot[k] : 0.0f, vnext, carry, gamma, lam);
    if (LAM) vnext = tr.vcur[k];
    tr.ret[k] = carry;
  }
}

// segment ends: every record that cuts the recurrence, plus each env's last record (whose carry starts from 0, :215)
__global__ void k_seg_ends(Traj tr, int64_t B, int64_t* __restrict__ ends, unsigned long long* __restrict__ count) {
  const int64_t last_row = (tr.horizon - 1) * tr.n_envs;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < B; k += (int64_t)gridDim.x * blockDim.x)
    if ((tr.flags[k] & FLAG_CUT) || k >= last_row) ends[atomicAdd(count, 1ull)] = k;
}

constexpr int SEG_EXACT_ROUNDS = 64;   // 2048 records in the reference's sequential order, then the affine-map scan

template <bool LAM>
__global__ void __launch_bounds__(128) k_seg_returns(Traj tr, float gamma, float lam, const int64_t* __restrict__ ends,
                                                     const unsigned long long* __restrict__ count) {
  const int lane = threadIdx.x & 31;
  const int64_t NE = tr.n_envs;
  const int64_t nseg = (int64_t)*count;
  const int64_t W = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t sgm = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); sgm < nseg; sgm += W) {
    const int64_t k_end = ends[sgm];
    const int64_t e = k_end % NE;
    int64_t t_hi = k_end / NE;             // lane l of a round owns record t_hi - l of env e
    float carry = 0.0f;                    // :215
    bool first = true, open = true;
    for (int round = 0; open; round++) {
      const int64_t t = t_hi - lane;
      const int64_t k = t * NE + e;
      const bool in = t >= 0;
      const float r = in ? tr.reward[k] : 0.0f;
      const unsigned f = in ? tr.flags[k] : 0u;
      const float vb = (in && (f & FLAG_BOOT)) ? tr.vboot[k] : 0.0f;
      float vnext = 0.0f;
      if (LAM && in && t + 1 < tr.horizon) vnext = tr.vcur[k + NE];
      // where the segment stops: the previous cut of this env (exclusive), or the start of the buffer
      unsigned cut = __ballot_sync(0xffffffffu, !in || ((f & FLAG_CUT) && !(first && lane == 0)));
      const int n = cut ? __ffs(cut) - 1 : 32;   // records of this round that belong to the segment
      open = cut == 0;
      float mine = 0.0f;
      if (round < SEG_EXACT_ROUNDS) {
        // the reference's order: one record after the other, operands handed from lane to lane (uniform control flow)
        for (int i = 0; i < n; i++) {
          const float ri = __shfl_sync(0xffffffffu, r, i);
          const unsigned fi = __shfl_sync(0xffffffffu, f, i);
          const float vbi = __shfl_sync(0xffffffffu, vb, i);
          const float vni = LAM ? __shfl_sync(0xffffffffu, vnext, i) : 0.0f;
          // only the segment's own end record (i == 0 of the first round) carries a flag inside the segment
          carry = returns_step<LAM>(ri, (first && i == 0) ? fi : 0u, vbi, vni, carry, gamma, lam);
          if (lane == i) mine = carry;
        }
      } else {
        // affine-map scan: R_l = a_l + b_l R_{l-1} with R_{-1} = carry; inclusive composition over lanes 0..l
        float a = r, b = __fmul_rn(gamma, LAM ? lam : 1.0f);
        if (LAM) a = __fadd_rn(r, __fmul_rn(gamma, __fmul_rn(1.0f - lam, vnext)));
        if (lane >= n) { a = 0.0f; b = 1.0f; }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const float ap = __shfl_up_sync(0xffffffffu, a, d), bp = __shfl_up_sync(0xffffffffu, b, d);
          if (lane >= d) { a = fmaf(b, ap, a); b *= bp; }
        }
        mine = fmaf(b, carry, a);
        carry = __shfl_sync(0xffffffffu, mine, n > 0 ? n - 1 : 0);
      }
      if (lane < n) tr.ret[k] = mine;
      first = false;
      t_hi -= 32;
    }
  }
}

// Batches narrower than this many envs take the segment-per-warp path (the per-env walk would leave the GPU idle:
// PPO::update(MemoryNN) is ONE column).
constexpr int64_t SEG_PATH_MAX_ENVS = 1024;

bool returns_needs_segment_list(const Traj& tr) { return tr.n_envs < SEG_PATH_MAX_ENVS; }

cudaError_t launch_returns(const float* params, const Traj& tr, const HyperDev& hp, cudaStream_t st, int* n_launches,
                           bool skip_bootstrap, int64_t* seg_ends, unsigned long long* seg_count) {
  using N = NetDefault;
  const int64_t B = tr.n_envs * tr.horizon;
  if (B == 0) return cudaSuccess;
  const bool lam = hp.gae_lambda < 1.0f;
  int64_t g = (B + 127) / 128;
  if (g > 148 * 16) g = 148 * 16;
  cudaError_t e = cudaSuccess;
  if (!skip_bootstrap) {   // wide nets bootstrap through wide_bootstrap (wide_update.cu)
    if (lam) k_bootstrap<N, true><<<(int)g, 128, mlp_smem_bytes<N>(128), st>>>(params, tr, B);
    else k_bootstrap<N, false><<<(int)g, 128, mlp_smem_bytes<N>(128), st>>>(params, tr, B);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (n_launches) *n_launches += 1;
  }
  if (returns_needs_segment_list(tr)) {
    if (!seg_ends || !seg_count) return cudaErrorInvalidValue;
    if ((e = cudaMemsetAsync(seg_count, 0, sizeof *seg_count, st)) != cudaSuccess) return e;
    k_seg_ends<<<(int)g, 128, 0, st>>>(tr, B, seg_ends, seg_count);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (lam) k_seg_returns<true><<<148 * 8, 128, 0, st>>>(tr, hp.gamma, hp.gae_lambda, seg_ends, seg_count);
    else k_seg_returns<false><<<148 * 8, 128, 0, st>>>(tr, hp.gamma, hp.gae_lambda, seg_ends, seg_count);
    if (n_launches) *n_launches += 2;
    return cudaGetLastError();
  }
  if (lam) k_returns<true><<<(int)((tr.n_envs + 127) / 128), 128, 0, st>>>(tr, hp.gamma, hp.gae_lambda);
  else k_returns<false><<<(int)((tr.n_envs + 127) / 128), 128, 0, st>>>(tr, hp.gamma, hp.gae_lambda);
  if (n_launches) *n_launches += 1;
  return cudaGetLastError();
}

}  // namespace b2

// ------------------------------------------------------------------------------------------------------------
// b2_returns_noalloc (include/b200ppo.h): the segment-per-warp returns on caller-owned device buffers with the segment
// list in a static device array -- for callers that cannot call cudaMalloc (graph capture, serving loops).  No
// counterpart in the reference.  Nothing above this line is changed by it.
// ------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace b2 {
namespace {
// [0, 8): the segment counter; [256, 256 + 8 * B2_NOALLOC_MAX_RECORDS): the segment ends (k_seg_ends writes at most one
// entry per record).  One copy per GPU that loads the library, for the life of the process.
constexpr size_t NOALLOC_COUNT_BYTES = 256;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_COUNT_BYTES + sizeof(int64_t) * B2_NOALLOC_MAX_RECORDS];

// In host code `g_noalloc_scratch` is only the symbol's host-side stand-in: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_dev[64] = {};

int noalloc_scratch(unsigned char** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_error(B2_ERR_NO_DEVICE, "b2_returns_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return set_error(B2_ERR_INVALID, "b2_returns_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_dev[dev]) {
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return set_error(B2_ERR_CUDA, "b2_returns_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
    g_noalloc_dev[dev] = p;
  }
  *out = g_noalloc_dev[dev];
  return B2_OK;
}
}  // namespace
}  // namespace b2

extern "C" int b2_returns_noalloc(const float* reward, const uint8_t* flags, const float* vboot, int64_t n_envs, int64_t horizon,
                                  float gamma, float* returns, void* stream) {
  using namespace b2;
  if (!reward || !flags || !vboot || !returns)
    return set_error(B2_ERR_INVALID, "b2_returns_noalloc: reward, flags, vboot and returns must be device pointers (got a NULL)");
  if (n_envs <= 0 || horizon <= 0)
    return set_error(B2_ERR_INVALID, "b2_returns_noalloc: n_envs (%lld) and horizon (%lld) must be positive", (long long)n_envs,
                     (long long)horizon);
  if (n_envs > B2_NOALLOC_MAX_RECORDS || horizon > B2_NOALLOC_MAX_RECORDS || n_envs * horizon > B2_NOALLOC_MAX_RECORDS)
    return set_error(B2_ERR_INVALID,
                     "b2_returns_noalloc: %lld x %lld records exceed B2_NOALLOC_MAX_RECORDS = %d (the static scratch array); "
                     "use b2_returns / b2_update_resident for larger batches",
                     (long long)n_envs, (long long)horizon, (int)B2_NOALLOC_MAX_RECORDS);
  unsigned char* scratch = nullptr;
  if (int rc = noalloc_scratch(&scratch)) return rc;
  unsigned long long* count = reinterpret_cast<unsigned long long*>(scratch);
  int64_t* ends = reinterpret_cast<int64_t*>(scratch + NOALLOC_COUNT_BYTES);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Traj tr{};
  const int64_t B = n_envs * horizon;
  tr.cap = B; tr.n_envs = n_envs; tr.horizon = horizon;
  tr.reward = const_cast<float*>(reward); tr.flags = const_cast<uint8_t*>(flags); tr.vboot = const_cast<float*>(vboot);
  tr.ret = returns;
  cudaError_t e;
  // the counter must start at zero; ends[] keeps the previous call's entries, and k_seg_returns reads only the first
  // `count` of them, all written by THIS call's k_seg_ends
  if ((e = cudaMemsetAsync(count, 0, sizeof *count, st)) != cudaSuccess)
    return set_error(B2_ERR_CUDA, "b2_returns_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
  int64_t g = (B + 127) / 128;
  if (g > 148 * 16) g = 148 * 16;
  k_seg_ends<<<(int)g, 128, 0, st>>>(tr, B, ends, count);
  if ((e = cudaGetLastError()) != cudaSuccess) return set_error(B2_ERR_CUDA, "b2_returns_noalloc: k_seg_ends: %s", cudaGetErrorString(e));
  k_seg_returns<false><<<148 * 8, 128, 0, st>>>(tr, gamma, 1.0f, ends, count);
  if ((e = cudaGetLastError()) != cudaSuccess) return set_error(B2_ERR_CUDA, "b2_returns_noalloc: k_seg_returns: %s", cudaGetErrorString(e));
  return B2_OK;
}
