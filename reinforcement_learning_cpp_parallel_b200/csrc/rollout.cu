// rollout.cu -- K1: fused batched cart-pole rollout.
//
// One LANE owns one environment while it is being advanced: the 4-component state, the episode counters
// and the RK accumulators stay in registers.  Per step the lane (1) forms the f32 observation, (2) runs the
// actor MLP out of shared-memory weights, (3) draws its Philox Gaussian, (4) evaluates the log-density,
// (5) integrates 50 x 6 RK46-NL stages, (6) emits the transition with one coalesced store per field (SoA,
// time-major), (7) applies the TIMEUP/reset rule.  There is no per-step message and no host involvement:
// this replaces env_run (src/main.cpp:12-148), respond_to_env/getAction (src/main.cpp:165-188,
// src/ppo_nn.cpp:308-331,99-125) and the MPI traffic.
//
// Scheduling.  Envs are packed into groups of 32 (one warp's worth).  The kernel is PERSISTENT: it launches
// only warps that are resident at once (up to 4 per SM sub-partition at <= 128 registers) and deals the
// G x T (group, step) work units out evenly -- warp w owns units [w U / W, (w+1) U / W) of the group-major
// sequence, W = the largest multiple of the sub-partition count that is <= G and resident.
// Measured (profiles/r1c_saturation.txt): two warps already saturate a sub-partition's FP64 pipe, so what
// limits one-group-per-warp scheduling at 65,536 envs is quantisation -- 2048 groups over 592
// sub-partitions is 3.46 on average but 4 on the busiest, 13.5 % of the machine idle.  Cutting along the
// time axis removes that: every warp gets the same number of env-steps (+-1).  A warp's range covers the
// tail of one group, whole groups, and the head of another; it runs the HEAD first, so the next warp
// (which also starts with its own head) finds the hand-off state ready when it reaches that group's tail:
// with equal speeds nobody waits.  Hand-off goes through the state arrays in global memory plus a
// per-group progress word (release/acquire).  All warps are co-resident by construction, and a waiting
// warp only ever depends on work that a lower-numbered warp does before anything else, so the spin
// cannot deadlock (and the wait is bounded all the same, see rollout_segment).
//
// Roofline: FP64 (B2_PREC_F64) or FP32 FMA issue, not HBM -- 41 B are written per env-step against
// ~12k-24k arithmetic instructions.
#include <stdlib.h>

#include "cartpole.cuh"
#include "engine.h"
#include "mlp.cuh"

namespace b2 {

constexpr int ROLLOUT_BLOCK = 128;  // 4 warps: one per SM sub-partition
constexpr int ROLLOUT_CTAS_PER_SM = 4;

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Advance group g (envs 32 g .. 32 g + 31) over steps [t0, t1) of this LAUNCH (absolute step = t_begin + t).
// VAR: the opt-in env variants of b200ppo.h B2_ENV_* (a.env_flags), double physics in the reference's operation order;
// a separate instantiation, so the shipped-env kernels are untouched by it.
template <typename Real, class N, bool TRACE, bool VAR>
__device__ __forceinline__ void rollout_segment(const RolloutArgs& a, const float* sm, float* hs, int64_t g, int t0,
                                                int t1) {
  using Dyn = CartPoleDyn<Real>;
  const int lane = threadIdx.x & 31;
  const int64_t NE = a.n_envs;
  const int64_t e = g * 32 + lane;
  if (t0 > 0) {  // hand-off: wait until the warp that owns the earlier steps has published the state
    // The wait is BOUNDED (~3 s): the no-deadlock argument above needs every warp of the grid to be resident, which another
    // stream or process holding SMs can break.  On expiry status bit 2 is raised (b2_rollout reports B2_ERR_STATE) and
    // the segment is skipped -- an error, never a hung GPU.
    int ok = 1;
    if (lane == 0) {
      int spins = 0;
      while (ld_acquire(a.progress + g) < t0) {
        if (++spins > (1 << 24)) { atomicOr(a.status, 4); ok = 0; break; }
        __nanosleep(200);
      }
    }
    ok = __shfl_sync(0xffffffffu, ok, 0);
    if (!ok) return;
  }
  if (e >= NE) return;  // ragged last group (no warp-level sync below this line, except under full_group)
  const bool full_group = g * 32 + 32 <= NE;

  using S = typename StateOf<Real>::type;   // double for B2_PREC_MIXED
  S u[4];
#pragma unroll
  for (int j = 0; j < 4; j++) u[j] = (S)__ldcg(a.state + j * NE + e);
  int es = __ldcg(a.ep_step + e);
  float er = __ldcg(a.ep_ret + e);
  int ecount = (a.t_begin + t0 == 0) ? 0 : __ldcg(a.ep_count + e);
  bool bad = false;
  const int64_t cap = a.tr.cap;

  for (int t = a.t_begin + t0; t < a.t_begin + t1; t++) {
    const int64_t k = (int64_t)t * NE + e;
    float o[6];
    Dyn::observe(u, o);
    const float mu = act_tanh(mlp_row<N>(sm, hs, o));                       // ppo_nn.h:98-102
    float act;
    if (a.forced_action) {
      act = a.forced_action[k];
    } else {
      const float eps = gauss_noise(a.seed, a.env_offset + (uint64_t)e, a.step0 + (uint64_t)t);
      act = (float)__dadd_rn((double)mu, __dmul_rn(a.hp.sigma_d, (double)eps));  // src/ppo_nn.cpp:107-108
    }
    const float lp = gauss_logprob(act, mu, a.hp.inv_var, a.hp.half_log_norm);   // :115
    bad |= !isfinite(act);                                               // src/main.cpp:50-55

    bool terminal = false;
    float r;
    if constexpr (VAR) {
      const CartPoleVar::Adv adv = CartPoleVar::advance(make_double4((double)u[0], (double)u[1], (double)u[2], (double)u[3]),
                                                        scale_action(act), a.env_flags, es + 1);
      u[0] = (S)adv.u.x; u[1] = (S)adv.u.y; u[2] = (S)adv.u.z; u[3] = (S)adv.u.w;
      terminal = adv.term != 0;                                          // src/main.cpp:81-85
      const double ud[4] = {adv.u.x, adv.u.y, adv.u.z, adv.u.w};
      r = CartPoleVar::reward(a.env_flags, ud);
    } else {
      Dyn::advance(u, (S)scale_action(act));                          // src/main.cpp:57-71
      r = Dyn::reward(u);                                                // :73
    }
    es += 1;
    er += r;
    const bool timeup = !terminal && (es == a.n_timestep);               // :86-88 (TERMINATE never fires in the shipped env, Q1)
    const uint8_t fl = terminal ? (uint8_t)B2_FLAG_TERMINAL
                                : (timeup ? (uint8_t)B2_FLAG_TIMEUP : ((t == a.T - 1) ? (uint8_t)B2_FLAG_TRUNC : (uint8_t)0));

#pragma unroll
    for (int c = 0; c < 6; c++) __stcs(a.tr.obs + c * cap + k, o[c]);   // streaming: written once, read by K3 later
    if (a.obs_aos) {                                                     // host-bound copy, record-major (b2_rollout_host)
      if (full_group) {
        // the warp's 32 records are 768 contiguous bytes: transpose through shuffles so that each of the 6 stores is one
        // fully coalesced 128-B line (PCIe writes of whole lines when the destination is pinned host memory)
        float* blk = a.obs_aos + 6 * ((int64_t)t * NE + g * 32);
#pragma unroll
        for (int i = 0; i < 6; i++) {
          const int idx = i * 32 + lane, src = idx / 6, c = idx - 6 * src;
          float v = __shfl_sync(0xffffffffu, o[0], src);
#pragma unroll
          for (int cc = 1; cc < 6; cc++) {
            const float w = __shfl_sync(0xffffffffu, o[cc], src);
            v = (c == cc) ? w : v;
          }
          __stcs(blk + idx, v);
        }
      } else {
        float2* rec = reinterpret_cast<float2*>(a.obs_aos + 6 * k);
        __stcs(rec, make_float2(o[0], o[1])); __stcs(rec + 1, make_float2(o[2], o[3])); __stcs(rec + 2, make_float2(o[4], o[5]));
      }
    }
    if (a.obs4_aos) {                                                    // compact host record: one 16-B store per record
      __stcs(reinterpret_cast<float4*>(a.obs4_aos) + k, make_float4(o[0], o[1], o[2], o[3]));
      a.rflags[k] = (uint8_t)(fl | (r != 0.0f ? B2_FLAG_REWARD : 0u));
    }
    __stcs(a.tr.action + k, act);
    __stcs(a.tr.logp + k, lp);
    __stcs(a.tr.mu + k, mu);
    __stcs(a.tr.reward + k, r);
    a.tr.flags[k] = fl;
    if (a.h_action) __stcs(a.h_action + k, act);                         // host-resident copies (b2_rollout_host, pinned buffers)
    if (a.h_logp) __stcs(a.h_logp + k, lp);
    if (a.h_mu) __stcs(a.h_mu + k, mu);
    if (a.h_reward) __stcs(a.h_reward + k, r);
    if (a.h_flags) a.h_flags[k] = fl;
    if (TRACE) {
      double4 s4 = make_double4((double)u[0], (double)u[1], (double)u[2], (double)u[3]);
      *reinterpret_cast<double4*>(a.tr.state_trace + 4 * k) = s4;
    }
    if (timeup || terminal) {                                            // :137-142 log, :27 reset
      if (ecount < a.max_eps) a.ep_log[(int64_t)ecount * NE + e] = er;
      ecount++;
      if constexpr (VAR) {
        double u0[4];
        CartPoleVar::init_state(a.env_flags, a.seed, a.env_offset + (uint64_t)e, a.step0 + (uint64_t)t + 1, u0);
        u[0] = (S)u0[0]; u[1] = (S)u0[1]; u[2] = (S)u0[2]; u[3] = (S)u0[3];
      } else {
        u[0] = u[1] = u[2] = u[3] = (S)0;                             // cart_pole.h:86
      }
      es = 0;
      er = 0.0f;
    }
  }
#pragma unroll
  for (int j = 0; j < 4; j++) __stcg(a.state + j * NE + e, (double)u[j]);
  __stcg(a.ep_step + e, es);
  __stcg(a.ep_ret + e, er);
  __stcg(a.ep_count + e, ecount);
  if (bad) atomicOr(a.status, 1);
}

template <typename Real, class N, bool TRACE, bool VAR = false>
__global__ void __launch_bounds__(ROLLOUT_BLOCK, ROLLOUT_CTAS_PER_SM) k_rollout(const RolloutArgs a) {
  extern __shared__ __align__(16) float sm[];
  float* hs = sm + NetSmem<N>::FLOATS;  // hidden-1 scratch, one column per thread
  stage_net<N>(sm, a.params);  // actor only: V is not needed while acting (src/ppo_nn.cpp:100)
  __syncthreads();

  const int wpb = blockDim.x >> 5;
  const int64_t w = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5);
  const int64_t W = (int64_t)gridDim.x * wpb;
  const int64_t T = a.t_end - a.t_begin;
  const int64_t G = (a.n_envs + 31) >> 5;
  const int64_t U = G * T;
  const int64_t u0 = w * U / W, u1 = (w + 1) * U / W;   // U W < 2^63 for any shape that fits in HBM
  if (u0 >= u1) return;
  const int64_t g_first = u0 / T, g_last = (u1 - 1) / T;
  const int ng = (int)(g_last - g_first) + 1;
  for (int i = 0; i < ng; i++) {
    // head of the last group first, then the whole groups, then the tail of the first group
    const int64_t g = (i == 0) ? g_last : ((i == ng - 1) ? g_first : g_first + i);
    const int64_t lo = g * T, hi = lo + T;
    const int t0 = (int)((u0 > lo ? u0 : lo) - lo);
    const int t1 = (int)((u1 < hi ? u1 : hi) - lo);
    rollout_segment<Real, N, TRACE, VAR>(a, sm, hs, g, t0, t1);
    if (t1 < (int)T) {  // publish: every lane's stores, then the progress word
      __threadfence();
      __syncwarp();
      if ((threadIdx.x & 31) == 0) st_release(a.progress + g, t1);
    }
  }
}

template <typename Real, bool TRACE, bool VAR = false>
static cudaError_t launch_t(const RolloutArgs& a, int sm_count, cudaStream_t st) {
  using N = NetDefault;
  auto kern = k_rollout<Real, N, TRACE, VAR>;
  const int smem = mlp_smem_bytes<N>(ROLLOUT_BLOCK);
  static b2::PerDeviceOnce once;           // per instantiation; attributes and occupancy are per device
  static std::atomic<int> per_sm_dev[64];
  int per_sm = once.need() ? 0 : per_sm_dev[b2::PerDeviceOnce::dev()].load();
  if (per_sm == 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    int n = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, ROLLOUT_BLOCK, smem);
    if (e != cudaSuccess) return e;
    if (n < 1) return cudaErrorLaunchOutOfResources;
    if (const char* env = getenv("B2_ROLLOUT_CTAS_PER_SM")) {  // experiment knob: fewer resident warps
      const int v = atoi(env);
      if (v >= 1 && v < n) n = v;
    }
    per_sm = n;
    per_sm_dev[b2::PerDeviceOnce::dev()].store(n);
    once.done();
  }
  const int64_t G = (a.n_envs + 31) / 32;
  const int wpb = ROLLOUT_BLOCK / 32;
  // Warps W = 4 * grid.  The no-wait ordering needs every warp to own at least one group's worth of steps
  // (W <= G), and the balance needs the same number of warps on every sub-partition (grid a multiple of the
  // SM count): take the largest such grid that is resident at once.  Tiny batches (G < one warp per
  // sub-partition) run one group per warp.
  int64_t k = G / ((int64_t)sm_count * wpb);
  if (k > per_sm) k = per_sm;
  const int64_t grid = k >= 1 ? (int64_t)sm_count * k : (G + wpb - 1) / wpb;
  cudaError_t e = cudaMemsetAsync(a.progress, 0, sizeof(int) * (size_t)G, st);
  if (e != cudaSuccess) return e;
  kern<<<(int)grid, ROLLOUT_BLOCK, smem, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_rollout(const RolloutArgs& a, int precision, bool trace, int sm_count, cudaStream_t st, int h1,
                           int h2) {
  if (h1 != NetDefault::H1 || h2 != NetDefault::H2) return cudaErrorInvalidValue;
  if (a.env_flags) {   // opt-in env variants: double physics only (b2_create refuses B2_PREC_F32 with env_flags)
    if (precision != B2_PREC_F64) return cudaErrorInvalidValue;
    return trace ? launch_t<double, true, true>(a, sm_count, st) : launch_t<double, false, true>(a, sm_count, st);
  }
  if (precision == B2_PREC_F64)
    return trace ? launch_t<double, true>(a, sm_count, st) : launch_t<double, false>(a, sm_count, st);
  if (precision == B2_PREC_MIXED)
    return trace ? launch_t<MixedPrec, true>(a, sm_count, st) : launch_t<MixedPrec, false>(a, sm_count, st);
  return trace ? launch_t<float, true>(a, sm_count, st) : launch_t<float, false>(a, sm_count, st);
}

}  // namespace b2
