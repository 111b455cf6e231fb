// engine.h -- internal state of one b2_handle and the kernel launch entry points.
// Host C++ only sees POD + cudaStream_t; every kernel lives in its own .cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/b200ppo.h"

namespace b2 {

int set_error(int code, const char* fmt, ...);

// Function attributes (cudaFuncSetAttribute) and occupancy answers are PER DEVICE: a process may hold engines on several
// GPUs (b2_config.device), so "already set" is remembered per call site AND per device ordinal, and safely across host
// threads.  Usage:  static b2::PerDeviceOnce once;  if (once.need()) { ...set attributes...; once.done(); }
struct PerDeviceOnce {
  std::atomic<uint64_t> mask{0};
  static int dev() { int d = 0; if (cudaGetDevice(&d) != cudaSuccess) d = 0; return d & 63; }
  bool need() const { return !((mask.load(std::memory_order_acquire) >> dev()) & 1ull); }
  void done() { mask.fetch_or(1ull << dev(), std::memory_order_release); }
};


// Debug allocator behind every device buffer of the engine.  With B2_GUARD=1 in the environment (read once) an
// allocation gets a 4-KiB guard zone on either side filled with 0xA5 and its payload filled with 0xFF (a NaN for every
// float, -1 for every integer: reading memory nobody wrote shows up in the results); b2_debug_guard_check() verifies
// every guard zone of every live allocation.  This is the memcheck / initcheck stand-in where compute-sanitizer cannot
// be run (tests/test_gpu_guard.py).  Without B2_GUARD it is plain cudaMalloc / cudaFree.
cudaError_t dmalloc_bytes(void** p, size_t bytes);
cudaError_t dfree(void* p);
template <class T>
inline cudaError_t dmalloc(T** p, size_t bytes) { return dmalloc_bytes(reinterpret_cast<void**>(p), bytes); }

// Device-resident trajectory buffer, SoA, flat index k = t * n_envs + e (time-major).
//   obs[c * cap + k]  c < 6          24 B/record
//   action, logp, mu, reward, ret    f32 each
//   flags                            u8
//   vboot[k]                         critic bootstrap value where flags & (TIMEUP|TRUNC)
//   vcur[k]                          V of the current epoch (critic pass -> actor pass scratch)
struct Traj {
  int64_t cap = 0;  // records
  float* obs = nullptr;
  float* action = nullptr;
  float* logp = nullptr;
  float* mu = nullptr;
  float* reward = nullptr;
  float* ret = nullptr;
  float* vboot = nullptr;
  float* vcur = nullptr;
  uint8_t* flags = nullptr;
  double* state_trace = nullptr;  // optional [k][4]
  int64_t n_envs = 0, horizon = 0;  // shape of what is currently stored
};

struct HyperDev {       // passed by value to kernels
  float sigma, inv_var, half_log_norm, loss3;
  float gamma, clip_lo, clip_hi;
  float gae_lambda;     // 1 = the reference's plain discounted returns (src/ppo_nn.cpp:231)
  double sigma_d;
  double lr, beta1, beta2, eps;
};

struct RolloutArgs {
  const float* params;
  double* state;        // [4][n_envs]
  int32_t* ep_step;     // [n_envs]
  float* ep_ret;        // [n_envs] running reward sum of the current episode
  int32_t* ep_count;    // [n_envs] episodes completed in this rollout
  float* ep_log;        // [max_eps][n_envs]
  int max_eps;
  Traj tr;
  const float* forced_action;  // [T][N] or null
  // b2_rollout_host with PINNED destination buffers: the kernel also stores every field straight into host memory (posted
  // PCIe writes, 37 B per record ~ 49 GB/s at the reference shape), so the trajectory is on the host when the kernel ends
  float *h_action, *h_logp, *h_mu, *h_reward;   // each null or [T][N] in pinned host memory
  uint8_t* h_flags;
  float* obs4_aos;             // null, or [T][N][4] device staging of (x, xdot, thetadot, theta): b2_rollout_host_compact
  uint8_t* rflags;             // null, or [T][N] device staging of flags | B2_FLAG_REWARD
  float* obs_aos;              // null, or [T][N][6] (device staging or pinned host): the observations ALSO as records (what MemoryNN holds) for b2_rollout_host,
                               // so that no transpose kernel sits between the simulation and the D2H copy
  int64_t n_envs;
  int T, n_timestep;    // T: horizon of the whole rollout (trajectory index and TRUNC flag use absolute t)
  int t_begin, t_end;   // this launch advances steps [t_begin, t_end) of it
  uint64_t seed, env_offset, step0;
  HyperDev hp;
  int* status;          // bit0: non-finite action seen
  int env_flags;        // B2_ENV_* (0: the shipped environment, hot kernels)
  int* progress;        // [ceil(n_envs/32)] steps of this rollout already published per env group (hand-off word)
};

// Gradient exchange over NVLink peer memory, fused into the reduce + Adam kernel (multi-GPU, one process per GPU; the
// DEFAULT whenever every rank can map every peer, B2_COMM_P2P=0 selects reduce -> ncclAllReduce -> Adam instead; see
// DESIGN.md s5 for the measurements).  Flag-in-data protocol (what
// NCCL calls LL): every rank owns, in memory its peers have mapped (cudaIpc),
//     slots[2][world][pstride] of {value, epoch} pairs (8 bytes, written and read as ONE 8-byte access).
// Step e (parity e & 1), every thread that finishes a gradient element i:
//   1. stores {g_i, e} into slot[rank][i] of EVERY rank -- posted NVLink stores, no fence, no flag, no counter;
//   2. polls its OWN copies slot[r][i] (local memory) until their tag is e, for r = 0..world-1;
//   3. adds the values in RANK ORDER (identical sum on every rank: weights stay bit-identical without a broadcast) and
//      applies Adam to element i.
// A thread waits only for the SAME thread of its peers, which pushes before it waits, so there is no circular wait as
// long as blocks are dispatched in index order; every wait is bounded (status bit 1 instead of a hung GPU): a thread whose
// wait expires raises the bit and does NOT apply Adam to its element, and every later step of the update finds the bit
// raised on entry (px_alive) and neither pushes nor applies, so its peers run into their own bounded waits and stop too;
// the host then reports B2_ERR_COMM (the parameters of the handle are unspecified from then on: reload a checkpoint)
// and clears the bit.  Two parities
// are enough: a rank pushes step e+2 only after its step e+1 kernel ended, i.e. after every peer pushed e+1, i.e. after
// every peer's step-e kernel -- the last reader of that parity -- ended.
//
// More than two ranks: REDUCE-SCATTER + ALL-GATHER shape (`scatter`), same flag-in-data slots.  Element i has an owner,
// rank (i / 64) % world.  A thread pushes {g_i, e} to the OWNER's slot only; the owner's thread for i waits for all
// `world` contributions, adds them in rank order and stores {sum_i, e} into sums[i] of EVERY rank; every thread then
// waits for its own sums[i] and applies Adam (redundantly, so that m, v and the weights stay complete and bit-identical
// on every rank).  Per rank and step that is 2 x 7/8 x P x 8 B on the wire instead of 7 x P x 8 B (1.9 MB instead of
// 7.6 MB for the 2 x 256 nets on 8 GPUs), at the price of a second NVLink hop; at world = 2 the all-to-all push moves
// the same bytes in one hop and stays.  The same thread (same block index) handles element i on every rank, so the
// wait graph is still index-symmetric.
constexpr int B2_MAX_PEERS = 8;
struct PeerXchg {
  int world = 0, rank = 0;      // world <= 1: single GPU, nothing below is used
  unsigned epoch = 0;           // this optimizer step (1-based, same on every rank)
  int pstride = 0;
  int scatter = 0;              // 1: reduce-scatter + all-gather (world > 2), 0: all-to-all push
  int64_t sum_off = 0;          // offset of sums[epoch & 1][pstride] inside every rank's buffer (scatter mode)
  int64_t par_off = 0;          // (epoch & 1) * world * pstride
  uint2* slots[B2_MAX_PEERS] = {};       // rank r's slot array (own entry: local pointer)
  float* out = nullptr;         // summed gradient (pstride floats, local): what b2_get_grads returns
  int* status = nullptr;        // bit1: a peer never arrived (bounded wait expired)
};

struct UpdateArgs {
  PeerXchg px;
  float* params;
  float* adam_m;
  float* adam_v;
  float* grads;         // [P] reduced gradient of this step
  float* partials;      // [grid][PSTRIDE]
  int64_t* adam_step;   // device counter
  double* loss_log;     // [steps][4] loss1, loss2, clip_frac, (spare)
  Traj tr;
  const int32_t* idx;   // minibatch gather indices or null
  int64_t first, count; // positions [first, first+count) of this minibatch (mapped through idx if set)
  int64_t denom;        // samples the means are taken over (global minibatch size across ranks)
  int step_slot;        // row of loss_log
  int grid;
  HyperDev hp;
};

#ifdef __CUDACC__
// steps 1-3 of PeerXchg for gradient element i with rank-local value v, in two halves so that a thread that owns several
// elements can push them all before it waits for any: px_push stores {v, epoch} into this rank's slot on every rank,
// px_pull waits for every rank's slot of element i and returns the rank-order sum
__device__ __forceinline__ int px_owner(const PeerXchg& x, int i) { return (i >> 6) % x.world; }
__device__ __forceinline__ void px_store(uint2* p, float v, unsigned tag) {
  asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(tag) : "memory");
}
__device__ __forceinline__ void px_push(const PeerXchg& x, int i, float v) {
  const int64_t mine = x.par_off + (int64_t)x.rank * x.pstride + i;
  if (x.scatter) { px_store(x.slots[px_owner(x, i)] + mine, v, x.epoch); return; }
  for (int r = 0; r < x.world; r++) px_store(x.slots[r] + mine, v, x.epoch);
}
// false once any wait of this update has expired (on this rank): the rest of the update is abandoned
__device__ __forceinline__ bool px_alive(const PeerXchg& x) { return !(*(volatile const int*)x.status & 2); }
// publish = false: a redundant reader of element i (k_wide_l1r: every row slice needs the summed W1 gradient, one slice
// pushed it) -- it waits and sums like the owner but leaves the second hop's stores to the thread that pushed
__device__ __forceinline__ float px_pull(const PeerXchg& x, int i, bool* ok, bool publish = true) {
  if (x.scatter && px_owner(x, i) != x.rank) {
    // not the owner: wait for the owner's sum of element i in this rank's sums[] (the second hop)
    const uint2* p = x.slots[x.rank] + x.sum_off + i;
    unsigned val, tag;
    long long spins = 0;
    do {
      asm volatile("ld.relaxed.sys.global.v2.b32 {%0, %1}, [%2];" : "=r"(val), "=r"(tag) : "l"(p) : "memory");
      if (tag == x.epoch) break;
      if (++spins > (1ll << 26)) { atomicOr(x.status, 2); *ok = false; break; }
    } while (true);
    const float s = __uint_as_float(val);
    x.out[i] = s;
    return s;
  }
  const uint2* base = x.slots[x.rank] + x.par_off + i;
  // every rank's slot is read at once (one L2 round trip when the peers are ahead, not `world` of them); a slot whose
  // tag is not this step's yet is re-read until it is
  unsigned val[B2_MAX_PEERS], tag[B2_MAX_PEERS];
#pragma unroll
  for (int r = 0; r < B2_MAX_PEERS; r++) {
    val[r] = 0u; tag[r] = x.epoch;
    if (r < x.world)
      asm volatile("ld.relaxed.sys.global.v2.b32 {%0, %1}, [%2];" : "=r"(val[r]), "=r"(tag[r]) : "l"(base + (int64_t)r * x.pstride) : "memory");
  }
  long long spins = 0;
  float s = 0.f;
#pragma unroll
  for (int r = 0; r < B2_MAX_PEERS; r++) {
    if (r < x.world) {
      while (tag[r] != x.epoch) {
        if (++spins > (1ll << 26)) { atomicOr(x.status, 2); *ok = false; break; }   // ~2 s: surface a lost peer instead of hanging the GPU
        asm volatile("ld.relaxed.sys.global.v2.b32 {%0, %1}, [%2];" : "=r"(val[r]), "=r"(tag[r]) : "l"(base + (int64_t)r * x.pstride) : "memory");
      }
      s = r == 0 ? __uint_as_float(val[r]) : __fadd_rn(s, __uint_as_float(val[r]));
    }
  }
  if (x.scatter && *ok && publish)   // the owner publishes the rank-order sum to every peer (its own copy is not read again)
    for (int r = 0; r < x.world; r++)
      if (r != x.rank) px_store(x.slots[r] + x.sum_off + i, s, x.epoch);
  x.out[i] = s;
  return s;
}
__device__ __forceinline__ float px_allreduce_elem(const PeerXchg& x, int i, float v, bool* ok) {
  *ok = px_alive(x);
  if (!*ok) return v;
  px_push(x, i, v);
  return px_pull(x, i, ok);
}
#endif

// ---- launchers (each returns cudaGetLastError()) ----
cudaError_t launch_rollout(const RolloutArgs& a, int precision, bool trace, int sm_count, cudaStream_t st, int hidden1,
                           int hidden2);
cudaError_t launch_act(const float* params, const float* obs_aos, const float* eps, int64_t n, uint64_t seed,
                       uint64_t call, float* action, float* logp, float* mu, const HyperDev& hp, int* status,
                       cudaStream_t st);
cudaError_t launch_value(const float* params, const float* obs_aos, int64_t n, float* v, cudaStream_t st);
cudaError_t launch_aos_to_soa(const float* obs_aos, float* obs_soa, int64_t cap, int64_t n, cudaStream_t st);
cudaError_t launch_soa_to_aos(const float* obs_soa, float* obs_aos, int64_t cap, int64_t n, cudaStream_t st);
// seg_ends / seg_count: scratch of the segment-per-warp path (needed iff returns_needs_segment_list(tr): few envs)
bool returns_needs_segment_list(const Traj& tr);
cudaError_t launch_returns(const float* params, const Traj& tr, const HyperDev& hp, cudaStream_t st, int* n_launches,
                           bool skip_bootstrap, int64_t* seg_ends, unsigned long long* seg_count);
int update_grid_for(int64_t count, int sm_count);
int update_grid_for_tc(int64_t count, int sm_count);                  // ppo_update_tc.cu: the bf16x3 tensor-core kernel
cudaError_t launch_ppo_grad_tc(const UpdateArgs& a, cudaStream_t st);
int update_partial_stride();
cudaError_t launch_ppo_grad(const UpdateArgs& a, cudaStream_t st);
cudaError_t launch_reduce(const UpdateArgs& a, bool apply_adam, cudaStream_t st);
cudaError_t launch_adam(const UpdateArgs& a, cudaStream_t st);
cudaError_t launch_env_step(double* state, int32_t* ep_step, const double* force, int64_t n, int precision,
                            int env_flags, uint8_t* terminated, cudaStream_t st);
cudaError_t launch_env_observe(const double* state, double* obs, double* reward, int64_t n, int env_flags, cudaStream_t st);
cudaError_t launch_env_init(double* state, int64_t n, int env_flags, uint64_t seed, uint64_t env_offset, uint64_t tick,
                            cudaStream_t st);
cudaError_t launch_flush(float* buf, int64_t n, cudaStream_t st);
cudaError_t measure_fma_peak(bool fp64, int sm_count, cudaStream_t st, double* tflops);

}  // namespace b2
