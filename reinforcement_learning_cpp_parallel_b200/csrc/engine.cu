// engine.cu -- the b2_handle and every extern "C" entry point of include/b200ppo.h.
//
// Host side of the engine: owns device memory and one stream per handle, turns the reference's
// per-step learner<->env messaging (src/main.cpp:12-148,190-354) into a handful of kernel launches
// per rollout/update, and never touches the data path on the CPU.  No libtorch, no CPU fallback.
#include "engine.h"
#include "wide.h"

#include <dlfcn.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <fstream>
#include <new>
#include <sstream>

#include "common.cuh"

// ---- NCCL is bound at run time (dlopen) so that the library loads on hosts without it ----
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId_t;
enum { NCCL_FLOAT32 = 7, NCCL_SUM = 0 };
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(ncclUniqueId_t*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_t, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
enum { NCCL_UINT8 = 1, NCCL_INT64 = 4 };
static NcclApi g_nccl;

namespace b2 {

static thread_local char g_err[512] = "";
int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

// ---- debug allocator (engine.h) ----
namespace {
constexpr size_t GUARD_BYTES = 4096;
struct GuardBook {
  std::mutex mu;
  std::map<void*, size_t> live;   // payload pointer -> payload bytes
};
GuardBook& guard_book() { static GuardBook b; return b; }
bool guard_on() {
  static const bool on = [] { const char* e = getenv("B2_GUARD"); return e && e[0] == '1'; }();
  return on;
}
}  // namespace

cudaError_t dmalloc_bytes(void** p, size_t bytes) {
  if (!guard_on()) return cudaMalloc(p, bytes);
  char* base = nullptr;
  const size_t padded = ((bytes + 255) / 256) * 256;
  cudaError_t e = cudaMalloc(&base, padded + 2 * GUARD_BYTES);
  if (e != cudaSuccess) return e;
  if ((e = cudaMemset(base, 0xA5, padded + 2 * GUARD_BYTES)) != cudaSuccess) return e;
  if ((e = cudaMemset(base + GUARD_BYTES, 0xFF, bytes)) != cudaSuccess) return e;   // poison: NaN / -1
  // cudaMemset on device memory returns before it has run, and the engine's streams are non-blocking: without this the
  // poison can land AFTER the first real write on E->stream (seen as a spurious "non-finite loss" in one run out of three)
  if ((e = cudaDeviceSynchronize()) != cudaSuccess) return e;
  *p = base + GUARD_BYTES;
  std::lock_guard<std::mutex> lk(guard_book().mu);
  guard_book().live[*p] = bytes;
  return cudaSuccess;
}
cudaError_t dfree(void* p) {
  if (!p || !guard_on()) return cudaFree(p);
  {
    std::lock_guard<std::mutex> lk(guard_book().mu);
    guard_book().live.erase(p);
  }
  return cudaFree(static_cast<char*>(p) - GUARD_BYTES);
}

static bool load_nccl() {
  if (g_nccl.lib) return true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib) break;
  }
  if (!g_nccl.lib) return false;
  g_nccl.GetUniqueId = (int (*)(ncclUniqueId_t*))dlsym(g_nccl.lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(ncclComm_t*, int, ncclUniqueId_t, int))dlsym(g_nccl.lib, "ncclCommInitRank");
  g_nccl.CommDestroy = (int (*)(ncclComm_t))dlsym(g_nccl.lib, "ncclCommDestroy");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t))dlsym(g_nccl.lib, "ncclAllReduce");
  g_nccl.AllGather = (int (*)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t))dlsym(g_nccl.lib, "ncclAllGather");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(g_nccl.lib, "ncclGetErrorString");
  return g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.AllReduce;
}

}  // namespace b2

using namespace b2;

constexpr int B2_MAX_CHUNKS = 32;

struct b2_engine {
  b2_config cfg;
  int P = 0, pstride = 0, sm_count = 148;
  // obs_vars = 4 (getobs(4), include/cart_pole.h:120-127): the caller's nets are S = 4 wide; on the device they live
  // inside the S = 6 layout with the two extra W1 columns, their Adam state and the stored cos/sin planes all exactly
  // zero -- 0 * x adds nothing to any FMA chain and a zero gradient leaves Adam's m, v and p untouched, so every
  // kernel computes the S = 4 net bit for bit.  S_ext / P_ext are the shapes at the C ABI.
  int S_ext = OBS, P_ext = 0;
  cudaStream_t stream = nullptr, copy_stream = nullptr;
  cudaEvent_t chunk_ev[B2_MAX_CHUNKS] = {}, copy_done = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // learner state
  float *params = nullptr, *adam_m = nullptr, *adam_v = nullptr, *grads = nullptr, *partials = nullptr;
  int64_t* adam_step = nullptr;
  double* loss_log = nullptr;
  int loss_log_cap = 0;
  std::vector<double> last_losses;  // [steps][4]
  int partial_rows = 0;
  // env state
  double* state = nullptr;
  int32_t *ep_step = nullptr, *ep_count = nullptr;
  uint8_t* env_term = nullptr;   // [n_envs] what advance() returned in the last b2_env_step
  uint64_t reset_calls = 0;      // tick of the random initial state drawn by b2_env_reset
  float *ep_ret = nullptr, *ep_log = nullptr;
  int* progress = nullptr;
  int max_eps = 0;
  uint64_t global_step = 0;
  int64_t total_env_steps = 0;
  // trajectories
  Traj tr;
  bool trace = false;
  bool have_traj = false;
  float* forced = nullptr;
  int64_t forced_cap = 0;
  // staging for the host-buffer entry points
  float* stage = nullptr;
  int64_t stage_floats = 0;
  uint8_t* stage_u8 = nullptr;
  int64_t stage_u8_n = 0;
  int* status = nullptr;
  int* status_host = nullptr;  // pinned
  // scratch of the segment-per-warp returns path (returns.cu): segment-end record indices + their count
  int64_t* seg_ends = nullptr;
  int64_t seg_cap = 0;
  unsigned long long* seg_count = nullptr;
  int last_update_path = 0;       // 0 FP32 FMA kernel, 1 bf16x3 tensor-core kernel, 2 wide nets (tcgen05)
  int64_t* denom_dev = nullptr;   // [2] global sample counts of a regular / the last minibatch (all-reduced over ranks)
  uint64_t act_calls = 0;
  int64_t launches = 0;
  float* flush_buf = nullptr;
  int64_t flush_n = 0;
  // comm
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  // gradient exchange over NVLink peer memory (engine.h PeerXchg); NCCL stays the fallback when the mapping fails
  bool p2p = false;
  char* xchg = nullptr;                    // slots[2][world][pstride] of {value, epoch}, mapped by every peer (cudaIpc)
  void* peer_base[B2_MAX_PEERS] = {};      // cudaIpcOpenMemHandle results (own entry: xchg)
  int px_epoch = 0;
  // wide hidden layers (tensor-core update path, wide.h); null for the reference's 64/32 nets
  WideNet* wide = nullptr;
  NetDims dims;
  HyperDev hp;
};

namespace {

#define CK(expr) B2_CUDA_CHECK(expr)
#define NEED(h) \
  if (!(h)) return set_error(B2_ERR_INVALID, "null handle")

void make_hyper(b2_engine* E) {
  const b2_config& c = E->cfg;
  HyperDev& h = E->hp;
  const float var = (float)(c.action_std * c.action_std);        // ppo_nn.h:124-126 action_var is f32
  h.sigma = (float)c.action_std;
  h.sigma_d = c.action_std;
  h.inv_var = 1.0f / var;                                        // src/ppo_nn.cpp:24
  h.half_log_norm = (logf(var) + (float)(log(2 * M_PI) * 1)) / 2;  // :42-44
  const float v1 = (float)pow(2 * M_PI * M_E, 1);                // :56 (Q8: (2 pi e)^k, not its log)
  h.loss3 = (float)(-0.01 * (double)(0.5f * (logf(var) + v1)));  // :64, :289
  h.gamma = (float)c.gamma;
  h.gae_lambda = c.gae_lambda > 0 ? (float)c.gae_lambda : 1.0f;   // 0 = unset = the reference's plain returns
  h.clip_lo = (float)(1 - c.eps_clip);
  h.clip_hi = (float)(1 + c.eps_clip);
  h.lr = c.lr; h.beta1 = c.beta1; h.beta2 = c.beta2; h.eps = 1e-8;  // libtorch AdamOptions default eps
}

int ensure_traj(b2_engine* E, int64_t cap) {
  Traj& t = E->tr;
  if (cap <= t.cap && (!E->trace || t.state_trace)) return B2_OK;
  if (cap > t.cap) {
    dfree(t.obs); dfree(t.action); dfree(t.logp); dfree(t.mu); dfree(t.reward);
    dfree(t.ret); dfree(t.vboot); dfree(t.vcur); dfree(t.flags); dfree(t.state_trace);
    t = Traj();
    t.cap = cap;
    CK(dmalloc(&t.obs, sizeof(float) * OBS * cap));
    CK(dmalloc(&t.action, sizeof(float) * cap));
    CK(dmalloc(&t.logp, sizeof(float) * cap));
    CK(dmalloc(&t.mu, sizeof(float) * cap));
    CK(dmalloc(&t.reward, sizeof(float) * cap));
    CK(dmalloc(&t.ret, sizeof(float) * cap));
    CK(dmalloc(&t.vboot, sizeof(float) * cap));
    CK(dmalloc(&t.vcur, sizeof(float) * cap));
    CK(dmalloc(&t.flags, cap));
    E->have_traj = false;
  }
  if (E->trace && !t.state_trace) CK(dmalloc(&t.state_trace, sizeof(double) * 4 * t.cap));
  return B2_OK;
}

int ensure_stage(b2_engine* E, int64_t floats, int64_t bytes_u8) {
  if (floats > E->stage_floats) {
    dfree(E->stage);
    E->stage = nullptr;
    CK(dmalloc(&E->stage, sizeof(float) * floats));
    E->stage_floats = floats;
  }
  if (bytes_u8 > E->stage_u8_n) {
    dfree(E->stage_u8);
    E->stage_u8 = nullptr;
    CK(dmalloc(&E->stage_u8, bytes_u8));
    E->stage_u8_n = bytes_u8;
  }
  return B2_OK;
}

int check_status(b2_engine* E, const char* where) {
  CK(cudaMemcpyAsync(E->status_host, E->status, sizeof(int), cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  if (*E->status_host & 4) {   // rollout.cu: a hand-off wait of the persistent rollout kernel expired
    cudaMemsetAsync(E->status, 0, sizeof(int), E->stream);
    return set_error(B2_ERR_STATE, "%s: the rollout kernel's hand-off wait expired (its warps were not all resident: is another "
                                   "stream or process holding SMs?); the trajectory of this rollout is incomplete", where);
  }
  if (*E->status_host & 1) {
    cudaMemsetAsync(E->status, 0, sizeof(int), E->stream);
    return set_error(B2_ERR_NONFINITE, "%s: non-finite action on device (cf. src/main.cpp:50-55)", where);
  }
  return B2_OK;
}

// K2 on whatever E->tr holds: bootstrap values at the flagged records, then the reverse scan.
int run_returns(b2_engine* E) {
  int nl = 0;
  const int64_t B = E->tr.n_envs * E->tr.horizon;
  if (E->wide) {
    int rc = wide_bootstrap(E->wide, E->params, E->tr, E->stream, &E->launches);
    if (rc) return rc;
    if (E->hp.gae_lambda < 1.0f) {   // lambda-returns need V at every record, not only at the flagged ones
      cudaError_t ce = launch_wide_forward(E->dims, E->params + E->dims.SIZE, E->tr.obs, E->tr.cap, 1, nullptr, B, 0, E->tr.vcur, E->stream);
      if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "values: %s", cudaGetErrorString(ce));
      E->launches += 1;
    }
  }
  if (returns_needs_segment_list(E->tr)) {
    if (!E->seg_count) CK(dmalloc(&E->seg_count, sizeof *E->seg_count));
    if (B > E->seg_cap) {
      dfree(E->seg_ends);
      E->seg_ends = nullptr;
      E->seg_cap = 0;
      CK(dmalloc(&E->seg_ends, sizeof(int64_t) * (size_t)B));
      E->seg_cap = B;
    }
  }
  cudaError_t ce = launch_returns(E->params, E->tr, E->hp, E->stream, &nl, E->wide != nullptr, E->seg_ends, E->seg_count);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "returns: %s", cudaGetErrorString(ce));
  E->launches += nl;
  return B2_OK;
}

// B2_UPDATE_TC=0 / 1 forces the FP32 / the tensor-core kernel (A/B measurements, parity tests of either on any shape)
bool use_tensor_update(int64_t global_count) {
  const char* e = getenv("B2_UPDATE_TC");   // read per update: tests flip it between engines of one process
  if (e && (e[0] == '0' || e[0] == '1')) return e[0] == '1';
  return global_count >= 32768;
}

int setup_peer_exchange(b2_engine* E);
void next_peer_step(b2_engine* E, UpdateArgs& a);
inline size_t xchg_slot_bytes(const b2_engine* E) {
  // slots[2][world][pstride] + sums[2][pstride] (engine.h PeerXchg)
  return ((sizeof(uint2) * 2 * ((size_t)E->world + 1) * (size_t)E->pstride + 255) / 256) * 256;
}

// K2 + K optimizer steps on whatever E->tr currently holds.
int run_update(b2_engine* E, b2_update_stats* stats) {
  Traj& t = E->tr;
  const int64_t B = t.n_envs * t.horizon;
  if (B <= 0) return set_error(B2_ERR_STATE, "update: empty trajectory");
  const int K = E->cfg.k_epochs, M = std::max(1, E->cfg.n_minibatches);
  const int steps = K * M;
  cudaError_t ce;
  int rc0 = run_returns(E);  // once, at entry (:207-236)
  if (rc0) return rc0;
  if (steps > E->loss_log_cap) {
    dfree(E->loss_log);
    E->loss_log = nullptr;
    CK(dmalloc(&E->loss_log, sizeof(double) * 4 * steps));
    E->loss_log_cap = steps;
  }
  CK(cudaMemsetAsync(E->loss_log, 0, sizeof(double) * 4 * steps, E->stream));
  const int64_t per = B / M;
  int rows_needed = E->wide ? 0 : std::max(update_grid_for(per + (B % M), E->sm_count), update_grid_for_tc(per + (B % M), E->sm_count));
  if (rows_needed > E->partial_rows) {
    dfree(E->partials);
    E->partials = nullptr;
    CK(dmalloc(&E->partials, sizeof(float) * (size_t)rows_needed * E->pstride));
    E->partial_rows = rows_needed;
  }
  // The means are over the GLOBAL (mini)batch (src/ppo_nn.cpp:302: one mean over the merged memory).  Ranks may hold
  // different numbers of samples (sharding.shard_envs hands the remainder to the low ranks), so the denominators are
  // all-reduced: [0] a regular minibatch, [1] the last one (which takes the remainder of B / M).
  int64_t denom_g[2] = {per, B - (int64_t)(M - 1) * per};
  if (E->world > 1) {
    if (!E->denom_dev) CK(dmalloc(&E->denom_dev, 2 * sizeof(int64_t)));
    CK(cudaMemcpyAsync(E->denom_dev, denom_g, sizeof denom_g, cudaMemcpyHostToDevice, E->stream));
    int ne = g_nccl.AllReduce(E->denom_dev, E->denom_dev, 2, NCCL_INT64, NCCL_SUM, E->comm, E->stream);
    if (ne != 0) return set_error(B2_ERR_COMM, "ncclAllReduce(sample counts): %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(ne) : "?");
    CK(cudaMemcpyAsync(denom_g, E->denom_dev, sizeof denom_g, cudaMemcpyDeviceToHost, E->stream));
    CK(cudaStreamSynchronize(E->stream));
  }
  UpdateArgs a{};
  a.params = E->params; a.adam_m = E->adam_m; a.adam_v = E->adam_v; a.grads = E->grads;
  a.partials = E->partials; a.adam_step = E->adam_step; a.loss_log = E->loss_log;
  a.tr = t; a.idx = nullptr; a.hp = E->hp;
  int slot = 0;
  for (int k = 0; k < K; k++) {           // :245
    for (int m = 0; m < M; m++, slot++) { // M == 1 is the reference (full batch, Q9)
      a.first = (int64_t)m * per;
      a.count = (m == M - 1) ? (B - a.first) : per;
      a.denom = denom_g[m == M - 1];      // samples of this minibatch over ALL ranks
      // default nets: the FP32 FMA kernel for reference-sized batches, the bf16x3 tensor-core kernel once the GLOBAL
      // minibatch is a real GEMM (>= 32,768 rows; decided on the global count so that every sharding takes the same path)
      const bool tc = !E->wide && use_tensor_update(a.denom);
      E->last_update_path = E->wide ? 2 : (tc ? 1 : 0);
      a.grid = tc ? update_grid_for_tc(a.count, E->sm_count) : update_grid_for(a.count, E->sm_count);
      a.step_slot = slot;
      // several GPUs: the reduce + Adam kernel exchanges every gradient element with its peers over NVLink peer memory
      // (engine.h PeerXchg); reduce -> ncclAllReduce -> Adam only if the peer mapping was refused or switched off
      if (E->world > 1 && E->p2p) next_peer_step(E, a);
      if (E->wide) {
        // (single GPU: the reduce + Adam of a step rides in the layer-1 kernel of the next one; the last step flushes itself)
        int rc = wide_step(E->wide, a, E->world == 1 || E->p2p, slot + 1 < steps, E->stream, &E->launches);
        if (rc) return rc;
        if (E->world > 1 && !E->p2p) {
          int ne = g_nccl.AllReduce(E->grads, E->grads, (size_t)E->pstride, NCCL_FLOAT32, NCCL_SUM, E->comm, E->stream);
          if (ne != 0) return set_error(B2_ERR_COMM, "ncclAllReduce: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(ne) : "?");
          if ((ce = launch_wide_adam(E->wide, a, E->stream)) != cudaSuccess) return set_error(B2_ERR_CUDA, "adam: %s", cudaGetErrorString(ce));
          E->launches += 2;
        }
        continue;
      }
      if ((ce = tc ? launch_ppo_grad_tc(a, E->stream) : launch_ppo_grad(a, E->stream)) != cudaSuccess) return set_error(B2_ERR_CUDA, "ppo_grad: %s", cudaGetErrorString(ce));
      if (E->world == 1 || E->p2p) {   // one kernel: reduce (+ peer-memory exchange) + Adam
        if ((ce = launch_reduce(a, true, E->stream)) != cudaSuccess) return set_error(B2_ERR_CUDA, "reduce: %s", cudaGetErrorString(ce));
        E->launches += 2;
      } else {
        if ((ce = launch_reduce(a, false, E->stream)) != cudaSuccess) return set_error(B2_ERR_CUDA, "reduce: %s", cudaGetErrorString(ce));
        int ne = g_nccl.AllReduce(E->grads, E->grads, (size_t)E->pstride, NCCL_FLOAT32, NCCL_SUM, E->comm, E->stream);
        if (ne != 0) return set_error(B2_ERR_COMM, "ncclAllReduce: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(ne) : "?");
        if ((ce = launch_adam(a, E->stream)) != cudaSuccess) return set_error(B2_ERR_CUDA, "adam: %s", cudaGetErrorString(ce));
        E->launches += 3;
      }
    }
  }
  E->last_losses.assign((size_t)4 * steps, 0.0);
  int64_t step_host = 0;
  CK(cudaMemcpyAsync(E->last_losses.data(), E->loss_log, sizeof(double) * 4 * steps, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaMemcpyAsync(&step_host, E->adam_step, sizeof(int64_t), cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  bool finite = true;
  for (double v : E->last_losses) finite &= std::isfinite(v);
  if (stats) {
    const double* L = &E->last_losses[(size_t)4 * (steps - 1)];
    stats->loss1 = L[0]; stats->loss2 = L[1]; stats->loss3 = (double)E->hp.loss3;
    stats->loss = L[0] + L[1] + (double)E->hp.loss3;  // :290, :302
    stats->clip_frac = L[2];
    stats->samples = B; stats->adam_step = step_host; stats->optimizer_steps = steps;
  }
  if (E->world > 1 && E->p2p) {
    int st = 0;
    CK(cudaMemcpyAsync(&st, E->status, sizeof(int), cudaMemcpyDeviceToHost, E->stream));
    CK(cudaStreamSynchronize(E->stream));
    if (st & 2) {
      // reported once: clear the bit so that the next update starts clean (the kernels of this one stopped pushing and
      // applying as soon as it was raised, engine.h px_alive)
      st &= ~2;
      CK(cudaMemcpyAsync(E->status, &st, sizeof(int), cudaMemcpyHostToDevice, E->stream));
      CK(cudaStreamSynchronize(E->stream));
      return set_error(B2_ERR_COMM, "update: a peer never published its gradient (peer-memory exchange timed out); the update was "
                                    "abandoned part-way, reload the parameters (b2_load_checkpoint / b2_set_params) before continuing");
    }
  }
  if (!finite) return set_error(B2_ERR_NONFINITE, "update: non-finite loss");
  return B2_OK;
}

// Map every peer's exchange buffer (engine.h PeerXchg).  The 64-byte cudaIpc handles travel through one ncclAllGather
// on the communicator that was just created; ranks then agree (all-reduce of a success flag) on whether EVERY mapping
// worked, so either all of them exchange gradients over peer memory or all of them keep calling ncclAllReduce.
// Default when every rank can map every peer; B2_COMM_P2P=0 in the environment forces ncclAllReduce (A/B measurements).
// Measured: 43.7 vs 47.1 ms per configs[4] update on 8 x B200, 40.1 vs 41.0 on 2.
int setup_peer_exchange(b2_engine* E) {
  E->p2p = false;
  const char* env = getenv("B2_COMM_P2P");
  const bool want = !(env && env[0] == '0') && E->world <= B2_MAX_PEERS && g_nccl.AllGather;
  const size_t buf_bytes = xchg_slot_bytes(E);
  CK(cudaMalloc(&E->xchg, buf_bytes + 256));   // plain: cudaIpcGetMemHandle wants the allocation base
  CK(cudaMemset(E->xchg, 0, buf_bytes + 256));
  CK(cudaDeviceSynchronize());   // (legacy-stream memset; the peers write into this buffer from THEIR streams, see b2_create)
  struct Rec { cudaIpcMemHandle_t h; int ok; int pad[3]; };
  static_assert(sizeof(Rec) == 80, "record size");
  Rec mine{};
  mine.ok = want && cudaIpcGetMemHandle(&mine.h, E->xchg) == cudaSuccess;
  cudaGetLastError();
  Rec* d_recs = nullptr;
  CK(dmalloc(&d_recs, sizeof(Rec) * E->world));
  CK(cudaMemcpyAsync(d_recs + E->rank, &mine, sizeof(Rec), cudaMemcpyHostToDevice, E->stream));
  std::vector<Rec> recs((size_t)E->world);
  int ne = g_nccl.AllGather ? g_nccl.AllGather(d_recs + E->rank, d_recs, sizeof(Rec), NCCL_UINT8, E->comm, E->stream) : 1;
  if (ne != 0) { dfree(d_recs); return B2_OK; }          // no exchange possible: NCCL all-reduce path
  CK(cudaMemcpyAsync(recs.data(), d_recs, sizeof(Rec) * E->world, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  bool ok = true;
  for (int r = 0; r < E->world; r++) ok = ok && recs[(size_t)r].ok;
  if (ok) {
    for (int r = 0; r < E->world && ok; r++) {
      if (r == E->rank) { E->peer_base[r] = E->xchg; continue; }
      if (cudaIpcOpenMemHandle(&E->peer_base[r], recs[(size_t)r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        E->peer_base[r] = nullptr;
        ok = false;
      }
    }
  }
  // agreement: sum of per-rank success flags must equal world
  float okf = ok ? 1.0f : 0.0f, sum = 0.0f;
  float* d_ok = reinterpret_cast<float*>(d_recs);
  CK(cudaMemcpyAsync(d_ok, &okf, sizeof(float), cudaMemcpyHostToDevice, E->stream));
  ne = g_nccl.AllReduce(d_ok, d_ok, 1, NCCL_FLOAT32, NCCL_SUM, E->comm, E->stream);
  if (ne == 0) {
    CK(cudaMemcpyAsync(&sum, d_ok, sizeof(float), cudaMemcpyDeviceToHost, E->stream));
    CK(cudaStreamSynchronize(E->stream));
  }
  dfree(d_recs);
  E->p2p = ne == 0 && (int)(sum + 0.5f) == E->world;
  if (!E->p2p)
    for (int r = 0; r < E->world; r++)
      if (r != E->rank && E->peer_base[r]) { cudaIpcCloseMemHandle(E->peer_base[r]); E->peer_base[r] = nullptr; }
  E->px_epoch = 0;
  return B2_OK;
}

// the PeerXchg of the next optimizer step (layout: slots[2][world][pstride] of {value, epoch})
void next_peer_step(b2_engine* E, UpdateArgs& a) {
  PeerXchg& x = a.px;
  x.world = E->world; x.rank = E->rank; x.pstride = E->pstride;
  x.epoch = (unsigned)++E->px_epoch;
  x.par_off = (int64_t)(x.epoch & 1u) * E->world * E->pstride;
  // more than two ranks: reduce-scatter + all-gather shape (B2_COMM_SCATTER=0 / 1 forces either, A/B measurements)
  const char* sc = getenv("B2_COMM_SCATTER");
  x.scatter = (sc && (sc[0] == '0' || sc[0] == '1')) ? (sc[0] == '1') : (E->world > 2);
  x.sum_off = (int64_t)2 * E->world * E->pstride + (int64_t)(x.epoch & 1u) * E->pstride;
  for (int r = 0; r < E->world; r++) x.slots[r] = reinterpret_cast<uint2*>(E->peer_base[r]);
  x.out = E->grads;
  x.status = E->status;
}

// host [n][S_ext] -> device AoS [n][6] (zero-filled tail when S_ext = 4) and back
int h2d_obs(b2_engine* E, float* dst, const float* src, int64_t n, cudaStream_t st) {
  if (E->S_ext == OBS) {
    CK(cudaMemcpyAsync(dst, src, sizeof(float) * OBS * n, cudaMemcpyHostToDevice, st));
  } else {
    CK(cudaMemsetAsync(dst, 0, sizeof(float) * OBS * n, st));
    CK(cudaMemcpy2DAsync(dst, sizeof(float) * OBS, src, sizeof(float) * E->S_ext, sizeof(float) * E->S_ext, (size_t)n,
                         cudaMemcpyHostToDevice, st));
  }
  return B2_OK;
}
int d2h_obs(b2_engine* E, float* dst, const float* src, int64_t n, cudaStream_t st) {
  if (E->S_ext == OBS) {
    CK(cudaMemcpyAsync(dst, src, sizeof(float) * OBS * n, cudaMemcpyDeviceToHost, st));
  } else {
    CK(cudaMemcpy2DAsync(dst, sizeof(float) * E->S_ext, src, sizeof(float) * OBS, sizeof(float) * E->S_ext, (size_t)n,
                         cudaMemcpyDeviceToHost, st));
  }
  return B2_OK;
}
// parameter-shaped vectors (weights, Adam m / v, gradients): caller layout (S_ext) <-> device layout (S = 6)
void widen_params(const b2_engine* E, const float* ext, float* in) {
  const NetDims de = NetDims::make(E->S_ext, E->dims.H1, E->dims.H2), &di = E->dims;
  for (int net = 0; net < 2; net++) {
    const float* e = ext + (size_t)net * de.SIZE;
    float* d = in + (size_t)net * di.SIZE;
    for (int j = 0; j < di.H1; j++)
      for (int c = 0; c < di.S; c++) d[di.oW1 + j * di.S + c] = c < de.S ? e[de.oW1 + j * de.S + c] : 0.0f;
    memcpy(d + di.ob1, e + de.ob1, sizeof(float) * (size_t)(di.SIZE - di.ob1));
  }
}
void narrow_params(const b2_engine* E, const float* in, float* ext) {
  const NetDims de = NetDims::make(E->S_ext, E->dims.H1, E->dims.H2), &di = E->dims;
  for (int net = 0; net < 2; net++) {
    float* e = ext + (size_t)net * de.SIZE;
    const float* d = in + (size_t)net * di.SIZE;
    for (int j = 0; j < di.H1; j++)
      for (int c = 0; c < de.S; c++) e[de.oW1 + j * de.S + c] = d[di.oW1 + j * di.S + c];
    memcpy(e + de.ob1, d + di.ob1, sizeof(float) * (size_t)(di.SIZE - di.ob1));
  }
}
int put_vec(b2_engine* E, float* dev, const float* host) {
  std::vector<float> tmp;
  if (E->S_ext != OBS) { tmp.resize((size_t)E->P); widen_params(E, host, tmp.data()); host = tmp.data(); }
  CK(cudaMemcpyAsync(dev, host, sizeof(float) * E->P, cudaMemcpyHostToDevice, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}
int get_vec(b2_engine* E, float* host, const float* dev) {
  if (E->S_ext == OBS) {
    CK(cudaMemcpyAsync(host, dev, sizeof(float) * E->P, cudaMemcpyDeviceToHost, E->stream));
    CK(cudaStreamSynchronize(E->stream));
    return B2_OK;
  }
  std::vector<float> tmp((size_t)E->P);
  CK(cudaMemcpyAsync(tmp.data(), dev, sizeof(float) * E->P, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  narrow_params(E, tmp.data(), host);
  return B2_OK;
}

int upload_batch(b2_engine* E, const b2_batch* b) {
  if (!b || !b->obs || !b->reward || !b->flags) return set_error(B2_ERR_INVALID, "batch: null field");
  const int64_t B = b->n_envs * b->horizon;
  if (B <= 0) return set_error(B2_ERR_INVALID, "batch: empty");
  int rc = ensure_traj(E, B);
  if (rc) return rc;
  if ((rc = ensure_stage(E, B * OBS, 0))) return rc;
  Traj& t = E->tr;
  if ((rc = h2d_obs(E, E->stage, b->obs, B, E->stream))) return rc;
  cudaError_t ce = launch_aos_to_soa(E->stage, t.obs, t.cap, B, E->stream);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "aos_to_soa: %s", cudaGetErrorString(ce));
  E->launches += 1;
  if (b->action) CK(cudaMemcpyAsync(t.action, b->action, sizeof(float) * B, cudaMemcpyHostToDevice, E->stream));
  if (b->logp) CK(cudaMemcpyAsync(t.logp, b->logp, sizeof(float) * B, cudaMemcpyHostToDevice, E->stream));
  CK(cudaMemcpyAsync(t.reward, b->reward, sizeof(float) * B, cudaMemcpyHostToDevice, E->stream));
  CK(cudaMemcpyAsync(t.flags, b->flags, B, cudaMemcpyHostToDevice, E->stream));
  t.n_envs = b->n_envs;
  t.horizon = b->horizon;
  E->have_traj = true;
  return B2_OK;
}

}  // namespace

extern "C" {

int b2_abi_version(void) { return B2_ABI_VERSION; }
const char* b2_last_error(void) { return g_err; }
int b2_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

static int create_body(b2_engine* E, const b2_config* cfg, bool is_default_net);

int b2_create(const b2_config* cfg, b2_handle* out) {
  if (!cfg || !out) return set_error(B2_ERR_INVALID, "b2_create: null argument");
  *out = nullptr;
  if (cfg->control_vars != 1) return set_error(B2_ERR_INVALID, "control_vars must be 1 (src/ppo_nn.cpp:107)");
  if (cfg->obs_vars != 6 && cfg->obs_vars != 4) return set_error(B2_ERR_INVALID, "obs_vars must be 6 or 4 (include/cart_pole.h:120-131)");
  const bool is_default_net = cfg->hidden1 == NetDefault::H1 && cfg->hidden2 == NetDefault::H2;
  if (!is_default_net && (cfg->hidden1 < 64 || cfg->hidden2 < 64 || cfg->hidden1 > 256 || cfg->hidden2 > 256 || cfg->hidden1 % 64 || cfg->hidden2 % 64))
    return set_error(B2_ERR_INVALID, "hidden widths %dx%d not built: 64x32 (reference) or both in {64,128,192,256}", cfg->hidden1, cfg->hidden2);
  if (cfg->n_envs < 1 || cfg->horizon < 1 || cfg->n_timestep < 1) return set_error(B2_ERR_INVALID, "n_envs, horizon, n_timestep must be >= 1");
  if (cfg->precision != B2_PREC_F64 && cfg->precision != B2_PREC_F32 && cfg->precision != B2_PREC_MIXED) return set_error(B2_ERR_INVALID, "unknown precision %d", cfg->precision);
  if (cfg->env_flags & ~(B2_ENV_SWINGUP | B2_ENV_RANDOM_INIT | B2_ENV_EARLY_TERM)) return set_error(B2_ERR_INVALID, "unknown env_flags 0x%x", cfg->env_flags);
  if (cfg->env_flags && cfg->precision != B2_PREC_F64) return set_error(B2_ERR_INVALID, "env variants (env_flags 0x%x) run the double-precision physics only", cfg->env_flags);
  if (cfg->k_epochs < 1 || cfg->n_minibatches < 1) return set_error(B2_ERR_INVALID, "k_epochs and n_minibatches must be >= 1");
  if (!(cfg->action_std > 0)) return set_error(B2_ERR_INVALID, "action_std must be > 0");
  if (!(cfg->gae_lambda >= 0.0 && cfg->gae_lambda <= 1.0)) return set_error(B2_ERR_INVALID, "gae_lambda must be in (0, 1] (0 = unset = 1, the reference's returns)");
  int ndev = b2_device_count();
  if (ndev <= 0) return set_error(B2_ERR_NO_DEVICE, "no CUDA device: this engine has no CPU path");
  if (cfg->device < 0 || cfg->device >= ndev) return set_error(B2_ERR_INVALID, "device %d out of range (%d visible)", cfg->device, ndev);
  CK(cudaSetDevice(cfg->device));
  b2_engine* E = new (std::nothrow) b2_engine();
  if (!E) return set_error(B2_ERR_INVALID, "out of host memory");
  // any failure past this point releases the engine and everything it allocated so far (b2_destroy keeps the message)
  const int rc = create_body(E, cfg, is_default_net);
  if (rc) { b2_destroy(E); return rc; }
  *out = E;
  return B2_OK;
}

static int create_body(b2_engine* E, const b2_config* cfg, bool is_default_net) {
  E->cfg = *cfg;
  E->dims = NetDims::make(OBS, cfg->hidden1, cfg->hidden2);
  E->P = E->dims.P;
  E->S_ext = cfg->obs_vars;
  E->P_ext = NetDims::make(E->S_ext, cfg->hidden1, cfg->hidden2).P;
  E->pstride = is_default_net ? update_partial_stride() : wide_pstride(cfg->hidden1, cfg->hidden2);
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, cfg->device));
  E->sm_count = prop.multiProcessorCount;
  make_hyper(E);
  CK(cudaStreamCreateWithFlags(&E->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&E->ev0));
  CK(cudaEventCreate(&E->ev1));
  const int64_t NE = cfg->n_envs;
  CK(dmalloc(&E->params, sizeof(float) * E->P));
  CK(dmalloc(&E->adam_m, sizeof(float) * E->P));
  CK(dmalloc(&E->adam_v, sizeof(float) * E->P));
  CK(dmalloc(&E->grads, sizeof(float) * E->pstride));
  CK(dmalloc(&E->adam_step, sizeof(int64_t)));
  CK(dmalloc(&E->status, sizeof(int)));
  CK(cudaMallocHost(&E->status_host, sizeof(int)));
  CK(cudaMemset(E->params, 0, sizeof(float) * E->P));
  CK(cudaMemset(E->adam_m, 0, sizeof(float) * E->P));
  CK(cudaMemset(E->adam_v, 0, sizeof(float) * E->P));
  CK(cudaMemset(E->grads, 0, sizeof(float) * E->pstride));
  CK(cudaMemset(E->adam_step, 0, sizeof(int64_t)));
  CK(cudaMemset(E->status, 0, sizeof(int)));
  // episodes one env can finish inside one rollout: TIMEUP every n_timestep steps; with early termination every step
  E->max_eps = (cfg->env_flags & B2_ENV_EARLY_TERM) ? cfg->horizon + 1 : cfg->horizon / cfg->n_timestep + 2;
  CK(dmalloc(&E->env_term, (size_t)cfg->n_envs));
  CK(cudaMemset(E->env_term, 0, (size_t)cfg->n_envs));
  CK(dmalloc(&E->state, sizeof(double) * 4 * NE));
  CK(dmalloc(&E->ep_step, sizeof(int32_t) * NE));
  CK(dmalloc(&E->ep_count, sizeof(int32_t) * NE));
  CK(dmalloc(&E->ep_ret, sizeof(float) * NE));
  CK(dmalloc(&E->ep_log, sizeof(float) * NE * E->max_eps));
  CK(dmalloc(&E->progress, sizeof(int) * ((NE + 31) / 32)));
  CK(cudaMemset(E->state, 0, sizeof(double) * 4 * NE));   // CartPole::reset: u = 0 (cart_pole.h:86)
  // Every cudaMemset of this function runs on the legacy default stream, and E->stream is NON-BLOCKING: nothing orders them
  // against the engine's kernels.  Without this wait the zero fill above could land AFTER the random initial states written
  // below (seen once in ~10 runs of test_random_init_reset_and_sharding as shards that disagree with the full engine).
  CK(cudaDeviceSynchronize());
  if (cfg->env_flags & B2_ENV_RANDOM_INIT) {
    cudaError_t ce = launch_env_init(E->state, NE, cfg->env_flags, cfg->seed, cfg->env_offset, 0, E->stream);
    if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "env init: %s", cudaGetErrorString(ce));
    E->launches += 1;
  }
  CK(cudaMemset(E->ep_step, 0, sizeof(int32_t) * NE));
  CK(cudaMemset(E->ep_count, 0, sizeof(int32_t) * NE));
  CK(cudaMemset(E->ep_ret, 0, sizeof(float) * NE));
  int rc = ensure_traj(E, NE * cfg->horizon);
  if (rc) return rc;
  if (!is_default_net) {
    const int64_t B0 = NE * cfg->horizon;
    rc = wide_create(&E->wide, cfg->hidden1, cfg->hidden2, (B0 + cfg->n_minibatches - 1) / cfg->n_minibatches + cfg->n_minibatches, E->sm_count);
    if (rc) return rc;
  }
  CK(cudaDeviceSynchronize());   // the remaining legacy-stream memsets (episode counters, wide_create's buffers) before anything on E->stream
  return b2_init_params(E, cfg->seed);
}

int b2_destroy(b2_handle E) {
  if (!E) return B2_OK;
  cudaSetDevice(E->cfg.device);
  if (E->stream) cudaStreamSynchronize(E->stream);
  if (E->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(E->comm);
  wide_destroy(E->wide);
  Traj& t = E->tr;
  void* ptrs[] = {E->params, E->adam_m, E->adam_v, E->grads, E->partials, E->adam_step, E->loss_log, E->state,
                  E->ep_step, E->ep_count, E->ep_ret, E->ep_log, t.obs, t.action, t.logp, t.mu, t.reward, t.ret,
                  t.vboot, t.vcur, t.flags, t.state_trace, E->forced, E->stage, E->stage_u8, E->status, E->flush_buf,
                  E->progress, E->env_term, E->seg_ends, E->seg_count, E->denom_dev};
  for (void* p : ptrs) dfree(p);
  for (int r = 0; r < B2_MAX_PEERS; r++)
    if (r != E->rank && E->peer_base[r]) cudaIpcCloseMemHandle(E->peer_base[r]);
  cudaFree(E->xchg);
  if (E->status_host) cudaFreeHost(E->status_host);
  if (E->copy_stream) {
    cudaStreamSynchronize(E->copy_stream);
    for (auto ev : E->chunk_ev) if (ev) cudaEventDestroy(ev);
    if (E->copy_done) cudaEventDestroy(E->copy_done);
    cudaStreamDestroy(E->copy_stream);
  }
  if (E->ev0) cudaEventDestroy(E->ev0);
  if (E->ev1) cudaEventDestroy(E->ev1);
  if (E->stream) cudaStreamDestroy(E->stream);
  delete E;
  return B2_OK;
}

int b2_get_config(b2_handle E, b2_config* out) {
  NEED(E);
  if (!out) return set_error(B2_ERR_INVALID, "null out");
  *out = E->cfg;
  return B2_OK;
}

int64_t b2_n_params(b2_handle E) { return E ? E->P_ext : 0; }

int b2_init_params(b2_handle E, uint64_t seed) {
  NEED(E);
  // torch::nn::Linear::reset_parameters: kaiming_uniform(a=sqrt(5)) on W and U(+-1/sqrt(fan_in)) on b,
  // both = U(-1/sqrt(fan_in), 1/sqrt(fan_in)).  Drawn host-side from Philox (init is not a hot path;
  // parity runs import the reference's weights through b2_set_params instead).
  const NetDims N = NetDims::make(E->S_ext, E->dims.H1, E->dims.H2);
  std::vector<float> p((size_t)E->P_ext);
  uint64_t ctr = 0;
  auto fill = [&](int off, int n, int fan_in) {
    const float bound = 1.0f / sqrtf((float)fan_in);
    for (int i = 0; i < n; i++, ctr++) {
      uint32_t x[4];
      philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), 0x1717u, 0, (uint32_t)seed, (uint32_t)(seed >> 32), x);
      float u = ((float)(x[0] >> 8) + 0.5f) * (1.0f / 16777216.0f);
      p[(size_t)off + i] = (2.0f * u - 1.0f) * bound;
    }
  };
  for (int net = 0; net < 2; net++) {
    int o = net * N.SIZE;
    fill(o + N.oW1, N.H1 * N.S, N.S); fill(o + N.ob1, N.H1, N.S);
    fill(o + N.oW2, N.H2 * N.H1, N.H1); fill(o + N.ob2, N.H2, N.H1);
    fill(o + N.oW3, N.H2, N.H2); fill(o + N.ob3, 1, N.H2);
  }
  return b2_set_params(E, p.data(), E->P_ext);
}

int b2_set_params(b2_handle E, const float* p, int64_t n) {
  NEED(E);
  if (!p || n != E->P_ext) return set_error(B2_ERR_INVALID, "set_params: expected %d floats, got %lld", E->P_ext, (long long)n);
  return put_vec(E, E->params, p);
}
int b2_get_params(b2_handle E, float* p, int64_t n) {
  NEED(E);
  if (!p || n != E->P_ext) return set_error(B2_ERR_INVALID, "get_params: expected %d floats", E->P_ext);
  return get_vec(E, p, E->params);
}
int b2_set_adam_state(b2_handle E, const float* m, const float* v, int64_t n, int64_t step) {
  NEED(E);
  if (!m || !v || n != E->P_ext || step < 0) return set_error(B2_ERR_INVALID, "set_adam_state: bad arguments");
  int rc = put_vec(E, E->adam_m, m);
  if (rc || (rc = put_vec(E, E->adam_v, v))) return rc;
  CK(cudaMemcpyAsync(E->adam_step, &step, sizeof(int64_t), cudaMemcpyHostToDevice, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}
int b2_get_adam_state(b2_handle E, float* m, float* v, int64_t n, int64_t* step) {
  NEED(E);
  if (n != E->P_ext) return set_error(B2_ERR_INVALID, "get_adam_state: expected %d floats", E->P_ext);
  int rc;
  if (m && (rc = get_vec(E, m, E->adam_m))) return rc;
  if (v && (rc = get_vec(E, v, E->adam_v))) return rc;
  if (step) CK(cudaMemcpyAsync(step, E->adam_step, sizeof(int64_t), cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}
int b2_get_grads(b2_handle E, float* g, int64_t n) {
  NEED(E);
  if (!g || n != E->P_ext) return set_error(B2_ERR_INVALID, "get_grads: expected %d floats", E->P_ext);
  return get_vec(E, g, E->grads);
}

// checkpoint: "B2CK" | version | P | step | [v2: global_step, act_calls, reset_calls, n_envs, then env state u[n][4],
// ep_step[n], ep_ret[n]] | params | m | v   (the reference has none).  Version 2 also carries the
// Philox counters and the environment state, so that save -> load -> rollout continues the noise stream and the
// episodes of the run that saved, instead of replaying its first iterations; version-1 files still load (weights and
// Adam state only).  The env block is restored only into an engine with the same n_envs.
int b2_save_checkpoint(b2_handle E, const char* path) {
  NEED(E);
  if (!path) return set_error(B2_ERR_INVALID, "null path");
  std::vector<float> p(E->P_ext), m(E->P_ext), v(E->P_ext);
  int64_t step = 0;
  int rc = b2_get_params(E, p.data(), E->P_ext);
  if (rc) return rc;
  if ((rc = b2_get_adam_state(E, m.data(), v.data(), E->P_ext, &step))) return rc;
  FILE* f = fopen(path, "wb");
  if (!f) return set_error(B2_ERR_IO, "cannot open %s for writing", path);
  int32_t hdr[2] = {2, E->P_ext};
  const int64_t NE = E->cfg.n_envs;
  std::vector<double> u((size_t)4 * NE);
  std::vector<int32_t> es((size_t)NE);
  std::vector<float> er((size_t)NE);
  if ((rc = b2_env_get_state(E, u.data(), es.data(), NE))) { fclose(f); return rc; }
  if (cudaMemcpy(er.data(), E->ep_ret, sizeof(float) * NE, cudaMemcpyDeviceToHost) != cudaSuccess) { fclose(f); return set_error(B2_ERR_CUDA, "checkpoint: ep_ret D2H"); }
  const uint64_t ctr[4] = {E->global_step, E->act_calls, E->reset_calls, (uint64_t)NE};
  bool ok = fwrite("B2CK", 1, 4, f) == 4 && fwrite(hdr, sizeof hdr, 1, f) == 1 && fwrite(&step, sizeof step, 1, f) == 1 &&
            fwrite(ctr, sizeof ctr, 1, f) == 1 && fwrite(u.data(), sizeof(double), u.size(), f) == u.size() &&
            fwrite(es.data(), sizeof(int32_t), es.size(), f) == es.size() && fwrite(er.data(), sizeof(float), er.size(), f) == er.size() &&
            fwrite(p.data(), sizeof(float), E->P_ext, f) == (size_t)E->P_ext && fwrite(m.data(), sizeof(float), E->P_ext, f) == (size_t)E->P_ext &&
            fwrite(v.data(), sizeof(float), E->P_ext, f) == (size_t)E->P_ext;
  ok = (fclose(f) == 0) && ok;
  return ok ? B2_OK : set_error(B2_ERR_IO, "short write to %s", path);
}
int b2_load_checkpoint(b2_handle E, const char* path) {
  NEED(E);
  if (!path) return set_error(B2_ERR_INVALID, "null path");
  FILE* f = fopen(path, "rb");
  if (!f) return set_error(B2_ERR_IO, "cannot open %s", path);
  char magic[4];
  int32_t hdr[2];
  int64_t step = 0;
  std::vector<float> p(E->P_ext), m(E->P_ext), v(E->P_ext);
  bool ok = fread(magic, 1, 4, f) == 4 && memcmp(magic, "B2CK", 4) == 0 && fread(hdr, sizeof hdr, 1, f) == 1 &&
            (hdr[0] == 1 || hdr[0] == 2) && hdr[1] == E->P_ext && fread(&step, sizeof step, 1, f) == 1;
  uint64_t ctr[4] = {0, 0, 0, 0};
  std::vector<double> u;
  std::vector<int32_t> es;
  std::vector<float> er;
  if (ok && hdr[0] == 2) {
    ok = fread(ctr, sizeof ctr, 1, f) == 1 && ctr[3] >= 1 && ctr[3] <= (1ull << 40);
    if (ok) {
      u.resize((size_t)4 * ctr[3]); es.resize((size_t)ctr[3]); er.resize((size_t)ctr[3]);
      ok = fread(u.data(), sizeof(double), u.size(), f) == u.size() && fread(es.data(), sizeof(int32_t), es.size(), f) == es.size() &&
           fread(er.data(), sizeof(float), er.size(), f) == er.size();
    }
  }
  ok = ok && fread(p.data(), sizeof(float), E->P_ext, f) == (size_t)E->P_ext && fread(m.data(), sizeof(float), E->P_ext, f) == (size_t)E->P_ext &&
            fread(v.data(), sizeof(float), E->P_ext, f) == (size_t)E->P_ext;
  fclose(f);
  if (!ok) return set_error(B2_ERR_IO, "%s is not a matching b200ppo checkpoint", path);
  int rc = b2_set_params(E, p.data(), E->P_ext);
  if (rc) return rc;
  if ((rc = b2_set_adam_state(E, m.data(), v.data(), E->P_ext, step))) return rc;
  if (hdr[0] == 2) {
    E->global_step = ctr[0]; E->act_calls = ctr[1]; E->reset_calls = ctr[2];
    if ((int64_t)ctr[3] == E->cfg.n_envs) {
      if ((rc = b2_env_set_state(E, u.data(), es.data(), E->cfg.n_envs))) return rc;
      CK(cudaMemcpy(E->ep_ret, er.data(), sizeof(float) * er.size(), cudaMemcpyHostToDevice));
      CK(cudaDeviceSynchronize());   // (pageable H2D on the legacy stream may still be in flight; E->stream is non-blocking)
    }
  }
  return B2_OK;
}

int b2_act(b2_handle E, const float* obs, int64_t n, const float* eps, float* action, float* logp, float* mu) {
  NEED(E);
  if (n == 0) return B2_OK;
  if (!obs || !action || !logp || n < 0) return set_error(B2_ERR_INVALID, "b2_act: null/negative argument");
  CK(cudaSetDevice(E->cfg.device));
  // staging: obs[n][6] | eps[n] | action[n] | logp[n] | mu[n]
  int rc = ensure_stage(E, n * (OBS + 4), 0);
  if (rc) return rc;
  float* d_obs = E->stage;
  float* d_eps = d_obs + n * OBS;
  float* d_act = d_eps + n;
  float* d_lp = d_act + n;
  float* d_mu = d_lp + n;
  if ((rc = h2d_obs(E, d_obs, obs, n, E->stream))) return rc;
  if (eps) CK(cudaMemcpyAsync(d_eps, eps, sizeof(float) * n, cudaMemcpyHostToDevice, E->stream));
  cudaError_t ce;
  if (E->wide) {
    ce = launch_wide_forward(E->dims, E->params, d_obs, 1, OBS, nullptr, n, 1, d_mu, E->stream);
    if (ce == cudaSuccess)
      ce = launch_wide_act_finish(d_mu, eps ? d_eps : nullptr, n, E->cfg.seed, E->act_calls++, d_act, d_lp, E->hp, E->status, E->stream);
    E->launches += 1;
  } else
    ce = launch_act(E->params, d_obs, eps ? d_eps : nullptr, n, E->cfg.seed, E->act_calls++, d_act, d_lp,
                    d_mu, E->hp, E->status, E->stream);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "act: %s", cudaGetErrorString(ce));
  E->launches += 1;
  CK(cudaMemcpyAsync(action, d_act, sizeof(float) * n, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaMemcpyAsync(logp, d_lp, sizeof(float) * n, cudaMemcpyDeviceToHost, E->stream));
  if (mu) CK(cudaMemcpyAsync(mu, d_mu, sizeof(float) * n, cudaMemcpyDeviceToHost, E->stream));
  return check_status(E, "b2_act");
}

int b2_value(b2_handle E, const float* obs, int64_t n, float* value) {
  NEED(E);
  if (n == 0) return B2_OK;
  if (!obs || !value || n < 0) return set_error(B2_ERR_INVALID, "b2_value: null/negative argument");
  CK(cudaSetDevice(E->cfg.device));
  int rc = ensure_stage(E, n * (OBS + 1), 0);
  if (rc) return rc;
  float* d_obs = E->stage;
  float* d_v = d_obs + n * OBS;
  if ((rc = h2d_obs(E, d_obs, obs, n, E->stream))) return rc;
  cudaError_t ce = E->wide ? launch_wide_forward(E->dims, E->params + E->dims.SIZE, d_obs, 1, OBS, nullptr, n, 0, d_v, E->stream)
                           : launch_value(E->params, d_obs, n, d_v, E->stream);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "value: %s", cudaGetErrorString(ce));
  E->launches += 1;
  CK(cudaMemcpyAsync(value, d_v, sizeof(float) * n, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}

int b2_update(b2_handle E, const b2_batch* batch, b2_update_stats* stats) {
  NEED(E);
  CK(cudaSetDevice(E->cfg.device));
  if (!batch || !batch->action || !batch->logp) return set_error(B2_ERR_INVALID, "b2_update: batch needs obs, action, logp, reward, flags");
  int rc = upload_batch(E, batch);
  if (rc) return rc;
  return run_update(E, stats);
}

int b2_returns(b2_handle E, const b2_batch* batch, float* returns) {
  NEED(E);
  CK(cudaSetDevice(E->cfg.device));
  if (!returns) return set_error(B2_ERR_INVALID, "b2_returns: null output");
  int rc = upload_batch(E, batch);
  if (rc) return rc;
  rc = run_returns(E);
  if (rc) return rc;
  CK(cudaMemcpyAsync(returns, E->tr.ret, sizeof(float) * batch->n_envs * batch->horizon, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}

int b2_last_update_path(b2_handle E) { return E ? E->last_update_path : -1; }

int b2_last_update_losses(b2_handle E, double* l1, double* l2, double* l3, double* lt, int n) {
  NEED(E);
  const int steps = (int)(E->last_losses.size() / 4);
  if (n > steps) return set_error(B2_ERR_INVALID, "only %d optimizer steps recorded", steps);
  for (int i = 0; i < n; i++) {
    const double* L = &E->last_losses[(size_t)4 * i];
    if (l1) l1[i] = L[0];
    if (l2) l2[i] = L[1];
    if (l3) l3[i] = (double)E->hp.loss3;
    if (lt) lt[i] = L[0] + L[1] + (double)E->hp.loss3;
  }
  return B2_OK;
}

int b2_env_reset(b2_handle E) {
  NEED(E);
  CK(cudaSetDevice(E->cfg.device));
  const int64_t NE = E->cfg.n_envs;
  CK(cudaMemsetAsync(E->state, 0, sizeof(double) * 4 * NE, E->stream));
  if (E->cfg.env_flags & B2_ENV_RANDOM_INIT) {   // a fresh draw per reset call; ticks >= 2^63 never collide with in-rollout resets
    E->reset_calls += 1;
    cudaError_t ce = launch_env_init(E->state, NE, E->cfg.env_flags, E->cfg.seed, E->cfg.env_offset,
                                     (1ull << 63) | E->reset_calls, E->stream);
    if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "env init: %s", cudaGetErrorString(ce));
    E->launches += 1;
  }
  CK(cudaMemsetAsync(E->env_term, 0, (size_t)NE, E->stream));
  CK(cudaMemsetAsync(E->ep_step, 0, sizeof(int32_t) * NE, E->stream));
  CK(cudaMemsetAsync(E->ep_ret, 0, sizeof(float) * NE, E->stream));
  CK(cudaMemsetAsync(E->ep_count, 0, sizeof(int32_t) * NE, E->stream));
  E->global_step = 0;
  return B2_OK;
}

int b2_env_set_state(b2_handle E, const double* u, const int32_t* ep_step, int64_t n) {
  NEED(E);
  if (!u || n != E->cfg.n_envs) return set_error(B2_ERR_INVALID, "env_set_state: expected %lld envs", (long long)E->cfg.n_envs);
  std::vector<double> soa((size_t)4 * n);
  for (int64_t e = 0; e < n; e++)
    for (int j = 0; j < 4; j++) soa[(size_t)j * n + e] = u[4 * e + j];
  CK(cudaMemcpyAsync(E->state, soa.data(), sizeof(double) * 4 * n, cudaMemcpyHostToDevice, E->stream));
  if (ep_step) CK(cudaMemcpyAsync(E->ep_step, ep_step, sizeof(int32_t) * n, cudaMemcpyHostToDevice, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}
int b2_env_get_state(b2_handle E, double* u, int32_t* ep_step, int64_t n) {
  NEED(E);
  if (n != E->cfg.n_envs) return set_error(B2_ERR_INVALID, "env_get_state: expected %lld envs", (long long)E->cfg.n_envs);
  std::vector<double> soa((size_t)4 * n);
  CK(cudaMemcpyAsync(soa.data(), E->state, sizeof(double) * 4 * n, cudaMemcpyDeviceToHost, E->stream));
  if (ep_step) CK(cudaMemcpyAsync(ep_step, E->ep_step, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  if (u)
    for (int64_t e = 0; e < n; e++)
      for (int j = 0; j < 4; j++) u[4 * e + j] = soa[(size_t)j * n + e];
  return B2_OK;
}

int b2_env_step(b2_handle E, const double* force, int64_t n) {
  NEED(E);
  if (!force || n != E->cfg.n_envs) return set_error(B2_ERR_INVALID, "env_step: expected %lld forces", (long long)E->cfg.n_envs);
  CK(cudaSetDevice(E->cfg.device));
  // staging (as doubles): force[n] | obs[n][6] | reward[n]
  int rc = ensure_stage(E, 2 * 8 * n, 0);
  if (rc) return rc;
  double* d_force = reinterpret_cast<double*>(E->stage);
  CK(cudaMemcpyAsync(d_force, force, sizeof(double) * n, cudaMemcpyHostToDevice, E->stream));
  cudaError_t ce = launch_env_step(E->state, E->ep_step, d_force, n, E->cfg.precision, E->cfg.env_flags, E->env_term, E->stream);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "env_step: %s", cudaGetErrorString(ce));
  E->launches += 1;
  E->total_env_steps += n;
  CK(cudaStreamSynchronize(E->stream));  // the caller may reuse `force` right away
  return B2_OK;
}

int b2_env_observe(b2_handle E, double* obs, double* reward, int64_t n) {
  NEED(E);
  if (n != E->cfg.n_envs) return set_error(B2_ERR_INVALID, "env_observe: expected %lld envs", (long long)E->cfg.n_envs);
  if (!obs && !reward) return B2_OK;
  CK(cudaSetDevice(E->cfg.device));
  int rc = ensure_stage(E, 2 * 8 * n, 0);
  if (rc) return rc;
  double* d_obs = reinterpret_cast<double*>(E->stage) + n;
  double* d_rew = d_obs + 6 * n;
  cudaError_t ce = launch_env_observe(E->state, obs ? d_obs : nullptr, reward ? d_rew : nullptr, n, E->cfg.env_flags, E->stream);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "env_observe: %s", cudaGetErrorString(ce));
  E->launches += 1;
  if (obs) CK(cudaMemcpyAsync(obs, d_obs, sizeof(double) * 6 * n, cudaMemcpyDeviceToHost, E->stream));
  if (reward) CK(cudaMemcpyAsync(reward, d_rew, sizeof(double) * n, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}

int b2_env_terminated(b2_handle E, uint8_t* terminated, int64_t n) {
  NEED(E);
  if (!terminated || n != E->cfg.n_envs) return set_error(B2_ERR_INVALID, "env_terminated: expected %lld envs", (long long)E->cfg.n_envs);
  CK(cudaSetDevice(E->cfg.device));
  CK(cudaMemcpyAsync(terminated, E->env_term, (size_t)n, cudaMemcpyDeviceToHost, E->stream));
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}

int b2_set_state_trace(b2_handle E, int enable) {
  NEED(E);
  E->trace = enable != 0;
  return ensure_traj(E, E->tr.cap);
}

// one k_rollout launch covering steps [t_begin, t_end) of a T-step rollout that started at E->global_step
static int rollout_window(b2_engine* E, int T, int t_begin, int t_end, bool forced, float* obs_aos = nullptr,
                          const b2_batch_out* host = nullptr, float* obs4_aos = nullptr, uint8_t* rflags = nullptr) {
  RolloutArgs a{};
  a.params = E->params; a.state = E->state; a.ep_step = E->ep_step; a.ep_ret = E->ep_ret;
  a.ep_count = E->ep_count; a.ep_log = E->ep_log; a.max_eps = E->max_eps;
  a.tr = E->tr;
  a.forced_action = forced ? E->forced : nullptr;
  a.obs_aos = obs_aos;
  a.obs4_aos = obs4_aos; a.rflags = rflags;
  if (host) { a.h_action = host->action; a.h_logp = host->logp; a.h_mu = host->mu; a.h_reward = host->reward; a.h_flags = host->flags; }
  a.n_envs = E->cfg.n_envs; a.T = T; a.t_begin = t_begin; a.t_end = t_end; a.n_timestep = E->cfg.n_timestep;
  a.seed = E->cfg.seed; a.env_offset = E->cfg.env_offset; a.step0 = E->global_step;
  a.hp = E->hp; a.status = E->status; a.progress = E->progress; a.env_flags = E->cfg.env_flags;
  cudaError_t ce;
  if (E->wide) {   // wide actors: one env step = observe + actor forward + step (rollout_wide.cu)
    ce = launch_rollout_wide(a, E->dims, E->cfg.precision, E->trace, E->stream, &E->launches);
  } else {
    ce = launch_rollout(a, E->cfg.precision, E->trace, E->sm_count, E->stream, E->cfg.hidden1, E->cfg.hidden2);
    E->launches += 1;
  }
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "rollout: %s", cudaGetErrorString(ce));
  if (E->S_ext != OBS) {   // getobs(4): the stored cos/sin planes must be exact zeros (see S_ext above)
    const int64_t k0 = (int64_t)t_begin * E->cfg.n_envs, n = (int64_t)(t_end - t_begin) * E->cfg.n_envs;
    for (int c = E->S_ext; c < OBS; c++) CK(cudaMemsetAsync(E->tr.obs + (size_t)c * E->tr.cap + k0, 0, sizeof(float) * n, E->stream));
  }
  return B2_OK;
}

static int rollout_prepare(b2_engine* E, int T, const float* forced_action) {
  CK(cudaSetDevice(E->cfg.device));
  if (T < 1 || T > E->cfg.horizon) return set_error(B2_ERR_INVALID, "rollout: T=%d outside [1, horizon=%d]", T, E->cfg.horizon);
  const int64_t NE = E->cfg.n_envs;
  int rc = ensure_traj(E, NE * T);
  if (rc) return rc;
  if (forced_action) {
    if (NE * T > E->forced_cap) {
      dfree(E->forced);
      E->forced = nullptr;
      CK(dmalloc(&E->forced, sizeof(float) * NE * T));
      E->forced_cap = NE * T;
    }
    CK(cudaMemcpyAsync(E->forced, forced_action, sizeof(float) * NE * T, cudaMemcpyHostToDevice, E->stream));
  }
  E->tr.n_envs = NE; E->tr.horizon = T;
  return B2_OK;
}

static void rollout_finish(b2_engine* E, int T) {
  E->global_step += (uint64_t)T;
  E->total_env_steps += E->cfg.n_envs * T;
  E->have_traj = true;
}

int b2_rollout(b2_handle E, int T, const float* forced_action) {
  NEED(E);
  int rc = rollout_prepare(E, T, forced_action);
  if (rc) return rc;
  if ((rc = rollout_window(E, T, 0, T, forced_action != nullptr))) return rc;
  rollout_finish(E, T);
  return B2_OK;
}

// chunk boundaries of the host pipeline (see b2_rollout_host): doubling head (2, 2, 4, .. steps), equal chunks, halving tail
// (copy_bound = false: the copies take less time than the simulation -- the compact record -- so no doubling head: short
// windows cost the persistent kernel its balance, 2 steps x 2,048 groups over 1,776 warps quantise to 3 units for 2.3)
static int host_chunk_schedule(int T, int n_chunks, int* bounds, bool copy_bound = true) {
  int nb = 0;
  const int cap = (T + n_chunks - 1) / n_chunks;
  int t = 0;
  bounds[0] = 0;
  for (int len = 2; copy_bound && len < cap && T - t > 2 * cap && nb < 6; len *= 2) { t += len; bounds[++nb] = t; if (nb == 1) { t += len; bounds[++nb] = t; } }
  while (T - t > cap && nb < B2_MAX_CHUNKS - 8) { t += cap; bounds[++nb] = t; }
  int r = T - t;                                   // the last <= cap steps (more only if out of events): halving tail
  while (r > 2 && nb < B2_MAX_CHUNKS - 1) { const int len = r / 2; t += len; r -= len; bounds[++nb] = t; }
  bounds[++nb] = T;
  return nb;
}

// Rollout with the trajectory delivered into HOST buffers.  The horizon is cut into n_chunks launches; while
// chunk i+1 is being simulated, chunk i (a contiguous k-range of every field, the buffer being time-major) is
// transposed to MemoryNN's AoS observation layout and copied out on a second stream, so the PCIe transfer
// (37 B per env-step) hides behind the simulation instead of following it.
int b2_rollout_host(b2_handle E, int T, int n_chunks, const b2_batch_out* out) {
  NEED(E);
  if (!out) return set_error(B2_ERR_INVALID, "rollout_host: null out");
  int rc = rollout_prepare(E, T, nullptr);
  if (rc) return rc;
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > T) n_chunks = T;
  if (n_chunks > B2_MAX_CHUNKS) n_chunks = B2_MAX_CHUNKS;
  const int64_t NE = E->cfg.n_envs;
  Traj& t = E->tr;
  // Two ways to get the trajectory into host buffers.  (a) DIRECT (B2_ROLLOUT_HOST_DIRECT=1, pinned destinations, 6-wide
  // observations): ONE launch whose stores go to the device trajectory AND straight into the host buffers (posted PCIe
  // writes, the observations transposed in the warp so that every store is a whole 128-B line).  SM-issued writes
  // sustain ~44 GB/s, so 37 B per record cost 7.05 ms at the reference shape whatever the kernel does.  (b) The chunked
  // copy-engine PIPELINE below (default): the DMA engine moves 57 GB/s (5.4 ms for the same bytes) behind the simulation.
  // Round 1 (6.3 ms of simulation): direct 7.1-7.2 ms, pipeline 7.3-7.4 ms -> direct was the default.  Round 2 (5.6 ms of
  // simulation, the 27-instruction stage): direct 7.45 ms = 1.13e9 env-steps/s, pipeline 6.7 ms = 1.25e9 -> the pipeline
  // is (profiles/r2/r2d_e2e_ab.txt).  Splitting the fields between the two (small fields direct, observations by DMA)
  // was 9.3 ms: SM writes and DMA contend for the link.
  {
    auto pinned = [](const void* p) {
      if (!p) return true;
      cudaPointerAttributes at;
      if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
      return at.type == cudaMemoryTypeHost;
    };
    const char* env = getenv("B2_ROLLOUT_HOST_DIRECT");
    const bool direct_ok = (env && env[0] == '1') && !E->wide && E->S_ext == OBS && pinned(out->obs) && pinned(out->action) &&
                           pinned(out->logp) && pinned(out->mu) && pinned(out->reward) && pinned(out->flags);
    if (direct_ok) {
      if ((rc = rollout_window(E, T, 0, T, false, out->obs, out))) return rc;
      rollout_finish(E, T);
      return check_status(E, "rollout");    // synchronises the stream: the host buffers are complete
    }
  }
  if (out->obs && (rc = ensure_stage(E, NE * T * OBS, 0))) return rc;
  if (!E->copy_stream) {
    CK(cudaStreamCreateWithFlags(&E->copy_stream, cudaStreamNonBlocking));
    for (auto& ev : E->chunk_ev) CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&E->copy_done, cudaEventDisableTiming));
  }
  // Chunk schedule.  Measured on the B200 box: pinned D2H runs at 57 GB/s, so the 37 B per record leave in 5.4 ms while
  // simulating them takes 6.4 ms at the reference shape -- the copies hide behind the simulation and the run costs
  // (all simulation) + (the LAST chunk's copy).  Hence equal chunks of T / n_chunks steps (long launches keep the
  // persistent kernel's launch quantisation small) and a halving tail (.., cap/2, cap/4, .., 2, 2 steps) so that only
  // ~2 steps' worth of bytes is still in flight when the simulation ends.
  // The copy engine is nearly as busy as the SMs (5.4 vs 5.6 ms at the reference shape), so the FIRST chunk matters as
  // much: nothing can leave before it is simulated.  Doubling head (2, 2, 4, .. steps), equal chunks, halving tail.
  int bounds[B2_MAX_CHUNKS + 1];
  const int nb = host_chunk_schedule(T, n_chunks, bounds);
  for (int c = 0; c < nb; c++) {
    const int t0 = bounds[c], t1 = bounds[c + 1];
    // 6-wide observations leave the kernel record-major as well (RolloutArgs::obs_aos): no transpose between the
    // simulation and the copy.  4-wide ones (S_ext = 4) keep the transpose: their stored cos/sin planes are zeroed first.
    const bool direct = out->obs && E->S_ext == OBS && !E->wide;
    if ((rc = rollout_window(E, T, t0, t1, false, direct ? E->stage : nullptr))) return rc;
    const int64_t k0 = (int64_t)t0 * NE, n = (int64_t)(t1 - t0) * NE;
    if (out->obs && !direct) {
      // The [c][k] -> [k][c] transpose runs on the MAIN stream, between two rollout launches: the persistent rollout
      // kernel owns every register file while it runs, so a transpose queued on the copy stream could only start
      // after the NEXT chunk's simulation had finished, delaying two thirds of the chunk's bytes by one chunk.
      cudaError_t ce = launch_soa_to_aos(t.obs + k0, E->stage + k0 * OBS, t.cap, n, E->stream);
      if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "soa_to_aos: %s", cudaGetErrorString(ce));
      E->launches += 1;
    }
    CK(cudaEventRecord(E->chunk_ev[c], E->stream));
    CK(cudaStreamWaitEvent(E->copy_stream, E->chunk_ev[c], 0));
    if (out->obs && (rc = d2h_obs(E, out->obs + k0 * E->S_ext, E->stage + k0 * OBS, n, E->copy_stream))) return rc;
    if (out->action) CK(cudaMemcpyAsync(out->action + k0, t.action + k0, sizeof(float) * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->logp) CK(cudaMemcpyAsync(out->logp + k0, t.logp + k0, sizeof(float) * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->mu) CK(cudaMemcpyAsync(out->mu + k0, t.mu + k0, sizeof(float) * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->reward) CK(cudaMemcpyAsync(out->reward + k0, t.reward + k0, sizeof(float) * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->flags) CK(cudaMemcpyAsync(out->flags + k0, t.flags + k0, (size_t)n, cudaMemcpyDeviceToHost, E->copy_stream));
  }
  rollout_finish(E, T);
  // the main stream must not overwrite the buffer (next rollout) before the copies have drained
  CK(cudaEventRecord(E->copy_done, E->copy_stream));
  CK(cudaStreamWaitEvent(E->stream, E->copy_done, 0));
  CK(cudaStreamSynchronize(E->copy_stream));
  return check_status(E, "rollout");
}

// The chunked copy-engine pipeline of b2_rollout_host with the compact record (b200ppo.h b2_batch_compact): the kernel
// also leaves (x, xdot, thetadot, theta) as one 16-B record per env-step and flags | reward bit in device staging,
// and 25 instead of 37 B per env-step cross PCIe -- less time than the simulation takes, so the run is compute-bound.
int b2_rollout_host_compact(b2_handle E, int T, int n_chunks, const b2_batch_compact* out) {
  NEED(E);
  if (!out) return set_error(B2_ERR_INVALID, "rollout_host_compact: null out");
  if (E->wide || E->cfg.env_flags || E->S_ext != OBS)
    return set_error(B2_ERR_INVALID, "rollout_host_compact: default nets, the shipped environment and obs_vars = 6 only");
  int rc = rollout_prepare(E, T, nullptr);
  if (rc) return rc;
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > T) n_chunks = T;
  if (n_chunks > B2_MAX_CHUNKS) n_chunks = B2_MAX_CHUNKS;
  const int64_t NE = E->cfg.n_envs;
  Traj& t = E->tr;
  if ((rc = ensure_stage(E, NE * T * 4, NE * T))) return rc;
  if (!E->copy_stream) {
    CK(cudaStreamCreateWithFlags(&E->copy_stream, cudaStreamNonBlocking));
    for (auto& ev : E->chunk_ev) CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&E->copy_done, cudaEventDisableTiming));
  }
  int bounds[B2_MAX_CHUNKS + 1];
  const int nb = host_chunk_schedule(T, n_chunks, bounds, false);
  for (int c = 0; c < nb; c++) {
    const int t0 = bounds[c], t1 = bounds[c + 1];
    if ((rc = rollout_window(E, T, t0, t1, false, nullptr, nullptr, E->stage, E->stage_u8))) return rc;
    const int64_t k0 = (int64_t)t0 * NE, n = (int64_t)(t1 - t0) * NE;
    CK(cudaEventRecord(E->chunk_ev[c], E->stream));
    CK(cudaStreamWaitEvent(E->copy_stream, E->chunk_ev[c], 0));
    if (out->obs4) CK(cudaMemcpyAsync(out->obs4 + k0 * 4, E->stage + k0 * 4, sizeof(float) * 4 * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->action) CK(cudaMemcpyAsync(out->action + k0, t.action + k0, sizeof(float) * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->logp) CK(cudaMemcpyAsync(out->logp + k0, t.logp + k0, sizeof(float) * n, cudaMemcpyDeviceToHost, E->copy_stream));
    if (out->rflags) CK(cudaMemcpyAsync(out->rflags + k0, E->stage_u8 + k0, (size_t)n, cudaMemcpyDeviceToHost, E->copy_stream));
  }
  rollout_finish(E, T);
  CK(cudaEventRecord(E->copy_done, E->copy_stream));
  CK(cudaStreamWaitEvent(E->stream, E->copy_done, 0));
  CK(cudaStreamSynchronize(E->copy_stream));
  return check_status(E, "rollout");
}

int b2_update_resident(b2_handle E, b2_update_stats* stats) {
  NEED(E);
  CK(cudaSetDevice(E->cfg.device));
  if (!E->have_traj) return set_error(B2_ERR_STATE, "update_resident: no trajectory on device (call b2_rollout first)");
  int rc = check_status(E, "rollout");
  if (rc) return rc;
  return run_update(E, stats);
}

int b2_train_iteration(b2_handle E, int T, b2_update_stats* stats) {
  int rc = b2_rollout(E, T, nullptr);
  if (rc) return rc;
  return b2_update_resident(E, stats);
}

int b2_traj_download(b2_handle E, b2_batch_out* out, double* state_trace) {
  NEED(E);
  CK(cudaSetDevice(E->cfg.device));
  if (!E->have_traj) return set_error(B2_ERR_STATE, "traj_download: nothing on device");
  if (!out) return set_error(B2_ERR_INVALID, "null out");
  Traj& t = E->tr;
  const int64_t B = t.n_envs * t.horizon;
  if (out->obs) {
    int rc = ensure_stage(E, B * OBS, 0);
    if (rc) return rc;
    cudaError_t ce = launch_soa_to_aos(t.obs, E->stage, t.cap, B, E->stream);
    if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "soa_to_aos: %s", cudaGetErrorString(ce));
    E->launches += 1;
    if ((rc = d2h_obs(E, out->obs, E->stage, B, E->stream))) return rc;
  }
  if (out->action) CK(cudaMemcpyAsync(out->action, t.action, sizeof(float) * B, cudaMemcpyDeviceToHost, E->stream));
  if (out->logp) CK(cudaMemcpyAsync(out->logp, t.logp, sizeof(float) * B, cudaMemcpyDeviceToHost, E->stream));
  if (out->mu) CK(cudaMemcpyAsync(out->mu, t.mu, sizeof(float) * B, cudaMemcpyDeviceToHost, E->stream));
  if (out->reward) CK(cudaMemcpyAsync(out->reward, t.reward, sizeof(float) * B, cudaMemcpyDeviceToHost, E->stream));
  if (out->flags) CK(cudaMemcpyAsync(out->flags, t.flags, B, cudaMemcpyDeviceToHost, E->stream));
  if (out->returns) CK(cudaMemcpyAsync(out->returns, t.ret, sizeof(float) * B, cudaMemcpyDeviceToHost, E->stream));
  if (state_trace) {
    if (!E->trace || !t.state_trace) return set_error(B2_ERR_STATE, "state trace was not enabled before the rollout");
    CK(cudaMemcpyAsync(state_trace, t.state_trace, sizeof(double) * 4 * B, cudaMemcpyDeviceToHost, E->stream));
  }
  return check_status(E, "rollout");
}

int b2_episode_stats(b2_handle E, int32_t* counts, float* returns, int max_eps, int64_t n) {
  NEED(E);
  if (n != E->cfg.n_envs || !counts) return set_error(B2_ERR_INVALID, "episode_stats: expected %lld envs", (long long)E->cfg.n_envs);
  CK(cudaMemcpyAsync(counts, E->ep_count, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, E->stream));
  std::vector<float> log;
  if (returns) {
    log.resize((size_t)n * E->max_eps);
    CK(cudaMemcpyAsync(log.data(), E->ep_log, sizeof(float) * n * E->max_eps, cudaMemcpyDeviceToHost, E->stream));
  }
  CK(cudaStreamSynchronize(E->stream));
  if (returns)
    for (int64_t e = 0; e < n; e++)
      for (int i = 0; i < max_eps; i++)
        returns[e * max_eps + i] = (i < counts[e] && i < E->max_eps) ? log[(size_t)i * n + e] : 0.0f;
  return B2_OK;
}

int64_t b2_total_env_steps(b2_handle E) { return E ? E->total_env_steps : 0; }

int b2_comm_unique_id(char id[B2_COMM_ID_BYTES]) {
  if (!id) return set_error(B2_ERR_INVALID, "null id");
  if (!load_nccl()) return set_error(B2_ERR_COMM, "libnccl.so.2 not found: %s", dlerror());
  ncclUniqueId_t u;
  int e = g_nccl.GetUniqueId(&u);
  if (e != 0) return set_error(B2_ERR_COMM, "ncclGetUniqueId failed (%d)", e);
  static_assert(sizeof(u) == B2_COMM_ID_BYTES, "id size");
  memcpy(id, &u, sizeof u);
  return B2_OK;
}

int b2_comm_init(b2_handle E, int rank, int world, const char id[B2_COMM_ID_BYTES]) {
  NEED(E);
  if (world < 1 || rank < 0 || rank >= world || !id) return set_error(B2_ERR_INVALID, "comm_init: bad rank/world");
  if (world == 1) { E->rank = 0; E->world = 1; return B2_OK; }
  if (!load_nccl()) return set_error(B2_ERR_COMM, "libnccl.so.2 not found");
  CK(cudaSetDevice(E->cfg.device));
  ncclUniqueId_t u;
  memcpy(&u, id, sizeof u);
  int e = g_nccl.CommInitRank(&E->comm, world, u, rank);
  if (e != 0) return set_error(B2_ERR_COMM, "ncclCommInitRank: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "?");
  E->rank = rank;
  E->world = world;
  return setup_peer_exchange(E);
}
int b2_comm_world(b2_handle E) { return E ? E->world : 0; }
int b2_comm_mode(b2_handle E) { return !E || E->world <= 1 ? 0 : (E->p2p ? 2 : 1); }

int b2_sync(b2_handle E) {
  NEED(E);
  CK(cudaStreamSynchronize(E->stream));
  return B2_OK;
}
int b2_timer_start(b2_handle E) {
  NEED(E);
  CK(cudaEventRecord(E->ev0, E->stream));
  return B2_OK;
}
int b2_timer_stop_ms(b2_handle E, float* ms) {
  NEED(E);
  CK(cudaEventRecord(E->ev1, E->stream));
  CK(cudaEventSynchronize(E->ev1));
  if (ms) CK(cudaEventElapsedTime(ms, E->ev0, E->ev1));
  return B2_OK;
}
int64_t b2_kernel_launches(b2_handle E) { return E ? E->launches : 0; }
void* b2_stream(b2_handle E) { return E ? (void*)E->stream : nullptr; }

int b2_flush_l2(b2_handle E) {
  NEED(E);
  if (!E->flush_buf) {
    E->flush_n = (int64_t)160 << 18;  // 160 MiB of floats: larger than the 126 MB L2
    CK(dmalloc(&E->flush_buf, sizeof(float) * E->flush_n));
  }
  cudaError_t ce = launch_flush(E->flush_buf, E->flush_n, E->stream);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "flush: %s", cudaGetErrorString(ce));
  return B2_OK;
}

int b2_measure_fma_peak(b2_handle E, int fp64, double* tflops) {
  NEED(E);
  if (!tflops) return set_error(B2_ERR_INVALID, "null out");
  CK(cudaSetDevice(E->cfg.device));
  cudaError_t ce = measure_fma_peak(fp64 != 0, E->sm_count, E->stream, tflops);
  if (ce != cudaSuccess) return set_error(B2_ERR_CUDA, "fma peak: %s", cudaGetErrorString(ce));
  return B2_OK;
}

int b2_host_alloc(void** p, size_t bytes) {
  if (!p) return set_error(B2_ERR_INVALID, "null out");
  CK(cudaMallocHost(p, bytes));
  return B2_OK;
}
int b2_selftest_gemm(b2_handle E, int M, int N, int K, const float* A, const float* B, float* C, int splits, int mn_major) {
  NEED(E);
  if (!A || !B || !C || M < 1 || N < 1 || K < 1) return set_error(B2_ERR_INVALID, "selftest_gemm: bad argument");
  CK(cudaSetDevice(E->cfg.device));
  float *dA = nullptr, *dB = nullptr, *dC = nullptr;
  CK(dmalloc(&dA, sizeof(float) * (size_t)M * K));
  CK(dmalloc(&dB, sizeof(float) * (size_t)N * K));
  CK(dmalloc(&dC, sizeof(float) * (size_t)M * N));
  CK(cudaMemcpyAsync(dA, A, sizeof(float) * (size_t)M * K, cudaMemcpyHostToDevice, E->stream));
  CK(cudaMemcpyAsync(dB, B, sizeof(float) * (size_t)N * K, cudaMemcpyHostToDevice, E->stream));
  int rc = wide_selftest_gemm(M, N, K, dA, dB, dC, splits, mn_major, E->sm_count, E->stream);
  E->launches += 4;
  if (!rc) {
    cudaError_t ce = cudaMemcpy(C, dC, sizeof(float) * (size_t)M * N, cudaMemcpyDeviceToHost);
    if (ce != cudaSuccess) rc = set_error(B2_ERR_CUDA, "selftest_gemm D2H: %s", cudaGetErrorString(ce));
  }
  dfree(dA); dfree(dB); dfree(dC);
  return rc;
}

// every guard zone of every live device allocation of this process (B2_GUARD=1): number of corrupted zones, 0 = clean,
// -1 = guard mode is off.  The message names the first offender.
int b2_debug_guard_check(void) {
  if (!guard_on()) return -1;
  std::lock_guard<std::mutex> lk(guard_book().mu);
  cudaDeviceSynchronize();
  std::vector<unsigned char> buf(GUARD_BYTES);
  int bad = 0;
  for (auto& kv : guard_book().live) {
    char* p = static_cast<char*>(kv.first);
    const size_t bytes = kv.second;
    for (int side = 0; side < 2; side++) {
      const char* zone = side == 0 ? p - GUARD_BYTES : p + bytes;
      const size_t n = GUARD_BYTES;   // after the payload: the alignment padding, then the guard zone (all 0xA5)
      if (cudaMemcpy(buf.data(), zone, n, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); continue; }
      for (size_t i = 0; i < n; i++)
        if (buf[i] != 0xA5) {
          if (!bad) set_error(B2_ERR_STATE, "guard zone %s a %zu-byte device allocation was overwritten (first bad byte at offset %zu)",
                              side == 0 ? "BEFORE" : "AFTER", bytes, i);
          bad++;
          break;
        }
    }
  }
  return bad;
}

int b2_host_free(void* p) {
  if (p) CK(cudaFreeHost(p));
  return B2_OK;
}

}  // extern "C"
