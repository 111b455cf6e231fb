/* include/b200ppo.h -- C ABI of libb200ppo.so, the B200-native PPO rollout-and-update engine.
 *
 * The reference (haotianh9/Reinforcement_learning_cpp_parallel) has no FFI: its boundary is the
 * C++ class API of include/ppo_nn.h plus the header-only env of include/cart_pole.h, both consumed
 * by src/main.cpp in one binary.  This header is what a binding for that path
 * would link against; each entry point cites the reference interface it replaces (paths relative
 * to the reference checkout).  include/compat/ppo_nn.h and include/compat/cart_pole.h are thin
 * C++ shims over this ABI that keep the reference's own names and call shapes.
 *
 * Conventions: plain pointers and sizes only; every function returns B2_OK (0) or a negative
 * b2_status and never exits or throws; b2_last_error() gives the message (thread-local).
 * Host pointers unless a name ends in _dev.  A handle owns its device memory and one CUDA
 * stream; calls on one handle are not re-entrant.  There is NO CPU fallback: without a CUDA
 * device b2_create fails with B2_ERR_NO_DEVICE.
 */
#ifndef B200PPO_H
#define B200PPO_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 2   /* 2: b2_config.gae_lambda takes the first 8 bytes of the former reserved[6] (same size) */

typedef enum {
  B2_OK = 0,
  B2_ERR_INVALID = -1,     /* bad argument / unsupported shape */
  B2_ERR_NO_DEVICE = -2,   /* no CUDA device: the engine has no CPU path */
  B2_ERR_CUDA = -3,        /* a CUDA runtime call failed */
  B2_ERR_NONFINITE = -4,   /* NaN/Inf action or loss detected on device (cf. src/main.cpp:50-55) */
  B2_ERR_IO = -5,          /* config / checkpoint file problem */
  B2_ERR_COMM = -6,        /* NCCL failure */
  B2_ERR_STATE = -7        /* call order (e.g. update before any rollout) */
} b2_status;

/* TRAIN_STATUS on the wire (include/ppo_nn.h:15) becomes a flag byte per record. */
#define B2_FLAG_TERMINAL 1u /* TERMINATE: return restarts from 0      (src/ppo_nn.cpp:229-230) */
#define B2_FLAG_TIMEUP 2u   /* TIMEUP: bootstrap with V(pre-step obs) (src/ppo_nn.cpp:226-228) */
#define B2_FLAG_TRUNC 4u    /* engine extension: rollout horizon cut an episode; bootstraps like TIMEUP */
#define B2_FLAG_REWARD 8u   /* b2_batch_compact::rflags only: the record's reward is 1 (getReward is 0 / 1, cart_pole.h:142) */

/* physics arithmetic of the rollout kernel */
#define B2_PREC_F64 0 /* double state + double Diff: the parity-gated path (reference is double) */
#define B2_PREC_F32 1 /* float everything: throughput path, drift vs reference is reported, not gated */
#define B2_PREC_MIXED 2 /* double state + RK accumulators, Diff in float: teacher-forced step within 1e-5 of the
                         reference's double states (gated), free-running drift reported; ~1.5x the F64 rollout rate */

/* environment variants the reference header documents but ships disabled (opt-in; 0 = the shipped cart-pole):
 *   B2_ENV_SWINGUP      the `#if SWINGUP` branches: Diff include/cart_pole.h:154-159, getReward :136-139,
 *                       is_over :100-101, reset range +-0.1 :80-81
 *   B2_ENV_RANDOM_INIT  reset draws u ~ U(-r, r)^4, r = 0.05 (0.1 with SWINGUP): the commented-out line :85, live in
 *                       archived/single_environment/cart_pole.h:87.  Philox stream keyed by (seed, global env id, step)
 *                       instead of the reference's std::random_device, so runs are reproducible and shard-invariant
 *   B2_ENV_EARLY_TERM   advance() stops after the substep at which is_over() holds and returns 1 (commented-out :115,
 *                       live in the archived header :119); the record gets B2_FLAG_TERMINAL (src/main.cpp:81-85)
 * Variants run the double-precision physics with the reference's own operation order (B2_PREC_F64 only). */
#define B2_ENV_SWINGUP 1
#define B2_ENV_RANDOM_INIT 2
#define B2_ENV_EARLY_TERM 4

typedef struct b2_engine* b2_handle;

typedef struct {
  /* -- the reference's 6-line config file (config:1-6; src/main.cpp:356-382, include/cart_MPI.h:13-28) */
  int control_vars;     /* action dim, must be 1 (src/ppo_nn.cpp:107) */
  int obs_vars;         /* observation dim, must be 6 (include/cart_pole.h:120-131) */
  int n_episodes;       /* Nepisodes */
  int n_timestep;       /* N_timestep: steps per episode before TIMEUP */
  int max_timestep;     /* Max_timestep */
  int update_timestep;  /* updateTimestep */
  /* -- hyper-parameters the reference hard-codes at src/main.cpp:196-203 */
  double action_std, lr, beta1, beta2, gamma, eps_clip;
  int k_epochs;
  /* -- engine-only knobs (no reference analogue) */
  int hidden1, hidden2;   /* 64, 32 (include/ppo_nn.h:85-96) */
  int64_t n_envs;         /* environments resident on THIS device */
  int horizon;            /* max T per b2_rollout call (sizes the trajectory buffer) */
  int n_minibatches;      /* 1 == the reference's full-batch epochs (Q9) */
  int precision;          /* B2_PREC_* */
  uint64_t seed;          /* Philox key */
  uint64_t env_offset;    /* global id of local env 0: same global seed => same trajectories for any sharding */
  int device;             /* CUDA ordinal */
  int env_flags;          /* B2_ENV_* (0 = the shipped environment) */
  double gae_lambda;      /* lambda-returns R_k = r_k + gamma ((1 - lambda) V(s_{k+1}) + lambda R_{k+1}) between flagged records;
                             1 (and 0 = unset) = the reference's plain discounted returns (src/ppo_nn.cpp:231) */
  int reserved[4];
} b2_config;

/* One trajectory batch in host memory: what MemoryNN holds (include/ppo_nn.h:20-29).
 * Record k is a transition: obs/action/logp from when the action was chosen,
 * reward/flags from the env's next message.  Layout is time-major [T][N] (k = t*n_envs + e); the
 * reference's merged memory (src/main.cpp:290-303) is the special case n_envs = 1, horizon = B. */
typedef struct {
  int64_t n_envs;
  int64_t horizon;
  const float* obs;     /* [B][obs_vars] */
  const float* action;  /* [B] raw network-space action (not the scaled force) */
  const float* logp;    /* [B] */
  const float* reward;  /* [B] */
  const uint8_t* flags; /* [B] B2_FLAG_* */
} b2_batch;

/* mutable twin of b2_batch for downloads; any pointer may be NULL */
typedef struct {
  float* obs; float* action; float* logp; float* mu; float* reward; uint8_t* flags; float* returns;
} b2_batch_out;

/* Compact host record of b2_rollout_host_compact: 25 B per env-step instead of the 37 B of MemoryNN's layout.  The
 * observation leaves as getobs(4) = (x, xdot, thetadot, theta) (cos / sin are functions of theta, cart_pole.h:128-129)
 * and the 0 / 1 reward rides in the flag byte.  Any field may be NULL. */
typedef struct {
  float* obs4;      /* [T][N][4] */
  float* action;    /* [T][N] */
  float* logp;      /* [T][N] */
  uint8_t* rflags;  /* [T][N] B2_FLAG_* | B2_FLAG_REWARD */
} b2_batch_compact;

typedef struct {
  double loss1, loss2, loss3, loss; /* means over the LAST optimizer step's samples (src/ppo_nn.cpp:292-297) */
  double clip_frac;
  int64_t samples;   /* B */
  int64_t adam_step; /* persists across updates (optimizer is built once, include/ppo_nn.h:166-170) */
  int optimizer_steps;
} b2_update_stats;

/* ---- library ---- */
int b2_abi_version(void);
const char* b2_last_error(void);
int b2_device_count(void);

/* ---- configuration: replaces read_parameters (src/main.cpp:356-382) + hard-coded hyper-parameters */
int b2_config_default(b2_config* cfg);
int b2_config_read(const char* path, b2_config* cfg); /* positional "<int> = comment" lines; unread fields keep their value */

/* ---- lifetime: replaces PPO::PPO (include/ppo_nn.h:146-173) + CartPole construction (src/main.cpp:20) */
int b2_create(const b2_config* cfg, b2_handle* out);
int b2_destroy(b2_handle h);
int b2_get_config(b2_handle h, b2_config* out);

/* ---- parameters.  Order = policy.parameters(): actor W1[H1][S] b1 W2[H2][H1] b2 W3[A][H2] b3, then critic */
int64_t b2_n_params(b2_handle h);
int b2_init_params(b2_handle h, uint64_t seed);              /* torch::nn::Linear default init U(+-1/sqrt(fan_in)) */
int b2_set_params(b2_handle h, const float* p, int64_t n);
int b2_get_params(b2_handle h, float* p, int64_t n);
int b2_set_adam_state(b2_handle h, const float* m, const float* v, int64_t n, int64_t step);
int b2_get_adam_state(b2_handle h, float* m, float* v, int64_t n, int64_t* step);
int b2_get_grads(b2_handle h, float* g, int64_t n);          /* gradient of the last optimizer step (after all-reduce) */
/* params + Adam m, v, step + the Philox counters (rollout step, act / reset calls) + env state, so that a resumed run
 * CONTINUES the noise stream and the episodes of the run that saved (env block restored when n_envs matches) */
int b2_save_checkpoint(b2_handle h, const char* path);
int b2_load_checkpoint(b2_handle h, const char* path);

/* ---- action query: replaces getAction -> PPO::select_action -> ActorCritic::act
 *      (src/ppo_nn.cpp:308-331,196-203,99-125).  eps NULL => engine Philox stream; mu may be NULL. */
int b2_act(b2_handle h, const float* obs, int64_t n, const float* eps, float* action, float* logp, float* mu);
/* ---- value estimate: replaces policy.critic->forward (src/ppo_nn.cpp:189,227) */
int b2_value(b2_handle h, const float* obs, int64_t n, float* value);

/* ---- update on a trajectory batch: replaces PPO::update(MemoryNN) (src/ppo_nn.cpp:205-307) */
int b2_update(b2_handle h, const b2_batch* batch, b2_update_stats* stats);
int b2_returns(b2_handle h, const b2_batch* batch, float* returns); /* the returns part only (:207-236) */
/* which evaluate/backward kernel the last update ran: 0 = FP32 FMA (k_ppo_grad; reference-sized batches), 1 = bf16x3
 * tensor cores (k_ppo_grad_tc; global minibatch >= 32,768 rows; B2_UPDATE_TC=0/1 in the environment forces either),
 * 2 = wide nets (tcgen05 GEMMs) */
int b2_last_update_path(b2_handle h);
/* loss terms of every optimizer step of the last update (n = optimizer_steps; arrays may be NULL) */
int b2_last_update_losses(b2_handle h, double* loss1, double* loss2, double* loss3, double* loss, int n);

/* ---- device-resident path: replaces env_run + NN_run messaging (src/main.cpp:12-148,190-354).
 *      b2_rollout advances every resident env T steps with the current policy (cart_pole.h semantics,
 *      action scaling src/main.cpp:57-60, TIMEUP at n_timestep :86-88); trajectories stay in HBM.
 *      forced_action (host, [T][N]) replaces sampling for replay-parity runs, else NULL. */
int b2_env_reset(b2_handle h);                                 /* CartPole::reset on all envs (include/cart_pole.h:77-90) */
int b2_env_set_state(b2_handle h, const double* u, const int32_t* ep_step, int64_t n); /* u: [n][4] = y1,y2,y3,y4 */
int b2_env_get_state(b2_handle h, double* u, int32_t* ep_step, int64_t n);
int b2_rollout(b2_handle h, int T, const float* forced_action);
/* cart_pole.h-shaped batched env for callers that drive the simulator themselves (no policy, no trajectory):
 * b2_env_step = CartPole::advance (include/cart_pole.h:107-118) on every resident env, force[e] in newtons
 * (already scaled, as env_run passes it, src/main.cpp:69-71); b2_env_observe = getobs (:120-133, 6 doubles per
 * env: x, xdot, thetadot, theta, cos, sin) and getReward (:135-147).  Either output may be NULL. */
int b2_env_step(b2_handle h, const double* force, int64_t n);
int b2_env_observe(b2_handle h, double* obs, double* reward, int64_t n);
/* what advance() returned for every env in the last b2_env_step (always 0 without B2_ENV_EARLY_TERM, cart_pole.h:117) */
int b2_env_terminated(b2_handle h, uint8_t* terminated, int64_t n);
/* b2_rollout whose trajectory lands in HOST buffers (what MemoryNN holds after a collection phase): the horizon is
 * simulated in n_chunks launches and each finished chunk is copied out on a second stream while the next one is
 * being simulated.  Fields of `out` may be NULL (skipped); `returns` is ignored (no returns exist yet).  Pinned
 * buffers (b2_host_alloc) are needed for the overlap; pageable ones work but serialise. */
int b2_rollout_host(b2_handle h, int T, int n_chunks, const b2_batch_out* out);
/* The same pipeline with the compact record (opt-in; the 37-B layout above stays what MemoryNN holds).  The transfer is
 * PCIe-bound: 25 B per env-step leave in less time than the simulation takes, 37 B do not.  Default
 * 6-64-32-1 nets and the shipped environment only. */
int b2_rollout_host_compact(b2_handle h, int T, int n_chunks, const b2_batch_compact* out);
int b2_update_resident(b2_handle h, b2_update_stats* stats);   /* returns + K epochs on the resident trajectory */
int b2_train_iteration(b2_handle h, int T, b2_update_stats* stats); /* rollout + update_resident */
int b2_traj_download(b2_handle h, b2_batch_out* out, double* state_trace /* [T][N][4] or NULL, only if tracing on */);
int b2_set_state_trace(b2_handle h, int enable);               /* record fp64 post-step states (parity tests) */
/* per-env episode bookkeeping of the last rollout (src/main.cpp:74,137-142): completed episodes and their
 * reward sums; returns[e * max_eps + i] for i < counts[e] */
int b2_episode_stats(b2_handle h, int32_t* counts, float* returns, int max_eps, int64_t n);
int64_t b2_total_env_steps(b2_handle h);

/* ---- multi-GPU: one process per GPU, envs sharded, gradients all-reduced (replaces include/cart_MPI.h) */
#define B2_COMM_ID_BYTES 128
int b2_comm_unique_id(char id[B2_COMM_ID_BYTES]);              /* rank 0 makes it, the caller broadcasts it */
int b2_comm_init(b2_handle h, int rank, int world, const char id[B2_COMM_ID_BYTES]);
int b2_comm_world(b2_handle h);
/* how gradients travel between ranks: 0 single GPU; 1 reduce -> ncclAllReduce -> Adam (three launches); 2 NVLink peer
 * memory: the reduce + Adam kernel itself stores every finished gradient element, tagged with the step number, into a
 * buffer of every peer (mapped with cudaIpc), waits for the peers' copies of the same element and adds them in rank order.
 * 2 is chosen when every rank could map every peer; B2_COMM_P2P=0 in the environment forces 1. */
int b2_comm_mode(b2_handle h);

/* ---- measurement helpers (CUDA events on the handle's stream) */
int b2_sync(b2_handle h);
int b2_timer_start(b2_handle h);
int b2_timer_stop_ms(b2_handle h, float* ms);
int64_t b2_kernel_launches(b2_handle h);                       /* kernels this handle has launched so far */
void* b2_stream(b2_handle h);                                  /* cudaStream_t */
int b2_flush_l2(b2_handle h);                                  /* overwrite a >L2 scratch buffer */
int b2_measure_fma_peak(b2_handle h, int fp64, double* tflops); /* CUDA-core FMA issue peak of this GPU (roofline denominator) */
int b2_host_alloc(void** p, size_t bytes);                     /* pinned host memory for the host-buffer entry points */
int b2_host_free(void* p);
/* debug: with B2_GUARD=1 in the environment every device buffer of the library is allocated between two 4-KiB guard
 * zones and poisoned with 0xFF; this checks every guard zone of every live allocation of the process.  Returns the
 * number of corrupted zones (0 = clean; b2_last_error names the first), -1 when guard mode is off. */
int b2_debug_guard_check(void);
/* diagnostic for the tensor-core path of wide nets: C[M][N] = A[M][K] . B[N][K]^T (host fp32 in/out) through the
 * bf16x3 tcgen05 GEMM that the 2x256 update uses; N % 32 == 0, N <= 256; splits > 1 exercises the split-K partials.
 * mn_major = 1: operands are given contraction-major, A[K][M] and B[K][N] (the layout of dW2 = D2^T H1), M, N % 64 == 0;
 * mn_major = 2: A[M][K] as usual, B contraction-major B[K][N] (the layout of Z1 = D2 W2 with W2 as stored), N % 64 == 0 */
int b2_selftest_gemm(b2_handle h, int M, int N, int K, const float* A, const float* B, float* C, int splits, int mn_major);

/* ---- allocation-free entry point ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The discounted returns of b2_returns' segment-per-warp path (k_seg_ends + k_seg_returns, csrc/returns.cu: the
 * recurrence of src/ppo_nn.cpp:207-236 at gae_lambda = 1) on DEVICE buffers of the caller, with the segment list -- the
 * scratch b2_returns keeps in a cudaMalloc'ed buffer of the engine -- in a file-scope `static __device__` array of the
 * library: 8 B per record + 256 B = 2 MiB + 256 B for B2_NOALLOC_MAX_RECORDS records, held by every GPU that loads the
 * library until the process exits.  No engine handle; no cudaMalloc / cudaFree / allocating library call, on the first
 * call or any later one (the first call on a device resolves the array's device address with cudaGetSymbolAddress).
 *   reward, flags, vboot, returns: device pointers, [horizon][n_envs] (flat index t * n_envs + e, as the engine's
 *   trajectory); vboot[k] = V(obs_k) is read only where flags[k] has B2_FLAG_TIMEUP or B2_FLAG_TRUNC (b2_value gives it).
 *   stream: a cudaStream_t, NULL = the default stream.  The call only ENQUEUES (memset of the counter + two kernels) and
 *   does not synchronize.
 * n_envs * horizon > B2_NOALLOC_MAX_RECORDS is rejected with B2_ERR_INVALID (b2_last_error says so); so are NULL
 * pointers and non-positive sizes.  Every call on a GPU shares the ONE array: the call clears what must start at zero
 * (the segment counter), the rest of the array keeps the previous call's contents and is overwritten before it is read;
 * calls that overlap on one GPU (two streams, two threads) race on it -- serialize them. */
#define B2_NOALLOC_MAX_RECORDS 262144
int b2_returns_noalloc(const float* reward, const uint8_t* flags, const float* vboot, int64_t n_envs, int64_t horizon,
                       float gamma, float* returns, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200PPO_H */
