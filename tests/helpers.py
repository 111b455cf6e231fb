"""Shared test plumbing: golden-vector loader and the ctypes view of the CPU oracle.

The oracle (oracle/liboracle.so, built from oracle/oracle.c) is TEST INFRASTRUCTURE: it is the
checker the CUDA engine is compared with.  Nothing under reinforcement_learning_cpp_parallel_b200/
imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
ORACLE_DIR = os.path.join(ROOT, "oracle")

_DT = {"f32": np.float32, "f64": np.float64, "u8": np.uint8, "i64": np.int64}


def gold(name):
    """tests/golden/<name>: raw little-endian array, dtype from the extension."""
    ext = name.rsplit(".", 1)[1]
    return np.fromfile(os.path.join(GOLD, name), dtype=_DT[ext])


class Dims(C.Structure):
    _fields_ = [("S", C.c_int), ("H1", C.c_int), ("H2", C.c_int), ("A", C.c_int)]


class Loss(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("loss1", "loss2", "loss3", "loss", "clip_frac")]


DEFAULT_DIMS = Dims(6, 64, 32, 1)
SIGMA, EPS_CLIP, GAMMA, LR, B1, B2 = 0.5, 0.2, 0.99, 0.0003, 0.9, 0.999  # src/main.cpp:196-203

_f = np.ctypeslib.ndpointer(np.float32, flags="C")
_d = np.ctypeslib.ndpointer(np.float64, flags="C")
_u8 = np.ctypeslib.ndpointer(np.uint8, flags="C")
_i32 = np.ctypeslib.ndpointer(np.int32, flags="C")

_oracle = None


def oracle():
    global _oracle
    if _oracle is not None:
        return _oracle
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
    L = C.CDLL(so)
    P = C.POINTER
    L.orc_n_params.argtypes = [P(Dims)]
    L.orc_net_params.argtypes = [P(Dims), C.c_int]
    L.orc_cartpole_advance.argtypes = [_d, C.c_double]
    L.orc_cartpole_obs.argtypes = [_d, _f]
    L.orc_cartpole_reward.argtypes = [_d]
    L.orc_cartpole_reward.restype = C.c_double
    L.orc_scale_action.argtypes = [C.c_float]
    L.orc_scale_action.restype = C.c_double
    L.orc_logprob.argtypes = [C.c_float, C.c_float, C.c_double]
    L.orc_logprob.restype = C.c_float
    L.orc_entropy_const.argtypes = [C.c_double]
    L.orc_entropy_const.restype = C.c_float
    L.orc_act.argtypes = [P(Dims), _f, _f, C.c_int, C.c_void_p, C.c_double, _f, _f, _f]
    L.orc_value.argtypes = [P(Dims), _f, _f, C.c_int, _f]
    L.orc_returns.argtypes = [P(Dims), _f, _f, _f, _u8, C.c_long, C.c_double, _f]
    L.orc_loss_grad.argtypes = [P(Dims), _f, _f, _f, _f, _f, C.c_void_p, C.c_long, C.c_double,
                                C.c_double, P(Loss), _f]
    L.orc_adam.argtypes = [_f, _f, _f, _f, C.c_long, C.c_long] + [C.c_double] * 4
    L.orc_update.argtypes = [P(Dims), _f, _f, _f, P(C.c_long), _f, _f, _f, _f, _u8, C.c_long, C.c_int,
                             C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                             C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_gauss.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
    L.orc_gauss.restype = C.c_float
    L.orc_rollout.argtypes = [P(Dims), _f, C.c_long, C.c_int, C.c_int, C.c_double, C.c_uint64,
                              C.c_uint64, C.c_uint64, C.c_void_p, _d, _i32, C.c_void_p] + [C.c_void_p] * 7
    L.orc_cartpole_advance_v.argtypes = [C.c_int, _d, C.c_int, C.c_double]
    L.orc_cartpole_reward_v.argtypes = [C.c_int, _d]
    L.orc_cartpole_reward_v.restype = C.c_double
    L.orc_cartpole_is_over.argtypes = [C.c_int, _d, C.c_int]
    L.orc_cartpole_init_state.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, _d]
    L.orc_rollout_v.argtypes = [P(Dims), _f, C.c_long, C.c_int, C.c_int, C.c_double, C.c_uint64,
                                C.c_uint64, C.c_uint64, C.c_void_p, C.c_int, _d, _i32, C.c_void_p] + [C.c_void_p] * 7
    _oracle = L
    return L


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def oracle_update(params, m, v, step, obs, action, logp, reward, flags, K, dims=DEFAULT_DIMS,
                  sigma=SIGMA, eps_clip=EPS_CLIP, gamma=GAMMA, lr=LR, b1=B1, b2=B2):
    """orc_update wrapper: returns (params, m, v, step, losses[K], last_grads, returns)."""
    L = oracle()
    params, m, v = params.copy(), m.copy(), v.copy()
    B = reward.shape[0]
    P = params.shape[0]
    losses = (Loss * K)()
    g = np.zeros(P, np.float32)
    R = np.zeros(B, np.float32)
    st = C.c_long(step)
    L.orc_update(C.byref(dims), params, m, v, C.byref(st), np.ascontiguousarray(obs), action, logp,
                 reward, flags, B, K, sigma, eps_clip, gamma, lr, b1, b2,
                 C.cast(losses, C.c_void_p), ptr(g), ptr(R))
    return params, m, v, st.value, losses, g, R


def oracle_rollout(params, N, T, n_timestep=100, seed=0x5EED, env0=0, step0=0, forced=None,
                   state=None, ep_step=None, dims=DEFAULT_DIMS, sigma=SIGMA):
    L = oracle()
    state = np.zeros((N, 4), np.float64) if state is None else state.copy()
    ep_step = np.zeros(N, np.int32) if ep_step is None else ep_step.copy()
    ep_ret = np.zeros(N, np.float64)
    out = dict(obs=np.zeros((T, N, 6), np.float32), action=np.zeros((T, N), np.float32),
               logp=np.zeros((T, N), np.float32), mu=np.zeros((T, N), np.float32),
               reward=np.zeros((T, N), np.float32), flags=np.zeros((T, N), np.uint8),
               state_trace=np.zeros((T, N, 4), np.float64))
    L.orc_rollout(C.byref(dims), params, N, T, n_timestep, sigma, seed, env0, step0,
                  ptr(forced), state, ep_step, ptr(ep_ret), ptr(out["obs"]), ptr(out["action"]),
                  ptr(out["logp"]), ptr(out["mu"]), ptr(out["reward"]), ptr(out["flags"]),
                  ptr(out["state_trace"]))
    out.update(state=state, ep_step=ep_step, ep_ret=ep_ret)
    return out


ENV_SWINGUP, ENV_RANDOM_INIT, ENV_EARLY_TERM = 1, 2, 4  # oracle.h ORC_ENV_* == b200ppo.h B2_ENV_*


def oracle_rollout_v(params, N, T, env_flags, n_timestep=100, seed=0x5EED, env0=0, step0=0, forced=None,
                     state=None, ep_step=None, dims=DEFAULT_DIMS, sigma=SIGMA):
    """orc_rollout_v: the rollout with the opt-in env variants (SWINGUP / random init / early termination)."""
    L = oracle()
    state = np.zeros((N, 4), np.float64) if state is None else state.copy()
    ep_step = np.zeros(N, np.int32) if ep_step is None else ep_step.copy()
    ep_ret = np.zeros(N, np.float64)
    out = dict(obs=np.zeros((T, N, 6), np.float32), action=np.zeros((T, N), np.float32),
               logp=np.zeros((T, N), np.float32), mu=np.zeros((T, N), np.float32),
               reward=np.zeros((T, N), np.float32), flags=np.zeros((T, N), np.uint8),
               state_trace=np.zeros((T, N, 4), np.float64))
    L.orc_rollout_v(C.byref(dims), params, N, T, n_timestep, sigma, seed, env0, step0,
                    ptr(forced), env_flags, state, ep_step, ptr(ep_ret), ptr(out["obs"]), ptr(out["action"]),
                    ptr(out["logp"]), ptr(out["mu"]), ptr(out["reward"]), ptr(out["flags"]),
                    ptr(out["state_trace"]))
    out.update(state=state, ep_step=ep_step, ep_ret=ep_ret)
    return out


def flags_of(case):
    """bit0 terminal, bit1 timeup (include/b200ppo.h B2_FLAG_*)."""
    return (gold(f"{case}_terminal.u8") | (gold(f"{case}_timeup.u8") << 1)).astype(np.uint8)


def relerr(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


# parameter blocks in policy.parameters() order (include/ppo_nn.h:83-123)
def param_slices(dims=DEFAULT_DIMS):
    S, H1, H2, A = dims.S, dims.H1, dims.H2, dims.A
    out, o = [], 0
    for net, nout in (("actor", A), ("critic", 1)):
        for name, n in (("W1", H1 * S), ("b1", H1), ("W2", H2 * H1), ("b2", H2), ("W3", nout * H2), ("b3", nout)):
            out.append((f"{net}.{name}", slice(o, o + n)))
            o += n
    return out
