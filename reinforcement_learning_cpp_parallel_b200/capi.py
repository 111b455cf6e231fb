"""ctypes binding of include/b200ppo.h.  No compute happens here; every method is one C-ABI call."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    return os.path.join(_HERE, "libb200ppo.so")


FLAG_TERMINAL, FLAG_TIMEUP, FLAG_TRUNC = 1, 2, 4
FLAG_REWARD = 8   # b2_batch_compact.rflags only
PREC_F64, PREC_F32, PREC_MIXED = 0, 1, 2
ENV_SWINGUP, ENV_RANDOM_INIT, ENV_EARLY_TERM = 1, 2, 4
COMM_ID_BYTES = 128


class B2Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200ppo error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    """b2_config (include/b200ppo.h)."""
    _fields_ = [("control_vars", C.c_int), ("obs_vars", C.c_int), ("n_episodes", C.c_int),
                ("n_timestep", C.c_int), ("max_timestep", C.c_int), ("update_timestep", C.c_int),
                ("action_std", C.c_double), ("lr", C.c_double), ("beta1", C.c_double),
                ("beta2", C.c_double), ("gamma", C.c_double), ("eps_clip", C.c_double),
                ("k_epochs", C.c_int), ("hidden1", C.c_int), ("hidden2", C.c_int),
                ("n_envs", C.c_int64), ("horizon", C.c_int), ("n_minibatches", C.c_int),
                ("precision", C.c_int), ("seed", C.c_uint64), ("env_offset", C.c_uint64),
                ("device", C.c_int), ("env_flags", C.c_int), ("gae_lambda", C.c_double), ("reserved", C.c_int * 4)]


class Batch(C.Structure):
    _fields_ = [("n_envs", C.c_int64), ("horizon", C.c_int64), ("obs", C.c_void_p),
                ("action", C.c_void_p), ("logp", C.c_void_p), ("reward", C.c_void_p),
                ("flags", C.c_void_p)]


class BatchOut(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("obs", "action", "logp", "mu", "reward", "flags", "returns")]


class BatchCompact(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("obs4", "action", "logp", "rflags")]


class UpdateStats(C.Structure):
    _fields_ = [("loss1", C.c_double), ("loss2", C.c_double), ("loss3", C.c_double),
                ("loss", C.c_double), ("clip_frac", C.c_double), ("samples", C.c_int64),
                ("adam_step", C.c_int64), ("optimizer_steps", C.c_int)]


# every symbol include/b200ppo.h declares: (restype, argtypes)
_H = C.c_void_p
_P = C.POINTER
SYMBOLS = {
    "b2_abi_version": (C.c_int, []),
    "b2_debug_guard_check": (C.c_int, []),
    "b2_last_update_path": (C.c_int, [_H]),
    "b2_last_error": (C.c_char_p, []),
    "b2_device_count": (C.c_int, []),
    "b2_config_default": (C.c_int, [_P(Config)]),
    "b2_config_read": (C.c_int, [C.c_char_p, _P(Config)]),
    "b2_create": (C.c_int, [_P(Config), _P(_H)]),
    "b2_destroy": (C.c_int, [_H]),
    "b2_get_config": (C.c_int, [_H, _P(Config)]),
    "b2_n_params": (C.c_int64, [_H]),
    "b2_init_params": (C.c_int, [_H, C.c_uint64]),
    "b2_set_params": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "b2_get_params": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "b2_set_adam_state": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    "b2_get_adam_state": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, _P(C.c_int64)]),
    "b2_get_grads": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "b2_save_checkpoint": (C.c_int, [_H, C.c_char_p]),
    "b2_load_checkpoint": (C.c_int, [_H, C.c_char_p]),
    "b2_act": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2_value": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_void_p]),
    "b2_update": (C.c_int, [_H, _P(Batch), _P(UpdateStats)]),
    "b2_returns": (C.c_int, [_H, _P(Batch), C.c_void_p]),
    "b2_last_update_losses": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "b2_env_reset": (C.c_int, [_H]),
    "b2_env_set_state": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64]),
    "b2_env_get_state": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64]),
    "b2_rollout": (C.c_int, [_H, C.c_int, C.c_void_p]),
    "b2_rollout_host": (C.c_int, [_H, C.c_int, C.c_int, _P(BatchOut)]),
    "b2_rollout_host_compact": (C.c_int, [_H, C.c_int, C.c_int, _P(BatchCompact)]),
    "b2_env_step": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "b2_env_observe": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64]),
    "b2_env_terminated": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "b2_update_resident": (C.c_int, [_H, _P(UpdateStats)]),
    "b2_train_iteration": (C.c_int, [_H, C.c_int, _P(UpdateStats)]),
    "b2_traj_download": (C.c_int, [_H, _P(BatchOut), C.c_void_p]),
    "b2_set_state_trace": (C.c_int, [_H, C.c_int]),
    "b2_episode_stats": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int, C.c_int64]),
    "b2_total_env_steps": (C.c_int64, [_H]),
    "b2_comm_unique_id": (C.c_int, [C.c_void_p]),
    "b2_comm_init": (C.c_int, [_H, C.c_int, C.c_int, C.c_void_p]),
    "b2_comm_world": (C.c_int, [_H]),
    "b2_comm_mode": (C.c_int, [_H]),
    "b2_sync": (C.c_int, [_H]),
    "b2_timer_start": (C.c_int, [_H]),
    "b2_timer_stop_ms": (C.c_int, [_H, _P(C.c_float)]),
    "b2_kernel_launches": (C.c_int64, [_H]),
    "b2_stream": (C.c_void_p, [_H]),
    "b2_flush_l2": (C.c_int, [_H]),
    "b2_measure_fma_peak": (C.c_int, [_H, C.c_int, _P(C.c_double)]),
    "b2_host_alloc": (C.c_int, [_P(C.c_void_p), C.c_size_t]),
    "b2_host_free": (C.c_int, [C.c_void_p]),
    "b2_selftest_gemm": (C.c_int, [_H, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "b2_returns_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]),
}
NOALLOC_MAX_RECORDS = 262144   # B2_NOALLOC_MAX_RECORDS

_lib = None


def lib():
    """Load libb200ppo.so (built in-tree by __graft_entry__.build()).  Fails loudly if absent."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no Python/CPU fallback for the engine)")
        L = C.CDLL(p)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise B2Error(rc, lib().b2_last_error().decode())


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, np.float32)


def default_config(**kw):
    c = Config()
    _ck(lib().b2_config_default(C.byref(c)))
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(k)
        setattr(c, k, v)
    return c


def read_config(path, **kw):
    c = default_config()
    _ck(lib().b2_config_read(os.fsencode(path), C.byref(c)))
    for k, v in kw.items():
        setattr(c, k, v)
    return c


class Engine:
    """One b2_handle.  Methods mirror the C ABI one to one and move data through HOST numpy buffers."""

    def __init__(self, cfg=None, **kw):
        self.L = lib()
        self.cfg = cfg if cfg is not None else default_config(**kw)
        h = _H()
        _ck(self.L.b2_create(C.byref(self.cfg), C.byref(h)))
        self.h = h
        self.P = int(self.L.b2_n_params(h))

    def close(self):
        if getattr(self, "h", None):
            self.L.b2_destroy(self.h)
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- parameters
    def set_params(self, p):
        p = _f32(p)
        _ck(self.L.b2_set_params(self.h, _vp(p), p.size))

    def get_params(self):
        p = np.empty(self.P, np.float32)
        _ck(self.L.b2_get_params(self.h, _vp(p), p.size))
        return p

    def init_params(self, seed):
        _ck(self.L.b2_init_params(self.h, seed))

    def set_adam_state(self, m, v, step):
        m, v = _f32(m), _f32(v)
        _ck(self.L.b2_set_adam_state(self.h, _vp(m), _vp(v), m.size, step))

    def get_adam_state(self):
        m = np.empty(self.P, np.float32)
        v = np.empty(self.P, np.float32)
        s = C.c_int64()
        _ck(self.L.b2_get_adam_state(self.h, _vp(m), _vp(v), self.P, C.byref(s)))
        return m, v, s.value

    def get_grads(self):
        g = np.empty(self.P, np.float32)
        _ck(self.L.b2_get_grads(self.h, _vp(g), g.size))
        return g

    def save_checkpoint(self, path):
        _ck(self.L.b2_save_checkpoint(self.h, os.fsencode(path)))

    def load_checkpoint(self, path):
        _ck(self.L.b2_load_checkpoint(self.h, os.fsencode(path)))

    # ---- ppo_nn.h-shaped calls
    def act(self, obs, eps=None, out=None):
        """getAction on n observations (src/ppo_nn.cpp:308-331): returns (action, logp, mu)."""
        obs = _f32(obs).reshape(-1, int(self.cfg.obs_vars))
        n = obs.shape[0]
        eps = _f32(eps)
        a, lp, mu = out if out is not None else (np.empty(n, np.float32), np.empty(n, np.float32),
                                                 np.empty(n, np.float32))
        _ck(self.L.b2_act(self.h, _vp(obs), n, _vp(eps), _vp(a), _vp(lp), _vp(mu)))
        return a, lp, mu

    def value(self, obs):
        obs = _f32(obs).reshape(-1, int(self.cfg.obs_vars))
        v = np.empty(obs.shape[0], np.float32)
        _ck(self.L.b2_value(self.h, _vp(obs), obs.shape[0], _vp(v)))
        return v

    def _batch(self, obs, action, logp, reward, flags, n_envs, horizon):
        keep = [_f32(obs), _f32(action), _f32(logp), _f32(reward),
                None if flags is None else np.ascontiguousarray(flags, np.uint8)]
        B = keep[3].size
        if n_envs is None:
            n_envs, horizon = 1, B
        assert n_envs * horizon == B
        b = Batch(n_envs, horizon, *[None if k is None else k.ctypes.data for k in keep])
        return b, keep

    def update(self, obs, action, logp, reward, flags, n_envs=None, horizon=None):
        """PPO::update(MemoryNN) (src/ppo_nn.cpp:205-307) on a host batch."""
        b, keep = self._batch(obs, action, logp, reward, flags, n_envs, horizon)
        st = UpdateStats()
        _ck(self.L.b2_update(self.h, C.byref(b), C.byref(st)))
        return st

    def returns(self, obs, reward, flags, n_envs=None, horizon=None):
        b, keep = self._batch(obs, None, None, reward, flags, n_envs, horizon)
        R = np.empty(keep[3].size, np.float32)
        _ck(self.L.b2_returns(self.h, C.byref(b), _vp(R)))
        return R

    def last_update_path(self):
        """0 FP32 FMA kernel, 1 bf16x3 tensor-core kernel, 2 wide nets."""
        return int(self.L.b2_last_update_path(self.h))

    def last_update_losses(self, n):
        out = [np.empty(n, np.float64) for _ in range(4)]
        _ck(self.L.b2_last_update_losses(self.h, *[_vp(o) for o in out], n))
        return out

    # ---- device-resident path
    def env_reset(self):
        _ck(self.L.b2_env_reset(self.h))

    def env_set_state(self, u, ep_step=None):
        u = np.ascontiguousarray(u, np.float64).reshape(-1, 4)
        es = None if ep_step is None else np.ascontiguousarray(ep_step, np.int32)
        _ck(self.L.b2_env_set_state(self.h, _vp(u), _vp(es), u.shape[0]))

    def env_get_state(self):
        n = int(self.cfg.n_envs)
        u = np.empty((n, 4), np.float64)
        es = np.empty(n, np.int32)
        _ck(self.L.b2_env_get_state(self.h, _vp(u), _vp(es), n))
        return u, es

    def env_step(self, force):
        """CartPole::advance on every resident env (include/cart_pole.h:107-118); force in newtons."""
        f = np.ascontiguousarray(force, np.float64).reshape(-1)
        _ck(self.L.b2_env_step(self.h, _vp(f), f.size))

    def env_observe(self):
        """getobs + getReward (include/cart_pole.h:120-147): (obs[n][6] f64, reward[n] f64)."""
        n = int(self.cfg.n_envs)
        o = np.empty((n, 6), np.float64)
        r = np.empty(n, np.float64)
        _ck(self.L.b2_env_observe(self.h, _vp(o), _vp(r), n))
        return o, r

    def env_terminated(self):
        """what advance() returned per env in the last env_step (include/cart_pole.h:115,117)."""
        n = int(self.cfg.n_envs)
        t = np.empty(n, np.uint8)
        _ck(self.L.b2_env_terminated(self.h, _vp(t), n))
        return t

    def set_state_trace(self, on=True):
        _ck(self.L.b2_set_state_trace(self.h, int(on)))

    def rollout(self, T, forced_action=None):
        fa = _f32(forced_action)
        if fa is not None:
            assert fa.size == T * int(self.cfg.n_envs)
        _ck(self.L.b2_rollout(self.h, T, _vp(fa)))

    def rollout_host(self, T, out, n_chunks=8):
        """b2_rollout_host: rollout T steps, trajectory delivered into the HOST arrays of `out` (pinned for overlap)."""
        bo = BatchOut(*[out[k].ctypes.data if k in out else None
                        for k in ("obs", "action", "logp", "mu", "reward", "flags")], None)
        _ck(self.L.b2_rollout_host(self.h, T, n_chunks, C.byref(bo)))
        return out

    def rollout_host_compact(self, T, out, n_chunks=8):
        """b2_rollout_host_compact: the 25-B record (obs4 [T][N][4], action, logp, rflags = flags | FLAG_REWARD)."""
        bc = BatchCompact(*[out[k].ctypes.data if k in out else None for k in ("obs4", "action", "logp", "rflags")])
        _ck(self.L.b2_rollout_host_compact(self.h, T, n_chunks, C.byref(bc)))
        return out

    def update_resident(self):
        st = UpdateStats()
        _ck(self.L.b2_update_resident(self.h, C.byref(st)))
        return st

    def train_iteration(self, T):
        st = UpdateStats()
        _ck(self.L.b2_train_iteration(self.h, T, C.byref(st)))
        return st

    def traj_download(self, T, fields=("obs", "action", "logp", "mu", "reward", "flags"), state_trace=False,
                      out=None):
        n = int(self.cfg.n_envs)
        shp = {"obs": ((T, n, int(self.cfg.obs_vars)), np.float32), "flags": ((T, n), np.uint8)}
        res = out if out is not None else {}
        for f in fields:
            if f not in res:
                s, dt = shp.get(f, ((T, n), np.float32))
                res[f] = np.empty(s, dt)
        bo = BatchOut(*[res[k].ctypes.data if k in fields else None
                        for k in ("obs", "action", "logp", "mu", "reward", "flags", "returns")])
        st = None
        if state_trace:
            st = res.setdefault("state_trace", np.empty((T, n, 4), np.float64))
        _ck(self.L.b2_traj_download(self.h, C.byref(bo), _vp(st)))
        return res

    def episode_stats(self, max_eps):
        n = int(self.cfg.n_envs)
        cnt = np.empty(n, np.int32)
        ret = np.empty((n, max_eps), np.float32)
        _ck(self.L.b2_episode_stats(self.h, _vp(cnt), _vp(ret), max_eps, n))
        return cnt, ret

    def total_env_steps(self):
        return int(self.L.b2_total_env_steps(self.h))

    # ---- multi-GPU
    @staticmethod
    def comm_unique_id():
        buf = (C.c_char * COMM_ID_BYTES)()
        _ck(lib().b2_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, rank, world, uid):
        buf = (C.c_char * COMM_ID_BYTES).from_buffer_copy(uid)
        _ck(self.L.b2_comm_init(self.h, rank, world, buf))

    def comm_mode(self):
        """0 single GPU, 1 ncclAllReduce, 2 NVLink peer-memory exchange fused into the Adam kernel."""
        return int(self.L.b2_comm_mode(self.h))

    # ---- measurement
    def sync(self):
        _ck(self.L.b2_sync(self.h))

    def timer_start(self):
        _ck(self.L.b2_timer_start(self.h))

    def timer_stop_ms(self):
        ms = C.c_float()
        _ck(self.L.b2_timer_stop_ms(self.h, C.byref(ms)))
        return ms.value

    def kernel_launches(self):
        return int(self.L.b2_kernel_launches(self.h))

    def flush_l2(self):
        _ck(self.L.b2_flush_l2(self.h))

    def measure_fma_peak(self, fp64):
        t = C.c_double()
        _ck(self.L.b2_measure_fma_peak(self.h, int(fp64), C.byref(t)))
        return t.value


def _selftest_gemm(self, A, B, splits=1, mn_major=False):
    """C = A @ B.T (A [M][K], B [N][K]) or, mn_major, C = A.T @ B (A [K][M], B [K][N]) through the bf16x3 tcgen05
    GEMM of the wide-net update (diagnostic)."""
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    if int(mn_major) in (2, 4):  # A [M][K] as usual, B contraction-major [K][N]: C = A @ B (4: on the CTA-pair kernel)
        M, K = A.shape
        N = B.shape[1]
        assert B.shape[0] == K
    elif int(mn_major) == 3:    # the CTA-pair kernel, usual operand layouts
        M, K = A.shape
        N = B.shape[0]
        assert B.shape[1] == K
    elif mn_major:
        K, M = A.shape
        N = B.shape[1]
        assert B.shape[0] == K
    else:
        M, K = A.shape
        N = B.shape[0]
        assert B.shape[1] == K
    out = np.empty((M, N), np.float32)
    _ck(self.L.b2_selftest_gemm(self.h, M, N, K, _vp(A), _vp(B), _vp(out), int(splits), int(mn_major)))
    return out


Engine.selftest_gemm = _selftest_gemm


def pinned_empty(shape, dtype):
    """numpy array over cudaMallocHost memory (freed when the array is collected)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    _ck(lib().b2_host_alloc(C.byref(p), max(n, 1)))
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[id(buf)] = (buf, p)
    return arr


_PINNED = {}
