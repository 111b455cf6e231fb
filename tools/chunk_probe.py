import os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi
N, T = 65536, 128
E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, seed=1))
E.init_params(0)
def t(f, n=6):
    f(); f()
    E.sync(); t0 = time.perf_counter()
    for _ in range(n): f()
    E.sync(); return (time.perf_counter() - t0) / n * 1e3
print("single launch           %.3f ms" % t(lambda: E.rollout(T)))
for c in (1, 2, 4, 8, 16, 32):
    print("chunked x%-2d, no copies   %.3f ms" % (c, t(lambda: E.rollout_host_compact(T, {}, n_chunks=c))))
outc = {"obs4": capi.pinned_empty((T, N, 4), np.float32), "action": capi.pinned_empty((T, N), np.float32),
        "logp": capi.pinned_empty((T, N), np.float32), "rflags": capi.pinned_empty((T, N), np.uint8)}
for c in (4, 8, 16):
    print("chunked x%-2d, compact     %.3f ms" % (c, t(lambda: E.rollout_host_compact(T, outc, n_chunks=c))))
out = {"obs": capi.pinned_empty((T, N, 6), np.float32), "action": capi.pinned_empty((T, N), np.float32),
       "logp": capi.pinned_empty((T, N), np.float32), "reward": capi.pinned_empty((T, N), np.float32), "flags": capi.pinned_empty((T, N), np.uint8)}
for c in (8, 16, 24):
    print("chunked x%-2d, 37-B record %.3f ms" % (c, t(lambda: E.rollout_host(T, out, n_chunks=c))))
