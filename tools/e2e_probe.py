"""Where the e2e rollout leg's time goes: b2_rollout_host with different output sets (wall clock per call)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi
N, T = 65536, 128
E = pkg.Engine(capi.default_config(n_envs=N, horizon=T))
full = {"obs": capi.pinned_empty((T, N, 6), np.float32), "action": capi.pinned_empty((T, N), np.float32),
        "logp": capi.pinned_empty((T, N), np.float32), "reward": capi.pinned_empty((T, N), np.float32),
        "flags": capi.pinned_empty((T, N), np.uint8)}
def run(label, out, chunks, reps=6):
    for _ in range(2): E.rollout_host(T, out, n_chunks=chunks)
    E.sync(); t0 = time.perf_counter()
    for _ in range(reps): E.rollout_host(T, out, n_chunks=chunks)
    E.sync(); dt = (time.perf_counter() - t0) / reps
    print(f"{label:34s} chunks {chunks:2d}: {dt * 1e3:7.3f} ms  {N * T / dt:.3e} env-steps/s")
for _ in range(3): E.rollout(T)
E.sync(); t0 = time.perf_counter()
for _ in range(6): E.rollout(T)
E.sync(); print(f"b2_rollout (resident, one launch): {(time.perf_counter() - t0) / 6 * 1e3:.3f} ms")
for ch in (8,):
    run("no outputs", {}, ch)
    run("small fields only (13 B)", {k: v for k, v in full.items() if k != "obs"}, ch)
    run("obs only (24 B)", {"obs": full["obs"]}, ch)
    run("all (37 B)", full, ch)
E.close()
