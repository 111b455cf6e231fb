#!/usr/bin/env python
"""Small invocations of every hot kernel for tools/sanitize.sh (compute-sanitizer memcheck / racecheck / synccheck /
initcheck).  Each case goes through the C ABI exactly as the tests do.

  python tools/sanitize_target.py rollout|update|wide|rollout_host|all
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import reinforcement_learning_cpp_parallel_b200 as pkg  # noqa: E402
from reinforcement_learning_cpp_parallel_b200 import capi  # noqa: E402


def synth(B, seed=0):
    rng = np.random.default_rng(seed)
    th = 0.2 * rng.standard_normal(B)
    obs = np.stack([rng.standard_normal(B), rng.standard_normal(B), rng.standard_normal(B), th, np.cos(th), np.sin(th)],
                   1).astype(np.float32)
    rw = (rng.random(B) < 0.8).astype(np.float32)
    fl = np.zeros(B, np.uint8)
    fl[99::100] = 2
    fl[-1] = 4
    return obs, rw, fl


def case_rollout():
    # 32,768 envs = 1,024 groups > 592 resident warps: the time-sliced grid with its release/acquire hand-off
    E = pkg.Engine(capi.default_config(n_envs=32768, horizon=8, seed=3, k_epochs=1))
    E.rollout(8)
    E.update_resident()       # k_bootstrap + k_returns (per-env walk) + k_ppo_grad + k_reduce at a resident shape
    E.sync()
    E.close()


def case_update():
    for B in (1000, 1999):    # ragged tiles of k_ppo_grad (128-sample tiles), the segment-per-warp returns path
        obs, rw, fl = synth(B, B)
        E = pkg.Engine(capi.default_config(k_epochs=2))
        a, lp, _ = E.act(obs)
        E.update(obs, a, lp, rw, fl)
        E.value(obs)
        E.close()
    E = pkg.Engine(capi.default_config(k_epochs=1, gae_lambda=0.9))     # lambda path: V at every record
    obs, rw, fl = synth(1500, 7)
    a, lp, _ = E.act(obs)
    E.update(obs, a, lp, rw, fl)
    E.close()


def case_wide():
    B = 1000                  # one optimizer step of the 256x256 nets: l1, forward GEMM, head, dX GEMM, dW2 GEMM + dw1, reduce + Adam
    obs, rw, fl = synth(B, 11)
    E = pkg.Engine(capi.default_config(n_envs=B, horizon=1, hidden1=256, hidden2=256, k_epochs=1, n_minibatches=1))
    E.init_params(4)
    a, lp, _ = E.act(obs)
    E.update(obs, a, lp, rw, np.full(B, 2, np.uint8))
    E.close()


def case_rollout_host():
    N, T = 1024, 16
    E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, seed=5))
    for pinned in (True, False):
        mk = (lambda s, d: capi.pinned_empty(s, d)) if pinned else (lambda s, d: np.empty(s, d))
        out = {"obs": mk((T, N, 6), np.float32), "action": mk((T, N), np.float32), "logp": mk((T, N), np.float32),
               "mu": mk((T, N), np.float32), "reward": mk((T, N), np.float32), "flags": mk((T, N), np.uint8)}
        E.rollout_host(T, out, n_chunks=4)
    E.close()


CASES = {"rollout": case_rollout, "update": case_update, "wide": case_wide, "rollout_host": case_rollout_host}

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    for name, fn in CASES.items():
        if which in ("all", name):
            fn()
            print(f"[sanitize_target] {name} done", flush=True)
