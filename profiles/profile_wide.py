"""BASELINE configs[4] per-GPU share: 2x256 MLPs, 524,288 samples, 32 minibatches x 10 epochs (4M samples over 8 GPUs).
usage: python profiles/profile_wide.py [samples] [minibatches] [epochs] [reps]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi

B = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
M = int(sys.argv[2]) if len(sys.argv) > 2 else 32
K = int(sys.argv[3]) if len(sys.argv) > 3 else 10
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
T = 128
N = B // T
E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, hidden1=256, hidden2=256, k_epochs=K, n_minibatches=M))
E.init_params(0)
rng = np.random.default_rng(0)
th = 0.3 * rng.standard_normal(B)
obs = np.stack([rng.standard_normal(B), rng.standard_normal(B), rng.standard_normal(B), th, np.cos(th), np.sin(th)], 1).astype(np.float32)
act, lp, mu = E.act(obs)
reward = (rng.random(B) < 0.9).astype(np.float32)
flags = np.zeros((T, N), np.uint8)
flags[99] = 2
flags[T - 1] = 4
t0 = time.perf_counter()
try:
    st = E.update(obs, act, lp, reward, flags.reshape(-1), n_envs=N, horizon=T)
    print("first update (incl. upload) %.1f ms, loss %.5f" % (1e3 * (time.perf_counter() - t0), st.loss))
except capi.B2Error as e:           # B2_WIDE_SKIP diagnostic runs compute garbage
    print("  (", e, ")")
    st = None
for _ in range(reps):
    l0 = E.kernel_launches()
    E.timer_start()
    try:
        st = E.update_resident()
    except capi.B2Error as e:       # B2_WIDE_SKIP diagnostic runs compute garbage (non-finite loss): timing still valid
        print("  (", e, ")")
    ms = E.timer_stop_ms()
    se = B * K / (ms * 1e-3)
    print("update_resident %.2f ms  %.3e sample-epochs/s  %.1f TFLOP/s algorithmic (x3 issued)  loss %.5f  launches %d"
          % (ms, se, se * 801792 / 1e12, st.loss if st else float("nan"), E.kernel_launches() - l0))
