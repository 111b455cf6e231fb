#!/usr/bin/env python
"""A/B of the default-net PPO update at BASELINE configs[2] (65,536 envs x 128 steps = 8.4 M samples, 4 full-batch
epochs): FP32 FMA kernel (B2_UPDATE_TC=0) vs bf16x3 tensor-core kernel (=1).  usage: python tools/update_ab.py [envs]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
T, K = 128, 4
for mode in ("0", "1", "0", "1"):
    os.environ["B2_UPDATE_TC"] = mode
    E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, k_epochs=K, seed=0x5EED))
    E.init_params(0)
    E.rollout(T)
    E.update_resident()
    ms = []
    for _ in range(3):
        E.rollout(T)
        E.timer_start()
        st = E.update_resident()
        ms.append(E.timer_stop_ms())
    m = min(ms)
    print(f"B2_UPDATE_TC={mode} path {E.last_update_path()}: {m:.2f} ms per {K}-epoch update = {N * T * K / (m * 1e-3):.4g} "
          f"sample-epochs/s, loss {st.loss:.6f}", flush=True)
    E.close()
