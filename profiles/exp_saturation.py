"""How many warps per SM sub-partition does k_rollout need to saturate its pipe?
block=128 gives every CTA one warp per sub-partition; n_envs = 148*128*k puts k CTAs on every SM.
usage: B2_ROLLOUT_BLOCK=128 python profiles/exp_saturation.py [f64|f32]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi

prec = pkg.PREC_F32 if (len(sys.argv) > 1 and sys.argv[1] == "f32") else pkg.PREC_F64
blk = int(os.environ.get("B2_ROLLOUT_BLOCK", "64"))
T = 128
for n in [148 * 128 * k for k in (1, 2, 4, 6)] + [65536, 131072, 524288]:
    E = pkg.Engine(capi.default_config(n_envs=n, horizon=T, precision=prec, k_epochs=1))
    E.init_params(0)
    for _ in range(2):
        E.rollout(T)
    E.sync()
    ms = []
    for _ in range(5):
        E.timer_start(); E.rollout(T); ms.append(E.timer_stop_ms())
    m = sorted(ms)[len(ms) // 2]
    print(f"block {blk} n_envs {n:7d} warps/SMSP {n / (148 * 128):5.2f}  {m:8.3f} ms  {n * T / m / 1e6:8.3f} G env-steps/s  "
          f"ms per (warp/SMSP) {m / (n / (148 * 128)):.3f}", flush=True)
    E.close()
