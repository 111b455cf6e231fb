"""Small fixed workload for ncu: BASELINE configs[2] shape, 2 rollouts + 1 one-epoch update.
usage: python profiles/profile_target.py [f64|f32]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi

prec = pkg.PREC_F32 if (len(sys.argv) > 1 and sys.argv[1] == "f32") else pkg.PREC_F64
E = pkg.Engine(capi.default_config(n_envs=65536, horizon=128, precision=prec, k_epochs=1))
E.init_params(0)
for _ in range(2):
    E.rollout(128)
st = E.update_resident()
E.sync()
print("ok", st.loss, E.kernel_launches())
