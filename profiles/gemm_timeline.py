"""Per-CTA timeline of k_gemm_x3 (clock64 stamps per role) at the shape of the configs[4] step GEMMs.
usage: B2_GEMM_DBG=<epi 0|1|2> python profiles/gemm_timeline.py [mn]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import reinforcement_learning_cpp_parallel_b200 as pkg
from reinforcement_learning_cpp_parallel_b200 import capi

E = pkg.Engine(capi.default_config(n_envs=4, horizon=4))
rng = np.random.default_rng(0)
if len(sys.argv) > 1 and sys.argv[1] == "mn":
    A = rng.standard_normal((32768, 256)).astype(np.float32)   # [K][M]
    B = rng.standard_normal((32768, 256)).astype(np.float32)
    E.selftest_gemm(A, B, splits=74, mn_major=True)
else:
    A = rng.standard_normal((32768, 256)).astype(np.float32)
    B = rng.standard_normal((256, 256)).astype(np.float32)
    E.selftest_gemm(A, B, splits=1)
