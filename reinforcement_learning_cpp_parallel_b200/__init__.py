"""B200-native PPO rollout-and-update engine (drop-in for the hot path of
haotianh9/Reinforcement_learning_cpp_parallel).

The product is ``libb200ppo.so`` (hand-written sm_100a CUDA kernels behind the C ABI of
``include/b200ppo.h``) plus the C++ shims under ``include/compat``.  This package is only the
thin ctypes view used by tests/ and bench.py; it contains no compute and no fallback: if the
shared library is missing, importing :mod:`capi` raises.
"""
from .capi import (B2Error, Batch, Config, ENV_EARLY_TERM, ENV_RANDOM_INIT, ENV_SWINGUP, Engine, FLAG_TERMINAL,
                   FLAG_TIMEUP, FLAG_TRUNC, PREC_F32, PREC_F64, PREC_MIXED, UpdateStats, lib, lib_path)

__all__ = ["B2Error", "Batch", "Config", "ENV_EARLY_TERM", "ENV_RANDOM_INIT", "ENV_SWINGUP", "Engine", "FLAG_TERMINAL", "FLAG_TIMEUP", "FLAG_TRUNC",
           "PREC_F32", "PREC_F64", "PREC_MIXED", "UpdateStats", "lib", "lib_path"]
