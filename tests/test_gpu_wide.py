"""GPU parity of the tensor-core (tcgen05, bf16x3) path used for wide hidden layers (BASELINE configs[4], 2x256)."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    E = pkg.Engine(capi.default_config(n_envs=4, horizon=4))
    yield E
    E.close()


# tolerance: bf16x3 keeps ~2^-16 relative per product; the gate for gradients is 1e-4 (north_star)
@pytest.mark.parametrize("M,N,K,splits", [(128, 256, 256, 1), (128, 64, 64, 1), (300, 256, 256, 1), (1000, 128, 192, 1),
                                          (256, 256, 1024, 4), (100, 64, 70, 1), (200, 96, 130, 2), (64, 256, 4096, 7)])
def test_gemm_bf16x3_matches_fp64(eng, M, N, K, splits):
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
    C = eng.selftest_gemm(A, B, splits)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    err = np.abs(C - ref).max() / np.abs(ref).max()
    rel = np.linalg.norm(C - ref) / np.linalg.norm(ref)
    print(f"gemm {M}x{N}x{K} splits {splits}: max err {err:.2e}, rel fro {rel:.2e}")
    assert rel < 2e-5, (rel, err)
