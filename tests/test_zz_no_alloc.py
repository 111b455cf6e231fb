"""b2_returns_noalloc (include/b200ppo.h): the allocation-free variant of the segment-per-warp returns, its segment
list in a static device array of the library.  No counterpart in the reference,
so the checker is the existing entry point b2_returns on the same batch.

Collected last on purpose: a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

from helpers import GAMMA, gold, relerr


def _batch(B, n_envs, seed):
    """[horizon][n_envs] records with TIMEUP / TERMINATE cuts (src/ppo_nn.cpp:226-231), ragged episode lengths"""
    rng = np.random.default_rng(seed)
    th = 0.2 * rng.standard_normal(B)
    obs = np.stack([rng.standard_normal(B), rng.standard_normal(B), rng.standard_normal(B), th, np.cos(th), np.sin(th)],
                   1).astype(np.float32)
    rw = (rng.random(B) < 0.8).astype(np.float32)
    fl = np.zeros(B, np.uint8)
    fl[rng.random(B) < 0.01] = 2        # TIMEUP: bootstrap from V(obs_k)
    fl[rng.random(B) < 0.004] = 1       # TERMINATE
    fl[(B // 2) - 1] = 4                # TRUNC
    return obs, rw, fl


def test_noalloc_rejects_oversize_and_null_without_a_gpu():
    """argument checks come before any CUDA call: sizes beyond the static array and NULL pointers -> B2_ERR_INVALID"""
    from reinforcement_learning_cpp_parallel_b200 import capi
    L = capi.lib()
    one = C.c_void_p(256)   # never dereferenced: the size check rejects first
    rc = L.b2_returns_noalloc(one, one, one, 1, capi.NOALLOC_MAX_RECORDS + 1, GAMMA, one, None)
    assert rc == -1 and b"B2_NOALLOC_MAX_RECORDS" in L.b2_last_error()
    rc = L.b2_returns_noalloc(one, one, one, 1 << 20, 1 << 20, GAMMA, one, None)
    assert rc == -1 and b"B2_NOALLOC_MAX_RECORDS" in L.b2_last_error()
    assert L.b2_returns_noalloc(None, one, one, 1, 16, GAMMA, one, None) == -1
    assert L.b2_returns_noalloc(one, one, one, 0, 16, GAMMA, one, None) == -1


@pytest.mark.gpu
def test_noalloc_returns_match_b2_returns_on_two_calls_in_a_row():
    """two calls in a row through the C ABI -- the second on a SMALLER batch of another shape, so stale segment ends of the
    first call sit in the shared scratch beyond what the second one writes -- each against b2_returns on the same batch:
    inside the 1e-5 gate of test_returns, and in fact bit for bit (same kernels, same order)"""
    import torch
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    L = capi.lib()
    E = pkg.Engine(capi.default_config(n_envs=1, horizon=2000))
    E.set_params(gold("c1_k5_params.f32"))
    dev = torch.device("cuda:0")
    for n_envs, horizon, seed in [(1, capi.NOALLOC_MAX_RECORDS, 11), (3, 700, 12), (1, 2000, 13)]:
        B = n_envs * horizon
        obs, rw, fl = _batch(B, n_envs, seed)
        ref = E.returns(obs, rw, fl, n_envs=n_envs, horizon=horizon)   # the existing entry point (engine-owned scratch)
        V = E.value(obs)                                               # bootstrap values the caller supplies
        d_rw, d_fl, d_v = (torch.from_numpy(a).to(dev) for a in (rw, fl, V))
        d_out = torch.full((B,), float("nan"), dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        rc = L.b2_returns_noalloc(d_rw.data_ptr(), d_fl.data_ptr(), d_v.data_ptr(), n_envs, horizon, GAMMA, d_out.data_ptr(), None)
        assert rc == 0, L.b2_last_error()
        torch.cuda.synchronize()
        out = d_out.cpu().numpy()
        assert np.isfinite(out).all()
        assert relerr(out, ref) < 1e-5
        np.testing.assert_allclose(out, ref, rtol=1e-5, atol=1e-5)
        np.testing.assert_array_equal(out, ref)
    E.close()
