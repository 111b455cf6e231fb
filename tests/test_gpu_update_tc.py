"""The tensor-core PPO update of the reference's 6-64-32-1 nets (csrc/ppo_update_tc.cu: warp-level bf16x3 MMAs), which
the engine takes once the global minibatch has >= 32,768 rows.  Gates: loss terms and the 12 gradient tensors within
1e-4 relative (north_star) of (a) the libtorch goldens minted from the unmodified reference, with the kernel FORCED onto
the reference-sized batches, and (b) the CPU oracle (double accumulators) on large batches, where it is the default."""
import ctypes as C

import numpy as np
import pytest

from helpers import (DEFAULT_DIMS, EPS_CLIP, GAMMA, SIGMA, Loss, flags_of, gold, oracle, oracle_update, param_slices,
                     relerr)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import reinforcement_learning_cpp_parallel_b200 as p
    return p


def make_engine(pkg, **kw):
    from reinforcement_learning_cpp_parallel_b200 import capi
    return pkg.Engine(capi.default_config(**kw))


def _batch(case):
    return (gold(f"{case}_obs.f32"), gold(f"{case}_action.f32"), gold(f"{case}_logp.f32"),
            gold(f"{case}_reward.f32"), flags_of(case))


@pytest.mark.parametrize("case,tag,K", [("c1", "c1_k1", 1), ("c1", "c1_k5", 5), ("c2", "c2_k1", 1), ("c2", "c2_k5", 5),
                                        ("c1", "c1_term_k1", 1), ("c1", "c1_clip_k1", 1), ("c1", "c1_clip_k5", 5)])
def test_tc_kernel_vs_reference_goldens(pkg, monkeypatch, case, tag, K):
    """B2_UPDATE_TC=1 forces the tensor-core kernel onto the reference's own batches (2,000 / 2,100 records: ragged
    16-sample warp tiles, most of the grid idle): loss terms, gradients, Adam state against libtorch."""
    monkeypatch.setenv("B2_UPDATE_TC", "1")
    obs, act, lp, rw, fl = _batch(case)
    if "term" in tag:
        fl = (gold("c1_term_terminal.u8") | (gold("c1_timeup.u8") << 1)).astype(np.uint8)
    if "clip" in tag:
        lp = gold("c1_clip_logp.f32")
    E = make_engine(pkg, k_epochs=K)
    E.set_params(gold("init_params.f32"))
    st = E.update(obs, act, lp, rw, fl)
    assert E.last_update_path() == 1
    assert st.adam_step == K == gold(f"{tag}_adam_step.i64")[0]
    l1, l2, l3, lt = E.last_update_losses(K)
    g1, g2, g3, gt = (gold(f"{tag}_{k}.f64") for k in ("loss1", "loss2", "loss3", "loss"))
    np.testing.assert_allclose(l1, g1, rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(l2, g2, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(lt, gt, rtol=1e-4, atol=2e-6)
    g, gg = E.get_grads(), gold(f"{tag}_grads.f32")
    worst = max(relerr(g[sl], gg[sl]) for _, sl in param_slices())
    print(f"{tag}: worst gradient tensor relerr {worst:.2e}")
    for name, sl in param_slices():
        assert relerr(g[sl], gg[sl]) < 1e-4, (name, relerr(g[sl], gg[sl]))
    assert relerr(E.get_params(), gold(f"{tag}_params.f32")) < (1e-5 if K == 1 else 5e-5)
    if "clip" in tag:
        assert st.clip_frac > 0.1
    E.close()


def _synthetic(B, seed):
    rng = np.random.default_rng(seed)
    th = 0.2 * rng.standard_normal(B)
    obs = np.stack([rng.standard_normal(B), rng.standard_normal(B), 2 * rng.standard_normal(B), th, np.cos(th), np.sin(th)],
                   1).astype(np.float32)
    rw = (rng.random(B) < 0.8).astype(np.float32)
    fl = np.zeros(B, np.uint8)
    fl[99::100] = 2
    fl[-1] = 2
    return obs, rw, fl


@pytest.mark.parametrize("B", [32768, 32768 + 5, 40000, 131072 + 17])
def test_tc_default_on_large_batches_vs_oracle(pkg, monkeypatch, B):
    """Large batches take the tensor-core kernel by default; one epoch against the CPU oracle (double accumulators)."""
    monkeypatch.delenv("B2_UPDATE_TC", raising=False)
    obs, rw, fl = _synthetic(B, B)
    p0 = gold("c1_k5_params.f32")
    z = np.zeros_like(p0)
    E = make_engine(pkg, k_epochs=1)
    E.set_params(p0)
    rng = np.random.default_rng(1)
    act, lp, _ = E.act(obs, eps=rng.standard_normal(B).astype(np.float32))
    lp = (lp + 0.35 * np.sin(np.arange(B))).astype(np.float32)          # a share of the ratios leaves the clip range
    st = E.update(obs, act, lp, rw, fl)
    assert E.last_update_path() == 1 and st.samples == B
    po, mo, vo, so, losses, go, Ro = oracle_update(p0, z, z, 0, obs.reshape(-1), act, lp, rw, fl, 1)
    g = E.get_grads()
    worst = max(relerr(g[sl], go[sl]) for _, sl in param_slices())
    print(f"B={B}: worst gradient tensor relerr vs oracle {worst:.2e}, clip_frac {st.clip_frac:.3f}")
    for name, sl in param_slices():
        assert relerr(g[sl], go[sl]) < 1e-4, (name, relerr(g[sl], go[sl]))
    l1, l2, l3, lt = E.last_update_losses(1)
    assert abs(l1[0] - losses[0].loss1) <= 1e-4 * abs(losses[0].loss1) + 2e-6
    assert abs(l2[0] - losses[0].loss2) <= 1e-4 * abs(losses[0].loss2) + 1e-6
    assert 0.02 < st.clip_frac < 0.9
    assert relerr(E.get_params(), po) < 1e-5
    E.close()


def test_tc_and_fp32_kernels_agree_over_epochs(pkg, monkeypatch):
    """The same 5-epoch update through either kernel: gradients of the last step and the weights stay together."""
    B = 50000
    obs, rw, fl = _synthetic(B, 3)
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("B2_UPDATE_TC", mode)
        E = make_engine(pkg, k_epochs=5)
        E.set_params(gold("init_params.f32"))
        act, lp, _ = E.act(obs, eps=np.random.default_rng(2).standard_normal(B).astype(np.float32))
        st = E.update(obs, act, lp, rw, fl)
        assert E.last_update_path() == int(mode)
        res[mode] = (E.get_grads(), E.get_params(), st.loss)
        E.close()
    assert relerr(res["1"][0], res["0"][0]) < 1e-4, relerr(res["1"][0], res["0"][0])
    assert relerr(res["1"][1], res["0"][1]) < 2e-5, relerr(res["1"][1], res["0"][1])
    assert abs(res["1"][2] - res["0"][2]) <= 1e-4 * abs(res["0"][2])


def test_tc_resident_update_after_rollout(pkg, monkeypatch):
    """Device-resident path at a time-major [T][N] shape: rollout, returns, tensor-core epochs; against the FP32 kernel."""
    monkeypatch.delenv("B2_UPDATE_TC", raising=False)
    N, T = 1024, 64
    out = {}
    for mode in (None, "0"):
        if mode is not None:
            monkeypatch.setenv("B2_UPDATE_TC", mode)
        E = make_engine(pkg, n_envs=N, horizon=T, seed=9, k_epochs=2)
        E.set_params(gold("c1_k5_params.f32"))
        E.rollout(T)
        st = E.update_resident()
        assert E.last_update_path() == (1 if mode is None else 0)
        out[mode] = (E.get_grads(), E.get_params(), st.loss)
        E.close()
    assert relerr(out[None][0], out["0"][0]) < 1e-4
    assert relerr(out[None][1], out["0"][1]) < 1e-5
