"""GPU parity tests: the CUDA engine, called through the C ABI (ctypes -> libb200ppo.so), against
(a) the golden vectors minted from the unmodified reference and (b) the CPU oracle on the same inputs.
Tolerances are the ones BASELINE.json's north_star states: env states / returns 1e-5 relative (the fp64
path is held to 1e-9), PPO loss and gradients 1e-4 relative."""
import ctypes as C

import numpy as np
import pytest

from helpers import (DEFAULT_DIMS, EPS_CLIP, GAMMA, SIGMA, Loss, flags_of, gold, oracle, oracle_rollout,
                     oracle_update, param_slices, relerr)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import reinforcement_learning_cpp_parallel_b200 as p
    return p


def make_engine(pkg, **kw):
    from reinforcement_learning_cpp_parallel_b200 import capi
    return pkg.Engine(capi.default_config(**kw))


def _case_tn(case):
    """golden arrays are env-major [N][T]; the engine is time-major [T][N]."""
    nenv = 1 if case == "c1" else 3
    a = gold(f"{case}_action.f32")
    per = a.size // nenv
    tn = lambda x, *tail: np.ascontiguousarray(x.reshape(nenv, per, *tail).swapaxes(0, 1))
    return nenv, per, tn


# ------------------------------------------------------------------------------------------- T3
@pytest.mark.parametrize("case", ["c1", "c2"])
def test_policy_forward_vs_reference(pkg, case):
    E = make_engine(pkg)
    E.set_params(gold("init_params.f32"))
    obs = gold(f"{case}_obs.f32").reshape(-1, 6)
    n = obs.shape[0]
    a, lp, mu = E.act(obs, eps=np.zeros(n, np.float32))
    np.testing.assert_allclose(mu, gold(f"{case}_mu.f32"), rtol=1e-6, atol=1e-6)
    np.testing.assert_array_equal(a, mu)  # eps = 0 -> action is the mean
    np.testing.assert_allclose(E.value(obs), gold(f"{case}_value.f32"), rtol=1e-6, atol=1e-6)
    # log-prob of the reference's own sampled actions: eps := (a - mu)/sigma reproduces them
    ga, gmu = gold(f"{case}_action.f32"), gold(f"{case}_mu.f32")
    eps = ((ga.astype(np.float64) - gmu.astype(np.float64)) / SIGMA).astype(np.float32)
    a2, lp2, _ = E.act(obs, eps=eps)
    np.testing.assert_allclose(a2, ga, rtol=0, atol=5e-7)  # eps is rounded to f32: <= 2 ulp
    np.testing.assert_allclose(lp2, gold(f"{case}_logp.f32"), rtol=1e-5, atol=2e-6)
    E.close()


def test_act_matches_oracle_on_random_inputs(pkg):
    L = oracle()
    rng = np.random.default_rng(7)
    n = 4099  # ragged: not a multiple of the block size
    obs = (rng.standard_normal((n, 6)) * np.array([1, 2, 2, 0.3, 1, 1])).astype(np.float32)
    eps = rng.standard_normal(n).astype(np.float32)
    params = (gold("c1_k80_params.f32"))
    E = make_engine(pkg)
    E.set_params(params)
    a, lp, mu = E.act(obs, eps=eps)
    om, oa, ol, ov = (np.zeros(n, np.float32) for _ in range(4))
    L.orc_act(C.byref(DEFAULT_DIMS), params, obs, n, eps.ctypes.data_as(C.c_void_p), SIGMA, om, oa, ol)
    L.orc_value(C.byref(DEFAULT_DIMS), params, obs, n, ov)
    np.testing.assert_allclose(mu, om, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(a, oa, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(lp, ol, rtol=1e-5, atol=5e-6)
    np.testing.assert_allclose(E.value(obs), ov, rtol=1e-6, atol=1e-6)
    # empty input is a no-op
    a0, lp0, mu0 = E.act(np.zeros((0, 6), np.float32), eps=np.zeros(0, np.float32))
    assert a0.size == 0
    E.close()


# ------------------------------------------------------------------------------------------ T1/T2
@pytest.mark.parametrize("case", ["c1", "c2"])
def test_env_replay_fp64(pkg, case):
    """Replay the reference's recorded actions: fp64 state, f32 obs, reward and flags for whole episodes."""
    nenv, T, tn = _case_tn(case)
    E = make_engine(pkg, n_envs=nenv, horizon=T, precision=pkg.PREC_F64)
    E.set_params(gold("init_params.f32"))
    E.set_state_trace(True)
    E.rollout(T, forced_action=tn(gold(f"{case}_action.f32")))
    out = E.traj_download(T, state_trace=True)
    st = tn(gold(f"{case}_state.f64"), 4)
    scale = np.maximum(np.abs(st).max(axis=-1, keepdims=True), 1e-300)
    err = np.abs(out["state_trace"] - st) / scale
    assert err.max() < 1e-9, err.max()          # north_star asks 1e-5; FMA contraction only costs ~1e-13
    np.testing.assert_allclose(out["obs"], tn(gold(f"{case}_obs.f32"), 6), rtol=1e-6, atol=1e-7)
    np.testing.assert_array_equal(out["reward"], tn(gold(f"{case}_reward.f32")))
    np.testing.assert_array_equal((out["flags"] >> 1) & 1, tn(gold(f"{case}_timeup.u8")))
    np.testing.assert_array_equal(out["flags"] & 1, tn(gold(f"{case}_terminal.u8")))
    np.testing.assert_allclose(out["mu"], tn(gold(f"{case}_mu.f32")), rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(out["logp"], tn(gold(f"{case}_logp.f32")), rtol=1e-5, atol=5e-6)
    # episode bookkeeping (src/main.cpp:74,137-142): T/100 episodes per env, reward sums
    cnt, ret = E.episode_stats(T // 100 + 1)
    assert (cnt == T // 100).all()
    ref = tn(gold(f"{case}_reward.f32")).reshape(T // 100, 100, nenv).sum(axis=1).T
    np.testing.assert_array_equal(ret[:, : T // 100], ref)
    E.close()


@pytest.mark.parametrize("kat", ["kat_a", "kat_b", "kat_c"])
def test_env_known_answers_fp64(pkg, kat):
    """Forces straight from the KAT files: choose actions a = F/10 so that src/main.cpp:57-60 scaling
    reproduces the force to f32, then compare against an oracle replay of the SAME f32 forces."""
    L = oracle()
    F = gold(f"{kat}_force.f64")
    a = (F / 10.0).astype(np.float32)
    T = a.size
    E = make_engine(pkg, n_envs=1, horizon=T, n_timestep=T + 1)
    E.set_state_trace(True)
    E.rollout(T, forced_action=a.reshape(T, 1))
    got = E.traj_download(T, state_trace=True)["state_trace"][:, 0]
    u = np.zeros(4)
    for k in range(T):
        L.orc_cartpole_advance(u, L.orc_scale_action(float(a[k])))
        assert np.abs(got[k] - u).max() <= 1e-9 * max(np.abs(u).max(), 1e-3), (k, got[k], u)
    E.close()


@pytest.mark.parametrize("case", ["c1"])
def test_env_replay_fp32_drift_is_bounded(pkg, case):
    """The f32 throughput path is NOT parity-gated; this only guards against gross bugs and
    records the drift."""
    nenv, T, tn = _case_tn(case)
    E = make_engine(pkg, n_envs=nenv, horizon=T, precision=pkg.PREC_F32)
    E.set_params(gold("init_params.f32"))
    E.set_state_trace(True)
    E.rollout(T, forced_action=tn(gold(f"{case}_action.f32")))
    out = E.traj_download(T, state_trace=True)
    st = tn(gold(f"{case}_state.f64"), 4)
    scale = np.maximum(np.abs(st).max(axis=-1, keepdims=True), 1e-3)
    err = (np.abs(out["state_trace"] - st) / scale).max(axis=(1, 2)).reshape(-1, 100)
    print("fp32 drift: first-step max %.3e, end-of-episode median %.3e max %.3e"
          % (err[:, 0].max(), np.median(err[:, -1]), err[:, -1].max()))
    assert err[:, 0].max() < 1e-4 and np.median(err[:, -1]) < 5e-2
    E.close()


# ------------------------------------------------------------------------------------------- T4
@pytest.mark.parametrize("case,flagcase", [("c1", "c1"), ("c2", "c2"), ("c1", "c1_term")])
def test_returns(pkg, case, flagcase):
    fl = flags_of(case)
    if flagcase.endswith("_term"):
        fl = (gold("c1_term_terminal.u8") | (gold("c1_timeup.u8") << 1)).astype(np.uint8)
    E = make_engine(pkg)
    params = gold("init_params.f32")
    E.set_params(params)
    obs, rw = gold(f"{case}_obs.f32"), gold(f"{case}_reward.f32")
    R = E.returns(obs, rw, fl)  # merged order == n_envs 1, horizon B
    ref = gold(f"{flagcase}_returns.f32")
    assert relerr(R, ref) < 1e-5
    np.testing.assert_allclose(R, ref, rtol=1e-5, atol=1e-5)
    Ro = np.zeros_like(R)
    oracle().orc_returns(C.byref(DEFAULT_DIMS), params, obs, rw, fl, rw.size, GAMMA, Ro)
    np.testing.assert_allclose(R, Ro, rtol=2e-6, atol=2e-6)
    if case == "c2":  # the same batch in time-major [T][N] layout gives the same numbers
        nenv, T, tn = _case_tn(case)
        R2 = E.returns(tn(obs, 6), tn(rw), tn(fl), n_envs=nenv, horizon=T)
        np.testing.assert_array_equal(R2.reshape(T, nenv), tn(R))
    E.close()


# ----------------------------------------------------------------------------------------- T5/T6
def _batch(case):
    return (gold(f"{case}_obs.f32"), gold(f"{case}_action.f32"), gold(f"{case}_logp.f32"),
            gold(f"{case}_reward.f32"), flags_of(case))


@pytest.mark.parametrize("case,tag,K", [("c1", "c1_k1", 1), ("c1", "c1_k5", 5), ("c1", "c1_k80", 80),
                                        ("c2", "c2_k1", 1), ("c2", "c2_k5", 5), ("c2", "c2_k80", 80),
                                        ("c1", "c1_term_k1", 1), ("c1", "c1_clip_k1", 1), ("c1", "c1_clip_k5", 5)])
def test_update_vs_reference(pkg, case, tag, K):
    obs, act, lp, rw, fl = _batch(case)
    if "term" in tag:
        fl = (gold("c1_term_terminal.u8") | (gold("c1_timeup.u8") << 1)).astype(np.uint8)
    if "clip" in tag:
        lp = gold("c1_clip_logp.f32")
    E = make_engine(pkg, k_epochs=K)
    E.set_params(gold("init_params.f32"))
    st = E.update(obs, act, lp, rw, fl)
    assert st.adam_step == K == gold(f"{tag}_adam_step.i64")[0] and st.optimizer_steps == K
    l1, l2, l3, lt = E.last_update_losses(K)
    g1, g2, g3, gt = (gold(f"{tag}_{k}.f64") for k in ("loss1", "loss2", "loss3", "loss"))
    np.testing.assert_allclose(l1, g1, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(l2, g2, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(l3, g3, rtol=0, atol=1e-6)
    np.testing.assert_allclose(lt, gt, rtol=1e-4, atol=1e-6)
    g, gg = E.get_grads(), gold(f"{tag}_grads.f32")
    tol = 1e-4 if K <= 5 else 2e-3  # 80 chained Adam steps amplify f32 rounding (same bound as the oracle test)
    for name, sl in param_slices():
        assert relerr(g[sl], gg[sl]) < tol, (name, relerr(g[sl], gg[sl]))
    assert relerr(E.get_params(), gold(f"{tag}_params.f32")) < (1e-5 if K <= 5 else 1e-4)
    m, v, step = E.get_adam_state()
    assert relerr(m, gold(f"{tag}_adam_m.f32")) < tol and relerr(v, gold(f"{tag}_adam_v.f32")) < tol
    if "clip" in tag:
        assert st.clip_frac > 0.1
    E.close()


def test_update_vs_oracle_tight(pkg):
    """Same batch, same weights: CUDA vs the CPU oracle (double accumulators) at K=1 -- tighter than 1e-4."""
    obs, act, lp, rw, fl = _batch("c2")
    p0 = gold("c1_k5_params.f32")  # weights that are not the init
    z = np.zeros_like(p0)
    E = make_engine(pkg, k_epochs=1)
    E.set_params(p0)
    E.update(obs, act, lp, rw, fl)
    po, mo, vo, so, losses, go, Ro = oracle_update(p0, z, z, 0, obs, act, lp, rw, fl, 1)
    g = E.get_grads()
    for name, sl in param_slices():
        assert relerr(g[sl], go[sl]) < 2e-5, (name, relerr(g[sl], go[sl]))
    assert relerr(E.get_params(), po) < 1e-6
    E.close()


def test_adam_step_counter_persists(pkg):
    obs, act, lp, rw, fl = _batch("c1")
    E = make_engine(pkg, k_epochs=2)
    E.set_params(gold("init_params.f32"))
    E.update(obs, act, lp, rw, fl)
    st = E.update(obs, act, lp, rw, fl)
    assert st.adam_step == 4 == gold("c1_k2x2_adam_step.i64")[0]
    assert relerr(E.get_params(), gold("c1_k2x2_params.f32")) < 1e-5
    assert relerr(E.get_grads(), gold("c1_k2x2_grads.f32")) < 1e-4
    E.close()


def test_update_ragged_batch_sizes(pkg):
    """Tile tails: B not a multiple of the 128-sample tile, B smaller than one tile, B == 1."""
    obs, act, lp, rw, fl = _batch("c1")
    p0 = gold("init_params.f32")
    z = np.zeros_like(p0)
    for B in (1, 77, 128, 129, 1000, 1999):
        f = fl[:B].copy()
        f[-1] |= 2  # close the last segment like the reference's batches always are
        E = make_engine(pkg, k_epochs=1)
        E.set_params(p0)
        E.update(obs[: 6 * B], act[:B], lp[:B], rw[:B], f)
        po, mo, vo, so, losses, go, Ro = oracle_update(p0, z, z, 0, obs[: 6 * B], act[:B], lp[:B], rw[:B], f, 1)
        assert relerr(E.get_grads(), go) < 5e-5, (B, relerr(E.get_grads(), go))
        l1, l2, l3, lt = E.last_update_losses(1)
        assert abs(l1[0] - losses[0].loss1) <= 1e-5 * abs(losses[0].loss1) + 1e-6
        assert abs(l2[0] - losses[0].loss2) <= 1e-5 * abs(losses[0].loss2) + 1e-6
        E.close()


def test_checkpoint_roundtrip(pkg, tmp_path):
    obs, act, lp, rw, fl = _batch("c1")
    E = make_engine(pkg, k_epochs=3)
    E.set_params(gold("init_params.f32"))
    E.update(obs, act, lp, rw, fl)
    path = str(tmp_path / "ck.b2")
    E.save_checkpoint(path)
    E2 = make_engine(pkg, k_epochs=3)
    E2.load_checkpoint(path)
    np.testing.assert_array_equal(E.get_params(), E2.get_params())
    a, b = E.get_adam_state(), E2.get_adam_state()
    np.testing.assert_array_equal(a[0], b[0]); np.testing.assert_array_equal(a[1], b[1]); assert a[2] == b[2] == 3
    E.update(obs, act, lp, rw, fl); E2.update(obs, act, lp, rw, fl)
    np.testing.assert_array_equal(E.get_params(), E2.get_params())  # resume is bit-exact
    E.close(); E2.close()
