"""Pins the CPU oracle (oracle/oracle.c) to vectors minted from the UNMODIFIED reference
(oracle/ref_harness.cpp -> tests/golden/).  CPU-only; runs in the `-m "not gpu"` suite."""
import ctypes as C

import numpy as np
import pytest

import helpers as H

from helpers import (B1, B2, DEFAULT_DIMS, EPS_CLIP, GAMMA, LR, SIGMA, Loss, flags_of, gold, oracle,
                     oracle_rollout, oracle_update, param_slices, relerr)


def test_param_count():
    L = oracle()
    assert L.orc_n_params(C.byref(DEFAULT_DIMS)) == 5122 == gold("init_params.f32").size


@pytest.mark.parametrize("kat", ["kat_a", "kat_b", "kat_c", "kat_fast_a", "kat_fast_b"])
def test_env_known_answers(kat):
    """cart_pole.h advance/getobs/getReward: fp64 states must match the header to ~1 ulp/step."""
    L = oracle()
    F = gold(f"{kat}_force.f64")
    S = gold(f"{kat}_state.f64").reshape(-1, 4)
    O = gold(f"{kat}_obs.f64").reshape(-1, 6)
    Rw = gold(f"{kat}_reward.f64")
    u = np.zeros(4)
    o = np.zeros(6, np.float32)
    for k in range(F.size):
        L.orc_cartpole_advance(u, float(F[k]))
        # same libm, same op order, no contraction: bit-exact
        assert np.array_equal(u, S[k]), (k, u, S[k])
        L.orc_cartpole_obs(u, o)
        assert np.array_equal(o, O[k].astype(np.float32))
        assert L.orc_cartpole_reward(u) == Rw[k]


def test_survey_quoted_vectors():
    """SURVEY.md s8(c) quotes these numbers from the unmodified header."""
    S = gold("kat_a_state.f64").reshape(-1, 4)
    assert S[0, 0] == 1.9512895542387541e-4 and S[0, 2] == -2.9283702350559661e-4
    assert S[4, 3] == -1.3111649698297223e-1
    assert gold("kat_b_state.f64").reshape(-1, 4)[-1, 0] == 18.189014582134952


@pytest.mark.parametrize("case", ["c1", "c2"])
def test_action_scaling_and_env_replay(case):
    """Replay the reference's recorded actions through the oracle env: force, obs, reward, state."""
    L = oracle()
    a = gold(f"{case}_action.f32")
    force = gold(f"{case}_force.f32")
    for k in range(a.size):
        assert np.float32(L.orc_scale_action(float(a[k]))) == force[k]
    nenv = 1 if case == "c1" else 3
    per = a.size // nenv
    obs = gold(f"{case}_obs.f32").reshape(nenv, per, 6)
    st = gold(f"{case}_state.f64").reshape(nenv, per, 4)
    rw = gold(f"{case}_reward.f32").reshape(nenv, per)
    tu = gold(f"{case}_timeup.u8").reshape(nenv, per)
    params = gold("init_params.f32")
    forced = np.ascontiguousarray(a.reshape(nenv, per).T)  # [T][N]
    out = oracle_rollout(params, nenv, per, forced=forced)
    assert np.array_equal(out["state_trace"], st.transpose(1, 0, 2))
    assert np.array_equal(out["obs"], obs.transpose(1, 0, 2))
    assert np.array_equal(out["reward"], rw.T)
    assert np.array_equal((out["flags"] >> 1) & 1, tu.T)


@pytest.mark.parametrize("case", ["c1", "c2"])
def test_policy_forward(case):
    L = oracle()
    params = gold("init_params.f32")
    obs = gold(f"{case}_obs.f32")
    a = gold(f"{case}_action.f32")
    n = a.size
    mu = np.zeros(n, np.float32)
    act = np.zeros(n, np.float32)
    lp = np.zeros(n, np.float32)
    v = np.zeros(n, np.float32)
    L.orc_act(C.byref(DEFAULT_DIMS), params, obs, n, None, SIGMA, mu, act, lp)
    L.orc_value(C.byref(DEFAULT_DIMS), params, obs, n, v)
    np.testing.assert_allclose(mu, gold(f"{case}_mu.f32"), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(v, gold(f"{case}_value.f32"), rtol=1e-6, atol=1e-6)
    # logprob of the RECORDED action under the recorded mean (src/ppo_nn.cpp:8-54)
    lp2 = np.array([L.orc_logprob(float(a[i]), float(gold(f"{case}_mu.f32")[i]), SIGMA) for i in range(n)],
                   np.float32)
    np.testing.assert_allclose(lp2, gold(f"{case}_logp.f32"), rtol=1e-6, atol=1e-6)


def test_entropy_constant():
    L = oracle()
    h = L.orc_entropy_const(SIGMA)
    assert abs(h - 7.8465867) < 1e-5
    assert abs(-0.01 * h - gold("c1_k1_loss3.f64")[0]) < 1e-7


@pytest.mark.parametrize("case,flagcase", [("c1", "c1"), ("c2", "c2"), ("c1", "c1_term")])
def test_returns(case, flagcase):
    L = oracle()
    params = gold("init_params.f32")
    obs = gold(f"{case}_obs.f32")
    rw = gold(f"{case}_reward.f32")
    fl = flags_of(case)
    if flagcase.endswith("_term"):
        fl = (gold("c1_term_terminal.u8") | (gold("c1_timeup.u8") << 1)).astype(np.uint8)
    R = np.zeros(rw.size, np.float32)
    L.orc_returns(C.byref(DEFAULT_DIMS), params, obs, rw, fl, rw.size, GAMMA, R)
    ref = gold(f"{flagcase}_returns.f32")
    np.testing.assert_allclose(R, ref, rtol=2e-6, atol=1e-6)
    assert relerr(R, ref) < 1e-6


def _batch(case):
    return (gold(f"{case}_obs.f32"), gold(f"{case}_action.f32"), gold(f"{case}_logp.f32"),
            gold(f"{case}_reward.f32"), flags_of(case))


@pytest.mark.parametrize("case,tag,K", [("c1", "c1_k1", 1), ("c1", "c1_k5", 5), ("c1", "c1_k80", 80),
                                        ("c2", "c2_k1", 1), ("c2", "c2_k5", 5), ("c2", "c2_k80", 80),
                                        ("c1", "c1_term_k1", 1), ("c1", "c1_clip_k1", 1),
                                        ("c1", "c1_clip_k5", 5)])
def test_update_against_reference(case, tag, K):
    """PPO::update: per-epoch loss terms, last-epoch gradients (12 tensors), Adam state, weights."""
    obs, act, lp, rw, fl = _batch(case)
    if "term" in tag:
        fl = (gold("c1_term_terminal.u8") | (gold("c1_timeup.u8") << 1)).astype(np.uint8)
    if "clip" in tag:
        lp = gold("c1_clip_logp.f32")
    p0 = gold("init_params.f32")
    z = np.zeros_like(p0)
    p, m, v, step, losses, g, R = oracle_update(p0, z, z, 0, obs, act, lp, rw, fl, K)
    assert step == K == gold(f"{tag}_adam_step.i64")[0]
    l1, l2, l3, lt = (gold(f"{tag}_{k}.f64") for k in ("loss1", "loss2", "loss3", "loss"))
    assert l1.size == K
    for k in range(K):
        assert abs(losses[k].loss1 - l1[k]) <= 1e-4 * abs(l1[k]) + 1e-6, (k, losses[k].loss1, l1[k])
        assert abs(losses[k].loss2 - l2[k]) <= 1e-4 * abs(l2[k]) + 1e-6
        assert abs(losses[k].loss3 - l3[k]) <= 1e-6
        assert abs(losses[k].loss - lt[k]) <= 1e-4 * abs(lt[k]) + 1e-6
    gg = gold(f"{tag}_grads.f32")
    tol = 1e-4 if K <= 5 else 2e-3  # 80 chained Adam steps amplify f32 rounding (sign-like update)
    for name, sl in param_slices():
        assert relerr(g[sl], gg[sl]) < tol, (name, relerr(g[sl], gg[sl]))
    assert relerr(p, gold(f"{tag}_params.f32")) < (1e-5 if K <= 5 else 1e-4)
    assert relerr(m, gold(f"{tag}_adam_m.f32")) < tol
    assert relerr(v, gold(f"{tag}_adam_v.f32")) < tol


def test_adam_step_counter_persists():
    """Two update() calls on one PPO: the Adam step keeps counting (optimizer built once)."""
    obs, act, lp, rw, fl = _batch("c1")
    p0 = gold("init_params.f32")
    z = np.zeros_like(p0)
    p, m, v, step, *_ = oracle_update(p0, z, z, 0, obs, act, lp, rw, fl, 2)
    p, m, v, step, losses, g, R = oracle_update(p, m, v, step, obs, act, lp, rw, fl, 2)
    assert step == 4 == gold("c1_k2x2_adam_step.i64")[0]
    assert relerr(p, gold("c1_k2x2_params.f32")) < 1e-5
    assert relerr(g, gold("c1_k2x2_grads.f32")) < 1e-4
    l1 = gold("c1_k2x2_loss1.f64")
    assert abs(losses[1].loss1 - l1[3]) <= 1e-4 * abs(l1[3]) + 1e-6


def test_clipping_is_exercised():
    """The *_clip goldens (perturbed old log-probs) put a large share of ratios outside [1-eps,1+eps] (T5)."""
    L = oracle()
    obs, act, _, rw, fl = _batch("c1")
    lp = gold("c1_clip_logp.f32")
    p = gold("init_params.f32")
    R = gold("c1_returns.f32")
    out = Loss()
    g = np.zeros(p.size, np.float32)
    L.orc_loss_grad(C.byref(DEFAULT_DIMS), p, obs, act, lp, R, None, rw.size, SIGMA, EPS_CLIP,
                    C.byref(out), g)
    assert out.clip_frac > 0.10, out.clip_frac


def test_gauss_stream_statistics():
    L = oracle()
    x = np.array([L.orc_gauss(0x5EED, e, 7) for e in range(20000)])
    assert abs(x.mean()) < 0.03 and abs(x.std() - 1) < 0.03
    assert L.orc_gauss(1, 2, 3) == L.orc_gauss(1, 2, 3) != L.orc_gauss(1, 2, 4)


# ---- widths other than the reference's 64/32: goldens minted by tests/golden/make_wide_golden.py (torch CPU) ----
@pytest.mark.parametrize("H1,H2", [(128, 64), (256, 256)])
def test_oracle_wide_nets_match_torch_golden(H1, H2):
    L = H.oracle()
    dims = H.Dims(6, H1, H2, 1)
    tag = f"wide{H1}x{H2}"
    params, obs = H.gold(f"{tag}_params.f32"), H.gold(f"{tag}_obs.f32").reshape(-1, 6)
    action, oldlp, R = H.gold(f"{tag}_action.f32"), H.gold(f"{tag}_oldlogp.f32"), H.gold(f"{tag}_R.f32")
    B = R.shape[0]
    assert L.orc_n_params(C.byref(dims)) == params.shape[0]
    v = np.zeros(B, np.float32)
    L.orc_value(C.byref(dims), params, obs, B, v)
    assert H.relerr(v, H.gold(f"{tag}_V.f32")) < 1e-6
    mu, a2, lp = (np.zeros(B, np.float32) for _ in range(3))
    L.orc_act(C.byref(dims), params, obs, B, H.ptr(np.zeros(B, np.float32)), H.SIGMA, mu, a2, lp)
    assert np.abs(mu - H.gold(f"{tag}_mu.f32")).max() < 1e-6
    g = np.zeros_like(params)
    out = H.Loss()
    L.orc_loss_grad(C.byref(dims), params, obs, action, oldlp, R, None, B, H.SIGMA, H.EPS_CLIP, C.byref(out), g)
    gl = H.gold(f"{tag}_loss.f64")
    assert abs(out.loss1 - gl[0]) < 1e-5 * abs(gl[0]) and abs(out.loss2 - gl[1]) < 1e-5 * abs(gl[1])
    assert abs(out.loss - gl[3]) < 1e-5 * abs(gl[3]) and abs(out.clip_frac - gl[4]) < 1e-7
    gg = H.gold(f"{tag}_grads.f32")
    for name, sl in H.param_slices(dims):
        assert H.relerr(g[sl], gg[sl]) < 1e-4, name


# ---- env variants: SWINGUP branches, early termination, scripted re-initialisation ----
# KATs minted from the reference's own headers by oracle/ref_env_variants.cpp (make -C oracle golden_variants):
# [E][T][8] = F, y1..y4, reward, terminate, step.
@pytest.mark.parametrize("name,fl", [("swingup", H.ENV_SWINGUP), ("early", H.ENV_EARLY_TERM),
                                     ("early_swingup", H.ENV_EARLY_TERM | H.ENV_SWINGUP)])
def test_env_variant_kat(name, fl):
    L = H.oracle()
    kat = H.gold(f"envvar_{name}.f64").reshape(8, 620, 8)
    init = H.gold(f"envvar_{name}_init.f64").reshape(8, 64, 4)
    n_term = 0
    for e in range(kat.shape[0]):
        u = init[e, 0].copy()
        slot, step = 0, 0
        for t in range(kat.shape[1]):
            F, ref_u, ref_r, ref_term, ref_step = kat[e, t, 0], kat[e, t, 1:5], kat[e, t, 5], kat[e, t, 6], kat[e, t, 7]
            step += 1
            term = L.orc_cartpole_advance_v(fl, u, step, F)
            assert step == int(ref_step)
            assert term == int(ref_term), (name, e, t)
            # same operations in the same order, no contraction: bit-identical to the reference header
            assert np.array_equal(u, ref_u), (name, e, t, u, ref_u)
            assert L.orc_cartpole_reward_v(fl, u) == ref_r
            if term:
                n_term += 1
                slot = (slot + 1) % 64
                u = init[e, slot].copy()
                step = 0
    if fl & H.ENV_EARLY_TERM:
        assert n_term > 10
    else:
        assert n_term == 0  # the shipped advance() never terminates (include/cart_pole.h:115,117)


def test_env_variant_default_flags_equal_shipped_env():
    """flags = 0 must be the shipped environment bit for bit."""
    L = H.oracle()
    rng = np.random.default_rng(3)
    for _ in range(20):
        u = rng.uniform(-0.05, 0.05, 4)
        F = float(rng.uniform(-10, 10))
        a, b = u.copy(), u.copy()
        L.orc_cartpole_advance(a, F)
        assert L.orc_cartpole_advance_v(0, b, 1, F) == 0
        assert np.array_equal(a, b)
        assert L.orc_cartpole_reward(a) == L.orc_cartpole_reward_v(0, b)


def test_env_variant_init_state_range_and_streams():
    L = H.oracle()
    u = np.zeros(4)
    L.orc_cartpole_init_state(0, 1, 2, 3, u)
    assert np.all(u == 0)  # shipped reset: zero state (include/cart_pole.h:86)
    seen = []
    for fl, r in ((H.ENV_RANDOM_INIT, 0.05), (H.ENV_RANDOM_INIT | H.ENV_SWINGUP, 0.1)):
        draws = np.zeros((2000, 4))
        for i in range(2000):
            L.orc_cartpole_init_state(fl, 0x5EED, i % 50, i // 50, u)
            draws[i] = u
        assert np.all(np.abs(draws) < r) and np.abs(draws).max() > 0.95 * r      # :81-84 ranges
        assert abs(draws.mean()) < 0.05 * r and len(np.unique(draws)) == draws.size
        seen.append(draws)
    assert np.allclose(seen[1], 2 * seen[0])  # same stream, scaled range


# ---- obs_vars = 4 (getobs(4), include/cart_pole.h:120-127; PPO(state_dim = 4)): goldens from the unmodified reference
# built with state_dim 4 (oracle/ref_harness.cpp golden_obs4) ----
D4 = H.Dims(4, 64, 32, 1)


def test_obs4_forward_returns_and_update_match_reference():
    L = oracle()
    p0 = gold("o4_init_params.f32")
    assert p0.size == L.orc_n_params(C.byref(D4)) == 4866
    obs = gold("o4_obs.f32").reshape(-1, 4)
    B = obs.shape[0]
    act, lp_old, rew = gold("o4_action.f32"), gold("o4_logp.f32"), gold("o4_reward.f32")
    fl = flags_of("o4")
    # the 4-vector is getobs' first four entries: x, xdot, thetadot, theta of the PREVIOUS post-step state
    st = gold("o4_state.f64").reshape(-1, 4)
    np.testing.assert_array_equal(obs[1:100], st[0:99][:, [0, 1, 3, 2]].astype(np.float32))
    mu, a2, lp = np.zeros(B, np.float32), np.zeros(B, np.float32), np.zeros(B, np.float32)
    eps = ((act.astype(np.float64) - gold("o4_mu.f32")) / SIGMA).astype(np.float32)
    L.orc_act(C.byref(D4), p0, obs, B, eps.ctypes.data_as(C.c_void_p), SIGMA, mu, a2, lp)
    np.testing.assert_allclose(mu, gold("o4_mu.f32"), rtol=0, atol=1e-6)
    v = np.zeros(B, np.float32)
    L.orc_value(C.byref(D4), p0, obs, B, v)
    np.testing.assert_allclose(v, gold("o4_value.f32"), rtol=1e-6, atol=1e-6)
    lp2 = np.array([L.orc_logprob(float(a), float(m), SIGMA) for a, m in zip(act, gold("o4_mu.f32"))], np.float32)
    np.testing.assert_allclose(lp2, lp_old, rtol=0, atol=2e-6)
    z = np.zeros_like(p0)
    for K in (1, 5):
        p, m, vv, step, losses, g, R = H.oracle_update(p0, z, z, 0, obs, act, lp_old, rew, fl, K, dims=D4)
        assert relerr(R, gold("o4_returns.f32")) < 1e-6
        assert step == K == int(gold(f"o4_k{K}_adam_step.i64")[0])
        assert relerr(g, gold(f"o4_k{K}_grads.f32")) < 1e-4
        assert relerr(p, gold(f"o4_k{K}_params.f32")) < 1e-6
        assert relerr(m, gold(f"o4_k{K}_adam_m.f32")) < 1e-4 and relerr(vv, gold(f"o4_k{K}_adam_v.f32")) < 1e-4
        np.testing.assert_allclose([l.loss for l in losses], gold(f"o4_k{K}_loss.f64"), rtol=1e-5)
