"""GPU parity of the opt-in environment variants (b200ppo.h B2_ENV_*): the SWINGUP branches of
include/cart_pole.h (:80-81,93-94,100-101,136-139,154-159), early termination inside advance() (:115, live in
archived/single_environment/cart_pole.h:119) and the random initial state (:85 / archived :87).

Two checkers: the known-answer vectors minted from the reference's own headers (tests/golden/envvar_*, made by
oracle/ref_env_variants.cpp) through the cart_pole.h-shaped entry points, and the C oracle's orc_rollout_v for the
fused rollout.  Tolerances: states 1e-9 relative (gate 1e-5); rewards, flags, terminations bit-exact."""
import ctypes as C

import numpy as np
import pytest

import helpers as H
from helpers import gold, oracle, oracle_rollout_v, relerr

pytestmark = pytest.mark.gpu

SW, RI, ET = H.ENV_SWINGUP, H.ENV_RANDOM_INIT, H.ENV_EARLY_TERM


@pytest.fixture(scope="module")
def pkg():
    import reinforcement_learning_cpp_parallel_b200 as p
    return p


def make_engine(pkg, **kw):
    from reinforcement_learning_cpp_parallel_b200 import capi
    return pkg.Engine(capi.default_config(**kw))


@pytest.mark.parametrize("name,fl", [("swingup", SW), ("early", ET), ("early_swingup", ET | SW)])
def test_env_step_against_reference_kat(pkg, name, fl):
    """b2_env_step / b2_env_observe / b2_env_terminated replay the KAT minted from the reference header."""
    kat = gold(f"envvar_{name}.f64").reshape(8, 620, 8)
    init = gold(f"envvar_{name}_init.f64").reshape(8, 64, 4)
    NE, T = kat.shape[0], kat.shape[1]
    E = make_engine(pkg, n_envs=NE, horizon=4, env_flags=fl)
    u = init[:, 0].copy()
    slot = np.zeros(NE, int)
    steps = np.zeros(NE, np.int32)
    E.env_set_state(u, steps)
    worst, n_term = 0.0, 0
    for t in range(T):
        E.env_step(kat[:, t, 0])
        obs, rew = E.env_observe()
        term = E.env_terminated()
        u, steps = E.env_get_state()
        ref_u = kat[:, t, 1:5]
        scale = np.maximum(np.abs(ref_u).max(axis=1, keepdims=True), 1e-3)
        worst = max(worst, float((np.abs(u - ref_u) / scale).max()))
        np.testing.assert_array_equal(term, kat[:, t, 6].astype(np.uint8), err_msg=f"{name} t={t}")
        np.testing.assert_array_equal(rew, kat[:, t, 5], err_msg=f"{name} t={t}")
        np.testing.assert_array_equal(steps, kat[:, t, 7].astype(np.int32))
        np.testing.assert_allclose(obs[:, [0, 1, 3, 2]], u, rtol=0, atol=0)       # getobs order (:123-126)
        if term.any():                                                            # the KAT's scripted re-initialisation
            for e in np.nonzero(term)[0]:
                n_term += 1
                slot[e] = (slot[e] + 1) % 64
                u[e] = init[e, slot[e]]
                steps[e] = 0
        # keep the engine on the reference's trajectory: drift must not flip a later termination
        E.env_set_state(np.where(term[:, None] != 0, u, ref_u), steps)
    print(f"{name}: worst per-step rel state err {worst:.2e}, {n_term} terminations")
    assert worst < 1e-9
    assert (n_term > 10) == bool(fl & ET)
    E.close()


@pytest.mark.parametrize("fl", [SW, ET, ET | RI, SW | ET | RI, RI])
def test_variant_rollout_matches_oracle(pkg, fl):
    """fused rollout under every variant, actions forced to the oracle's: states, rewards, flags, obs, mu, logp."""
    N, T, nts = 160, 130, 60
    params = gold("init_params.f32")
    # a noisy policy makes poles fall: many terminations per env with EARLY_TERM
    ref = oracle_rollout_v(params, N, T, fl, n_timestep=nts, seed=11, env0=40, sigma=H.SIGMA,
                           state=_init_states(fl, 11, 40, N))
    E = make_engine(pkg, n_envs=N, horizon=T, n_timestep=nts, seed=11, env_offset=40, env_flags=fl)
    E.set_params(params)
    u0, _ = E.env_get_state()
    np.testing.assert_array_equal(u0, _init_states(fl, 11, 40, N))              # b2_create draws tick 0
    E.set_state_trace(True)
    E.rollout(T, forced_action=ref["action"])
    out = E.traj_download(T, state_trace=True)
    scale = np.maximum(np.abs(ref["state_trace"]).max(axis=-1, keepdims=True), 1e-3)
    err = float((np.abs(out["state_trace"] - ref["state_trace"]) / scale).max())
    n_term = int((ref["flags"] & 1).sum())
    print(f"flags {fl}: max rel state err {err:.2e}; {n_term} TERMINAL, {int((ref['flags'] & 2 != 0).sum())} TIMEUP")
    assert err < 1e-9
    np.testing.assert_array_equal(out["flags"], ref["flags"])
    np.testing.assert_array_equal(out["reward"], ref["reward"])
    np.testing.assert_allclose(out["obs"], ref["obs"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out["mu"], ref["mu"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(out["logp"], ref["logp"], rtol=0, atol=2e-5)
    if fl & ET:
        assert n_term > N // 4
    else:
        assert n_term == 0
    # final env state and episode clocks
    u, steps = E.env_get_state()
    np.testing.assert_array_equal(steps, ref["ep_step"])
    assert relerr(u, ref["state"]) < 1e-9
    # episode log: completed episodes per env = TERMINAL + TIMEUP records
    cnt, _ = E.episode_stats(T + 1)
    np.testing.assert_array_equal(cnt, ((ref["flags"] & 3) != 0).sum(axis=0))
    E.close()


def _init_states(fl, seed, env0, N, tick=0):
    L = oracle()
    u = np.zeros((N, 4))
    for e in range(N):
        L.orc_cartpole_init_state(fl, seed, env0 + e, tick, u[e])
    return u


def test_random_init_reset_and_sharding(pkg):
    """b2_env_reset redraws (tick 2^63 | call); a shard with env_offset reproduces the same global envs."""
    fl = RI | ET
    E = make_engine(pkg, n_envs=64, horizon=8, seed=5, env_flags=fl)
    E.env_reset()
    u1, _ = E.env_get_state()
    np.testing.assert_array_equal(u1, _init_states(fl, 5, 0, 64, tick=(1 << 63) | 1))
    E.env_reset()
    u2, _ = E.env_get_state()
    np.testing.assert_array_equal(u2, _init_states(fl, 5, 0, 64, tick=(1 << 63) | 2))
    assert np.abs(u1).max() < 0.05 and not np.array_equal(u1, u2)
    E.close()
    # free-running (sampled) rollouts: two half shards == one engine, record for record
    params = gold("init_params.f32")
    full = make_engine(pkg, n_envs=64, horizon=40, n_timestep=25, seed=5, env_flags=fl)
    full.set_params(params)
    full.rollout(40)
    a = full.traj_download(40)
    full.close()
    for half in (0, 1):
        S = make_engine(pkg, n_envs=32, horizon=40, n_timestep=25, seed=5, env_offset=32 * half, env_flags=fl)
        S.set_params(params)
        S.rollout(40)
        b = S.traj_download(40)
        for k in ("obs", "action", "logp", "reward", "flags"):
            np.testing.assert_array_equal(b[k], a[k][:, 32 * half:32 * half + 32], err_msg=k)
        S.close()
    assert (a["flags"] & 1).sum() > 0


def test_variant_update_restarts_returns_at_terminal(pkg):
    """PPO::update on a trajectory with TERMINATE records (src/ppo_nn.cpp:229-230): returns + one epoch vs the oracle."""
    fl = ET | RI
    N, T = 48, 90
    params = gold("init_params.f32")
    E = make_engine(pkg, n_envs=N, horizon=T, n_timestep=50, seed=21, env_flags=fl, k_epochs=1)
    E.set_params(params)
    E.rollout(T)
    tr = E.traj_download(T)
    assert (tr["flags"] & 1).sum() > 10
    E.update_resident()
    R = E.traj_download(T, fields=("returns",))["returns"]
    L = oracle()
    Rref = np.empty_like(R)
    for e in range(N):
        r = np.zeros(T, np.float32)
        L.orc_returns(C.byref(H.DEFAULT_DIMS), params, np.ascontiguousarray(tr["obs"][:, e]),
                      np.ascontiguousarray(tr["reward"][:, e]), np.ascontiguousarray(tr["flags"][:, e]), T, H.GAMMA, r)
        Rref[:, e] = r
    assert relerr(R, Rref) < 1e-5
    # a TERMINAL record's return is its own reward (the discounted sum restarts)
    term = (tr["flags"] & 1) != 0
    np.testing.assert_array_equal(R[term], tr["reward"][term])
    flat = lambda a, *t: np.ascontiguousarray(a.reshape(N * T, *t))
    g = np.zeros_like(params)
    Ls = H.Loss()
    L.orc_loss_grad(C.byref(H.DEFAULT_DIMS), params, flat(tr["obs"], 6), flat(tr["action"]), flat(tr["logp"]),
                    flat(Rref), None, N * T, H.SIGMA, H.EPS_CLIP, C.byref(Ls), g)
    assert relerr(E.get_grads(), g) < 1e-4
    E.close()


def test_variants_refuse_f32_physics(pkg):
    from reinforcement_learning_cpp_parallel_b200 import capi
    with pytest.raises(capi.B2Error) as ei:
        make_engine(pkg, n_envs=32, horizon=4, env_flags=SW, precision=pkg.PREC_F32)
    assert ei.value.code == -1


# ---- obs_vars = 4: PPO(state_dim = 4) + getobs(4) (include/cart_pole.h:120-127).  Goldens o4_* come from the unmodified
# reference built with state_dim 4 (oracle/ref_harness.cpp golden_obs4). ----
def test_obs4_act_value_update_match_reference(pkg):
    p0 = gold("o4_init_params.f32")
    obs = gold("o4_obs.f32").reshape(-1, 4)
    B = obs.shape[0]
    act, lp_old, rew, fl = gold("o4_action.f32"), gold("o4_logp.f32"), gold("o4_reward.f32"), H.flags_of("o4")
    for K in (1, 5):
        E = make_engine(pkg, obs_vars=4, n_envs=1, horizon=B, k_epochs=K)
        assert E.P == 4866 == p0.size
        E.set_params(p0)
        np.testing.assert_array_equal(E.get_params(), p0)
        eps = ((act.astype(np.float64) - gold("o4_mu.f32")) / H.SIGMA).astype(np.float32)
        a, lp, mu = E.act(obs, eps=eps)
        np.testing.assert_allclose(mu, gold("o4_mu.f32"), rtol=0, atol=1e-6)
        np.testing.assert_allclose(E.value(obs), gold("o4_value.f32"), rtol=1e-6, atol=1e-6)
        R = E.returns(obs, rew, fl, n_envs=1, horizon=B)
        assert relerr(R, gold("o4_returns.f32")) < 1e-5
        st = E.update(obs, act, lp_old, rew, fl, n_envs=1, horizon=B)
        assert st.adam_step == K
        assert relerr(E.get_grads(), gold(f"o4_k{K}_grads.f32")) < 1e-4
        assert relerr(E.get_params(), gold(f"o4_k{K}_params.f32")) < 1e-5
        m, v, step = E.get_adam_state()
        assert relerr(m, gold(f"o4_k{K}_adam_m.f32")) < 1e-4 and relerr(v, gold(f"o4_k{K}_adam_v.f32")) < 1e-4
        l = E.last_update_losses(K)
        np.testing.assert_allclose(l[3], gold(f"o4_k{K}_loss.f64"), rtol=1e-4)
        E.close()


def test_obs4_rollout_replays_reference_episode(pkg):
    """fused rollout with 4-wide observations: replay the reference's recorded actions from the zero state; the stored
    observations are getobs(4) of the reference, mu/logp those of its state_dim-4 actor."""
    p0 = gold("o4_init_params.f32")
    T = 100
    obs = gold("o4_obs.f32").reshape(-1, 4)[:T]
    E = make_engine(pkg, obs_vars=4, n_envs=1, horizon=T, n_timestep=100)
    E.set_params(p0)
    E.set_state_trace(True)
    E.rollout(T, forced_action=gold("o4_action.f32")[:T])
    out = E.traj_download(T, state_trace=True)
    assert out["obs"].shape == (T, 1, 4)
    st = gold("o4_state.f64").reshape(-1, 4)[:T]
    assert relerr(out["state_trace"][:, 0], st) < 1e-9
    np.testing.assert_allclose(out["obs"][:, 0], obs, rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out["mu"][:, 0], gold("o4_mu.f32")[:T], rtol=0, atol=2e-6)
    np.testing.assert_allclose(out["logp"][:, 0], gold("o4_logp.f32")[:T], rtol=0, atol=2e-5)
    np.testing.assert_array_equal(out["reward"][:, 0], gold("o4_reward.f32")[:T])
    np.testing.assert_array_equal(out["flags"][:, 0], H.flags_of("o4")[:T])
    # resident update on that trajectory == host update on the same records (the padded columns stay exactly zero)
    E.update_resident()
    g_res = E.get_grads()
    E2 = make_engine(pkg, obs_vars=4, n_envs=1, horizon=T)
    E2.set_params(p0)
    E2.update(out["obs"].reshape(T, 4), out["action"].ravel(), out["logp"].ravel(), out["reward"].ravel(),
              out["flags"].ravel(), n_envs=1, horizon=T)
    assert relerr(g_res, E2.get_grads()) < 1e-6
    E.close()
    E2.close()
