"""GPU tests of the device-resident rollout path: free-running Philox sampling vs the oracle, sharding
invariance, full-size (BASELINE configs[2]) property checks, and end-to-end learning."""
import ctypes as C

import numpy as np
import pytest

from helpers import DEFAULT_DIMS, GAMMA, SIGMA, gold, oracle, oracle_rollout, relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import reinforcement_learning_cpp_parallel_b200 as p
    return p


def make_engine(pkg, **kw):
    from reinforcement_learning_cpp_parallel_b200 import capi
    return pkg.Engine(capi.default_config(**kw))


def test_noise_stream_matches_oracle(pkg):
    """eps recovered from the rollout, (a - mu)/sigma, equals the oracle's Philox/Box-Muller draw."""
    L = oracle()
    N, T = 257, 5
    E = make_engine(pkg, n_envs=N, horizon=T, seed=0x5EED, env_offset=1000)
    E.set_params(gold("init_params.f32"))
    E.rollout(T)
    out = E.traj_download(T)
    eps = (out["action"].astype(np.float64) - out["mu"]) / SIGMA
    ref = np.array([[L.orc_gauss(0x5EED, 1000 + e, t) for e in range(N)] for t in range(T)])
    np.testing.assert_allclose(eps, ref, rtol=0, atol=3e-6)
    # a second rollout continues the per-env step counter
    E.rollout(T)
    out2 = E.traj_download(T)
    eps2 = (out2["action"].astype(np.float64) - out2["mu"]) / SIGMA
    ref2 = np.array([[L.orc_gauss(0x5EED, 1000 + e, T + t) for e in range(N)] for t in range(T)])
    np.testing.assert_allclose(eps2, ref2, rtol=0, atol=3e-6)
    E.close()


def test_noise_statistics(pkg):
    N, T = 1 << 16, 16
    E = make_engine(pkg, n_envs=N, horizon=T, precision=pkg.PREC_F32)
    E.rollout(T)
    out = E.traj_download(T, fields=("action", "mu"))
    eps = ((out["action"].astype(np.float64) - out["mu"]) / SIGMA).ravel()
    n = eps.size
    assert abs(eps.mean()) < 5 / np.sqrt(n) and abs(eps.var() - 1) < 5 * np.sqrt(2 / n)
    assert abs((eps ** 3).mean()) < 0.02 and abs((eps ** 4).mean() - 3) < 0.05
    from scipy import stats
    assert stats.kstest(eps[:: 7][:200000], "norm").statistic < 0.005
    # neighbouring envs / steps are uncorrelated
    e2 = eps.reshape(T, N)
    assert abs(np.corrcoef(e2[:, :-1].ravel(), e2[:, 1:].ravel())[0, 1]) < 0.005
    assert abs(np.corrcoef(e2[:-1].ravel(), e2[1:].ravel())[0, 1]) < 0.005
    E.close()


def test_free_running_rollout_tracks_oracle(pkg):
    """Same weights, same Philox key: GPU fp64 rollout vs oracle rollout.  Divergence (actions differ in
    the last f32 bit, the pendulum amplifies) stays far below 1e-5 over a 100-step episode."""
    N, T = 96, 100
    params = gold("c1_k80_params.f32")
    E = make_engine(pkg, n_envs=N, horizon=T, seed=99, env_offset=5)
    E.set_params(params)
    E.set_state_trace(True)
    E.rollout(T)
    out = E.traj_download(T, state_trace=True)
    ref = oracle_rollout(params, N, T, seed=99, env0=5)
    scale = np.maximum(np.abs(ref["state_trace"]).max(axis=-1, keepdims=True), 1e-2)
    err = np.abs(out["state_trace"] - ref["state_trace"]) / scale
    print("free-running: max rel state err %.3e" % err.max())
    assert err.max() < 1e-3 and np.median(err.max(axis=(1, 2))) < 1e-5
    assert (out["reward"] == ref["reward"]).mean() > 0.999
    np.testing.assert_array_equal(out["flags"], ref["flags"])
    E.close()


def test_sharding_invariance(pkg):
    """Philox is keyed by the GLOBAL env id: 1 shard of 512 envs == 4 shards of 128, bit for bit."""
    T = 40
    params = gold("c1_k5_params.f32")
    full = make_engine(pkg, n_envs=512, horizon=T, seed=3)
    full.set_params(params)
    full.rollout(T)
    A = full.traj_download(T)
    for r in range(4):
        sh = make_engine(pkg, n_envs=128, horizon=T, seed=3, env_offset=128 * r)
        sh.set_params(params)
        sh.rollout(T)
        Bq = sh.traj_download(T)
        for k in ("obs", "action", "logp", "reward", "flags"):
            np.testing.assert_array_equal(Bq[k], A[k][:, 128 * r: 128 * (r + 1)])
        sh.close()
    full.close()


def test_horizon_truncation_and_state_carry(pkg):
    """T = 128 with 100-step episodes (BASELINE configs[2] shape): TIMEUP at step 100, TRUNC at the horizon,
    state carried into the next rollout; two 64-step calls == one 128-step call."""
    N = 64
    params = gold("init_params.f32")
    a = make_engine(pkg, n_envs=N, horizon=128, seed=11)
    a.set_params(params)
    a.rollout(128)
    A = a.traj_download(128)
    fl = A["flags"]
    assert (fl[99] == pkg.FLAG_TIMEUP).all() and (fl[127] == pkg.FLAG_TRUNC).all()
    assert (np.delete(fl, [99, 127], axis=0) == 0).all()
    assert (A["obs"][100] == np.array([0, 0, 0, 0, 1, 0], np.float32)).all()  # reset state (cart_pole.h:86)
    u, es = a.env_get_state()
    assert (es == 28).all()
    b = make_engine(pkg, n_envs=N, horizon=128, seed=11)
    b.set_params(params)
    b.rollout(64)
    B1 = b.traj_download(64)
    b.rollout(64)
    B2 = b.traj_download(64)
    for k in ("obs", "action", "logp", "reward"):
        np.testing.assert_array_equal(np.concatenate([B1[k], B2[k]]), A[k])
    u2, es2 = b.env_get_state()
    np.testing.assert_array_equal(u, u2)
    a.close(); b.close()


def test_full_size_properties(pkg):
    """BASELINE configs[2]: 65,536 envs x 128 steps.  Too big for the oracle, so check size-independent
    properties: determinism, flag pattern, reward domain, log-prob identity, the returns recurrence, and a
    sampled slice against the oracle."""
    N, T = 65536, 128
    params = gold("c1_k80_params.f32")
    E = make_engine(pkg, n_envs=N, horizon=T, seed=0x5EED, k_epochs=1)
    E.set_params(params)
    E.rollout(T)
    A = E.traj_download(T)
    assert set(np.unique(A["reward"])) <= {0.0, 1.0}
    assert (A["flags"][99] == 2).all() and (A["flags"][127] == 4).all() and A["flags"].astype(np.int64).sum() == 6 * N
    assert np.isfinite(A["obs"]).all() and np.isfinite(A["action"]).all()
    lp = -0.5 * ((A["action"].astype(np.float64) - A["mu"]) ** 2) / SIGMA ** 2 - 0.5 * (np.log(SIGMA ** 2) + np.log(2 * np.pi))
    np.testing.assert_allclose(A["logp"], lp, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(A["obs"][..., 4] ** 2 + A["obs"][..., 5] ** 2, 1.0, atol=1e-6)
    # oracle on a strided sample of envs (same global ids)
    ids = np.arange(0, N, 4099)
    for e in ids[:8]:
        ref = oracle_rollout(params, 1, T, seed=0x5EED, env0=int(e))
        err = np.abs(A["obs"][:, e] - ref["obs"][:, 0]).max()
        assert err < 1e-3, (e, err)
        assert (A["reward"][:, e] == ref["reward"][:, 0]).mean() > 0.98
    # determinism
    E2 = make_engine(pkg, n_envs=N, horizon=T, seed=0x5EED, k_epochs=1)
    E2.set_params(params)
    E2.rollout(T)
    A2 = E2.traj_download(T, fields=("action", "reward"))
    np.testing.assert_array_equal(A2["action"], A["action"])
    np.testing.assert_array_equal(A2["reward"], A["reward"])
    # returns recurrence on the full buffer (Q5): R_t = r_t + gamma * (flag ? V(obs_t) : R_{t+1})
    st = E.update_resident()
    assert np.isfinite(st.loss) and st.samples == N * T
    R = E.traj_download(T, fields=("returns",))["returns"]
    g = np.float32(GAMMA)
    nxt = np.empty_like(R)
    nxt[:-1] = R[1:]
    flagged = A["flags"] != 0
    plain = ~flagged
    plain[-1] = False
    np.testing.assert_array_equal(R[plain], (A["reward"] + g * nxt)[plain])
    sel = np.argwhere(flagged)[:: 997]
    obs_sel = A["obs"][sel[:, 0], sel[:, 1]]
    E3 = make_engine(pkg)
    E3.set_params(params)
    V = E3.value(obs_sel)
    np.testing.assert_allclose(R[sel[:, 0], sel[:, 1]], A["reward"][sel[:, 0], sel[:, 1]] + g * V, rtol=1e-6, atol=1e-6)
    E.close(); E2.close(); E3.close()


def test_time_sliced_handoff_is_bit_exact(pkg):
    """The persistent rollout kernel cuts (group, step) units along the TIME axis once there are >= 592 env groups
    (csrc/rollout.cu: a group's steps are then shared by two warps through a release/acquire hand-off in global memory).
    Small engines (one group per warp, no hand-off) over the same GLOBAL env ids must give the same trajectory bit for
    bit -- sharding invariance already promises it -- at the head, the middle and the tail of the full-size grid."""
    N, T = 65536, 128
    params = gold("c1_k80_params.f32")
    full = make_engine(pkg, n_envs=N, horizon=T, seed=0x5EED)
    full.set_params(params)
    full.rollout(T)
    A = full.traj_download(T)
    uA, esA = full.env_get_state()
    cntA, retA = full.episode_stats(2)
    full.close()
    for off in (0, 20000, 65024):
        sh = make_engine(pkg, n_envs=512, horizon=T, seed=0x5EED, env_offset=off)
        sh.set_params(params)
        sh.rollout(T)
        Bq = sh.traj_download(T)
        for k in ("obs", "action", "logp", "mu", "reward", "flags"):
            np.testing.assert_array_equal(Bq[k], A[k][:, off: off + 512], err_msg=f"{k} @ env_offset {off}")
        u, es = sh.env_get_state()
        np.testing.assert_array_equal(u, uA[off: off + 512])
        np.testing.assert_array_equal(es, esA[off: off + 512])
        cnt, ret = sh.episode_stats(2)
        np.testing.assert_array_equal(cnt, cntA[off: off + 512])
        np.testing.assert_array_equal(ret, retA[off: off + 512])
        sh.close()


def test_time_sliced_handoff_forced_action_replay(pkg):
    """32,768 envs (1,024 groups > 592 sub-partitions: time-sliced grid) replaying the oracle's actions on 256 sampled
    envs, zeros elsewhere: fp64 states of the sampled envs against orc_rollout at 1e-9, rewards and flags exact."""
    N, T = 32768, 100
    params = gold("init_params.f32")
    ids = (np.arange(256) * 127 + 5) % N
    assert np.unique(ids).size == 256
    ref = oracle_rollout(params, 256, T, seed=21)
    forced = np.zeros((T, N), np.float32)
    forced[:, ids] = ref["action"]
    E = make_engine(pkg, n_envs=N, horizon=T, seed=21)
    E.set_params(params)
    E.set_state_trace(True)
    E.rollout(T, forced_action=forced)
    out = E.traj_download(T, fields=("reward", "flags", "obs"), state_trace=True)
    got = out["state_trace"][:, ids]
    scale = np.maximum(np.abs(ref["state_trace"]).max(axis=-1, keepdims=True), 1e-300)
    err = (np.abs(got - ref["state_trace"]) / scale).max()
    assert err < 1e-9, err
    np.testing.assert_array_equal(out["reward"][:, ids], ref["reward"])
    np.testing.assert_array_equal(out["flags"][:, ids], ref["flags"])
    np.testing.assert_allclose(out["obs"][:, ids], ref["obs"], rtol=1e-6, atol=1e-7)
    # the envs that were fed a = 0 all saw the same force: identical trajectories, whichever warp advanced them
    rest = np.setdiff1d(np.arange(N), ids)
    assert (out["state_trace"][:, rest] == out["state_trace"][:, rest[:1]]).all()
    E.close()


def test_training_learns_cartpole(pkg):
    """T10: reference default shape (2000 samples per update, K=80, 100-step episodes) reaches the reference's
    plateau (trailing mean episode reward >= 95 of 100, cf. results/carpole1node) within ~60 updates."""
    N, T = 20, 100
    E = make_engine(pkg, n_envs=N, horizon=T, seed=2024, k_epochs=80)
    hist = []
    for it in range(80):
        E.train_iteration(T)
        cnt, ret = E.episode_stats(2)
        assert (cnt == 1).all()
        hist.append(ret[:, 0].mean())
        if len(hist) >= 5 and np.mean(hist[-5:]) >= 95:
            break
    print("episode reward per update:", np.round(hist, 1))
    assert np.mean(hist[-5:]) >= 95, hist
    E.close()


@pytest.mark.parametrize("pinned,direct", [(True, False), (True, True), (False, False)])
def test_rollout_host_pipelined_equals_resident_rollout(pkg, monkeypatch, pinned, direct):
    """b2_rollout_host delivers exactly what b2_rollout + b2_traj_download deliver and carries env state on.  Default:
    chunked launches + the copy-engine pipeline (doubling head, equal chunks, halving tail), for chunk counts that do and
    do not divide the horizon, pinned or pageable buffers.  B2_ROLLOUT_HOST_DIRECT=1 with pinned buffers: the kernel
    stores straight into host memory (one launch; 200 envs = 6 full warps + a ragged one, so both observation store
    paths run)."""
    from reinforcement_learning_cpp_parallel_b200 import capi
    if direct:
        monkeypatch.setenv("B2_ROLLOUT_HOST_DIRECT", "1")
    else:
        monkeypatch.delenv("B2_ROLLOUT_HOST_DIRECT", raising=False)
    alloc = capi.pinned_empty if pinned else (lambda shape, dt: np.empty(shape, dt))
    N, T = 200, 37
    ref = make_engine(pkg, n_envs=N, horizon=T, seed=5)
    ref.init_params(3)
    ref.rollout(T)
    a = ref.traj_download(T)
    ref.rollout(T)
    a2 = ref.traj_download(T)
    for chunks in (1, 2, 4, 37):
        E = make_engine(pkg, n_envs=N, horizon=T, seed=5)
        E.init_params(3)
        out = {"obs": alloc((T, N, 6), np.float32), "action": alloc((T, N), np.float32), "logp": alloc((T, N), np.float32),
               "mu": alloc((T, N), np.float32), "reward": alloc((T, N), np.float32), "flags": alloc((T, N), np.uint8)}
        E.rollout_host(T, out, n_chunks=chunks)
        for k in out:
            np.testing.assert_array_equal(out[k], a[k], err_msg=f"{k} chunks={chunks}")
        E.rollout_host(T, out, n_chunks=chunks)
        for k in out:
            np.testing.assert_array_equal(out[k], a2[k], err_msg=f"{k} second rollout chunks={chunks}")
        np.testing.assert_array_equal(E.env_get_state()[0], ref.env_get_state()[0])
        E.close()
    ref.close()


def test_rollout_host_four_wide_observations(pkg):
    """obs_vars = 4 takes the copy-engine pipeline (the stored cos/sin planes are zeroed before the transpose)."""
    from reinforcement_learning_cpp_parallel_b200 import capi
    N, T = 96, 20
    ref = make_engine(pkg, n_envs=N, horizon=T, seed=8, obs_vars=4)
    ref.init_params(2)
    ref.rollout(T)
    a = ref.traj_download(T)
    ref.close()
    E = make_engine(pkg, n_envs=N, horizon=T, seed=8, obs_vars=4)
    E.init_params(2)
    out = {"obs": capi.pinned_empty((T, N, 4), np.float32), "action": capi.pinned_empty((T, N), np.float32),
           "flags": capi.pinned_empty((T, N), np.uint8)}
    E.rollout_host(T, out, n_chunks=4)
    for k in out:
        np.testing.assert_array_equal(out[k], a[k], err_msg=k)
    E.close()


def test_rollout_host_compact_record(pkg):
    """b2_rollout_host_compact (25 B per env-step): obs4 = the first four observation components, the reward rides in
    the flag byte; everything bit-identical to the resident rollout of a twin engine (same seed), ragged env count."""
    from reinforcement_learning_cpp_parallel_b200 import capi
    N, T = 1000 + 7, 40
    A = make_engine(pkg, n_envs=N, horizon=T, seed=11)
    B = make_engine(pkg, n_envs=N, horizon=T, seed=11)
    p = gold("init_params.f32")
    A.set_params(p); B.set_params(p)
    A.rollout(T)
    ref = A.traj_download(T)
    out = {"obs4": capi.pinned_empty((T, N, 4), np.float32), "action": capi.pinned_empty((T, N), np.float32),
           "logp": capi.pinned_empty((T, N), np.float32), "rflags": capi.pinned_empty((T, N), np.uint8)}
    for k in out:
        out[k][...] = 0x55 if out[k].dtype == np.uint8 else np.nan
    B.rollout_host_compact(T, out, n_chunks=5)
    assert np.array_equal(out["obs4"], ref["obs"][..., :4])
    assert np.array_equal(out["action"], ref["action"]) and np.array_equal(out["logp"], ref["logp"])
    assert np.array_equal(out["rflags"] & 7, ref["flags"])
    assert np.array_equal((out["rflags"] >> 3) & 1, (ref["reward"] != 0).astype(np.uint8))
    # the device-resident trajectory is the same one: an update on it matches the twin's
    sa, sb = A.update_resident(), B.update_resident()
    assert sa.loss == sb.loss and np.array_equal(A.get_params(), B.get_params())
    A.close(); B.close()
