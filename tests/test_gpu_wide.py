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


# ------------------------------------------------------------------------------------------------------------
# wide nets through the C ABI: action query, value estimate, returns bootstrap, update (BASELINE configs[4])
# ------------------------------------------------------------------------------------------------------------
import ctypes as C

WIDE = [(128, 64), (256, 256)]


def wide_engine(H1, H2, n=256, **kw):
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    return pkg.Engine(capi.default_config(n_envs=n, horizon=1, hidden1=H1, hidden2=H2, **kw))


def wide_case(H1, H2):
    tag = f"wide{H1}x{H2}"
    g = lambda n: H.gold(f"{tag}_{n}")
    return dict(dims=H.Dims(6, H1, H2, 1), params=g("params.f32"), obs=g("obs.f32").reshape(-1, 6), action=g("action.f32"),
                oldlp=g("oldlogp.f32"), R=g("R.f32"), mu=g("mu.f32"), V=g("V.f32"), grads=g("grads.f32"), loss=g("loss.f64"))


def synth_batch(dims, B, seed, params=None):
    """random weights + a batch with ratios around 1 (some clipped), episodes of 50 steps ending in TIMEUP"""
    rng = np.random.default_rng(seed)
    L = H.oracle()
    P = L.orc_n_params(C.byref(dims))
    if params is None:
        params = np.zeros(P, np.float32)
        for name, sl in H.param_slices(dims):
            fan = {"W1": 6, "b1": 6, "W2": dims.H1, "b2": dims.H1, "W3": dims.H2, "b3": dims.H2}[name.split(".")[1]]
            params[sl] = rng.uniform(-1, 1, sl.stop - sl.start).astype(np.float32) / np.sqrt(fan)
    th = 0.3 * rng.standard_normal(B)
    obs = np.stack([rng.standard_normal(B), rng.standard_normal(B), rng.standard_normal(B), th, np.cos(th), np.sin(th)],
                   axis=1).astype(np.float32)
    mu, act, lp = (np.zeros(B, np.float32) for _ in range(3))
    eps = rng.standard_normal(B).astype(np.float32)
    L.orc_act(C.byref(dims), params, obs, B, H.ptr(eps), H.SIGMA, mu, act, lp)
    oldlp = (lp + 0.1 * rng.standard_normal(B)).astype(np.float32)
    reward = (rng.random(B) < 0.9).astype(np.float32)
    flags = np.zeros(B, np.uint8)
    flags[49::50] = 2
    flags[-1] = 2
    return params, obs, act, oldlp, reward, flags


@pytest.mark.parametrize("H1,H2", WIDE)
def test_wide_act_value_match_golden(H1, H2):
    c = wide_case(H1, H2)
    with wide_engine(H1, H2) as E:
        assert E.P == c["params"].size
        E.set_params(c["params"])
        v = E.value(c["obs"])
        assert np.abs(v - c["V"]).max() < 2e-6 * max(1.0, np.abs(c["V"]).max())
        eps = np.linspace(-2, 2, c["R"].size).astype(np.float32)
        a, lp, mu = E.act(c["obs"], eps)
        assert np.abs(mu - c["mu"]).max() < 1e-6
        mo, ao, lo = (np.zeros_like(a) for _ in range(3))
        H.oracle().orc_act(C.byref(c["dims"]), c["params"], c["obs"], a.size, H.ptr(eps), H.SIGMA, mo, ao, lo)
        assert np.abs(a - ao).max() < 1e-6 and np.abs(lp - lo).max() < 2e-5


@pytest.mark.parametrize("H1,H2", WIDE)
def test_wide_gradient_matches_torch_golden(H1, H2):
    """one epoch on the torch-minted batch: loss terms and all 12 gradient tensors within 1e-4 (north_star gate)"""
    c = wide_case(H1, H2)
    B = c["R"].size
    with wide_engine(H1, H2, k_epochs=1) as E:
        E.set_params(c["params"])
        # returns fed directly: every record terminal => R_k = r_k (src/ppo_nn.cpp:229-231)
        st = E.update(c["obs"], c["action"], c["oldlp"], c["R"], np.ones(B, np.uint8))
        g = E.get_grads()
        gl = c["loss"]
        assert abs(st.loss1 - gl[0]) < 1e-4 * abs(gl[0]) + 1e-6 and abs(st.loss2 - gl[1]) < 1e-4 * abs(gl[1])
        assert abs(st.loss - gl[3]) < 1e-4 * abs(gl[3]) and abs(st.clip_frac - gl[4]) < 1e-6
        worst = max((H.relerr(g[sl], c["grads"][sl]), name) for name, sl in H.param_slices(c["dims"]))
        print("wide grad worst tensor:", worst)
        assert worst[0] < 1e-4, worst
        assert st.adam_step == 1


@pytest.mark.parametrize("H1,H2,B,K", [(256, 256, 1000, 4), (128, 64, 333, 3), (192, 128, 2048, 2), (128, 256, 777, 3), (64, 256, 130, 3)])
def test_wide_update_matches_oracle(H1, H2, B, K):
    """K full-batch epochs incl. returns with TIMEUP bootstraps, ragged B: weights, Adam state, per-epoch losses"""
    dims = H.Dims(6, H1, H2, 1)
    params, obs, act, oldlp, reward, flags = synth_batch(dims, B, 5 + B)
    z = np.zeros_like(params)
    po, mo, vo, so, losses, go, Ro = H.oracle_update(params, z, z, 0, obs, act, oldlp, reward, flags, K, dims=dims)
    with wide_engine(H1, H2, k_epochs=K) as E:
        E.set_params(params)
        R = E.returns(obs, reward, flags)
        assert H.relerr(R, Ro) < 1e-5
        st = E.update(obs, act, oldlp, reward, flags)
        l1, l2, l3, lt = E.last_update_losses(K)
        for k in range(K):
            assert abs(lt[k] - losses[k].loss) < 1e-4 * abs(losses[k].loss) + 1e-6, (k, lt[k], losses[k].loss)
        assert H.relerr(E.get_grads(), go) < 1e-4
        p = E.get_params()
        m, v, step = E.get_adam_state()
        assert step == so == K
        assert H.relerr(p, po) < 1e-5 and H.relerr(m, mo) < 1e-4 and H.relerr(v, vo) < 2e-4


def test_wide_update_chunked_and_minibatched():
    """a batch larger than the activation chunk (forces gradient accumulation over chunks) and 3 minibatches per epoch"""
    H1, H2, B, K, M = 256, 256, 1500, 2, 3
    dims = H.Dims(6, H1, H2, 1)
    params, obs, act, oldlp, reward, flags = synth_batch(dims, B, 77)
    L = H.oracle()
    R = np.zeros(B, np.float32)
    L.orc_returns(C.byref(dims), params, obs, reward, flags, B, H.GAMMA, R)
    p, m, v = params.copy(), np.zeros_like(params), np.zeros_like(params)
    per, step = B // M, 0
    for k in range(K):
        for mb in range(M):
            lo = mb * per
            hi = B if mb == M - 1 else lo + per
            idx = np.arange(lo, hi, dtype=np.int64)
            g, out = np.zeros_like(params), H.Loss()
            L.orc_loss_grad(C.byref(dims), p, obs, act, oldlp, R, idx.ctypes.data_as(C.c_void_p), idx.size, H.SIGMA,
                            H.EPS_CLIP, C.byref(out), g)
            step += 1
            L.orc_adam(p, g, m, v, p.size, step, H.LR, H.B1, H.B2, 1e-8)
    # engine created for 256 records => activation chunk of 384 samples; minibatches of 500 need 2 chunks each
    with wide_engine(H1, H2, n=256, k_epochs=K, n_minibatches=M) as E:
        E.set_params(params)
        st = E.update(obs, act, oldlp, reward, flags)
        assert st.optimizer_steps == K * M and st.adam_step == K * M
        assert H.relerr(E.get_grads(), g) < 1e-4
        assert H.relerr(E.get_params(), p) < 1e-5


@pytest.mark.parametrize("merge", ["1", "0"], ids=["merged_l1r", "reduce_own_launch"])
def test_wide_minibatched_updates_carry_state(merge, monkeypatch):
    """3 ragged minibatches x 2 epochs, TWO updates in a row on one engine (the second starts from the first one's weights and
    Adam moments): the default sequence defers every step's reduce + Adam into the next step's layer-1 kernel (k_wide_l1r)
    and flushes after the last one, so the state the next update, get_params and get_adam_state see must be complete;
    B2_WIDE_MERGE=0 is the sequence with k_wide_reduce as its own launch"""
    monkeypatch.setenv("B2_WIDE_MERGE", merge)
    H1, H2, B, K, M = 256, 256, 1000, 2, 3
    dims = H.Dims(6, H1, H2, 1)
    params, obs, act, oldlp, reward, flags = synth_batch(dims, B, 99)
    L = H.oracle()
    p, m, v = params.copy(), np.zeros_like(params), np.zeros_like(params)
    per, step = B // M, 0
    with wide_engine(H1, H2, n=B, k_epochs=K, n_minibatches=M) as E:
        E.set_params(params)
        for upd in range(2):
            R = np.zeros(B, np.float32)
            L.orc_returns(C.byref(dims), p, obs, reward, flags, B, H.GAMMA, R)   # bootstraps use the current critic
            for k in range(K):
                for mb in range(M):
                    lo = mb * per
                    hi = B if mb == M - 1 else lo + per
                    idx = np.arange(lo, hi, dtype=np.int64)
                    g, out = np.zeros_like(params), H.Loss()
                    L.orc_loss_grad(C.byref(dims), p, obs, act, oldlp, R, idx.ctypes.data_as(C.c_void_p), idx.size, H.SIGMA,
                                    H.EPS_CLIP, C.byref(out), g)
                    step += 1
                    L.orc_adam(p, g, m, v, p.size, step, H.LR, H.B1, H.B2, 1e-8)
            st = E.update(obs, act, oldlp, reward, flags)
            assert st.optimizer_steps == K * M and st.adam_step == step
            assert abs(st.loss - out.loss) < 1e-4 * abs(out.loss) + 1e-6
            assert H.relerr(E.get_grads(), g) < 1e-4
            em, ev, es = E.get_adam_state()
            assert es == step
            assert H.relerr(E.get_params(), p) < 1e-5 and H.relerr(em, m) < 1e-4 and H.relerr(ev, v) < 2e-4


@pytest.mark.parametrize("env", [{"B2_WIDE_DETERMINISTIC": "1"}, {"B2_WIDE_FUSE_REDUCE": "1"}, {"B2_WIDE_FUSE_M": "1"},
                                 {"B2_WIDE_FUSED": "0"}, {"B2_WIDE_FUSED": "0", "B2_WIDE_DETERMINISTIC": "1"}, {"B2_WIDE_FUSE_DX": "0"},
                                 {"B2_WIDE_MERGE": "0"}],
                         ids=["deterministic", "fused_reduce_tail", "fuse_m", "unfused_forward_head", "unfused_deterministic", "dx_own_launch",
                              "reduce_own_launch"])
def test_wide_update_opt_in_modes_match_oracle(env, monkeypatch):
    """the opt-in variants of the step (read from the environment when the engine is created / the step is launched) are
    held to the same oracle gates as the default; the fixed-order mode must also be bit-identical run to run"""
    H1, H2, B, K = 256, 256, 1000, 3
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    dims = H.Dims(6, H1, H2, 1)
    params, obs, act, oldlp, reward, flags = synth_batch(dims, B, 4242)
    z = np.zeros_like(params)
    po, mo, vo, so, losses, go, Ro = H.oracle_update(params, z, z, 0, obs, act, oldlp, reward, flags, K, dims=dims)
    runs = []
    for _ in range(2):
        with wide_engine(H1, H2, k_epochs=K) as E:
            E.set_params(params)
            E.update(obs, act, oldlp, reward, flags)
            p = E.get_params()
            m, v, step = E.get_adam_state()
            assert step == so == K
            assert H.relerr(E.get_grads(), go) < 1e-4
            assert H.relerr(p, po) < 1e-5 and H.relerr(m, mo) < 1e-4 and H.relerr(v, vo) < 2e-4
            runs.append(p.copy())
    if "B2_WIDE_DETERMINISTIC" in env:
        assert np.array_equal(runs[0], runs[1])


@pytest.mark.parametrize("H1,H2", WIDE)
def test_wide_rollout_matches_oracle(H1, H2):
    """A wide policy collects its own data on the device (csrc/rollout_wide.cu: observe + actor forward + step per env
    step).  Same env semantics and Philox streams as the fused kernel: replaying the oracle's actions reproduces its fp64
    states to 1e-9, rewards / flags exactly and mu / log-prob to the forward-pass gate; free-running, the noise stream is
    the oracle's draw for draw; TIMEUP / TRUNC / reset / episode bookkeeping as k_rollout."""
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    dims = H.Dims(6, H1, H2, 1)
    rng = np.random.default_rng(H1)
    P = H.oracle().orc_n_params(C.byref(dims))
    params = (rng.standard_normal(P) / np.sqrt(H1)).astype(np.float32)
    N, T = 67, 128
    ref = H.oracle_rollout(params, N, T, seed=31, env0=9, dims=dims)
    E = pkg.Engine(capi.default_config(n_envs=N, horizon=T, hidden1=H1, hidden2=H2, seed=31, env_offset=9, k_epochs=1))
    E.set_params(params)
    E.set_state_trace(True)
    E.rollout(T, forced_action=ref["action"])
    out = E.traj_download(T, state_trace=True)
    scale = np.maximum(np.abs(ref["state_trace"]).max(axis=-1, keepdims=True), 1e-300)
    assert (np.abs(out["state_trace"] - ref["state_trace"]) / scale).max() < 1e-9
    np.testing.assert_array_equal(out["reward"], ref["reward"])
    np.testing.assert_array_equal(out["flags"], ref["flags"])
    assert (out["flags"][99] == 2).all() and (out["flags"][127] == 4).all()
    np.testing.assert_allclose(out["obs"], ref["obs"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out["mu"], ref["mu"], rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(out["logp"], ref["logp"], rtol=1e-5, atol=1e-5)
    cnt, ret = E.episode_stats(2)
    assert (cnt == 1).all()
    np.testing.assert_array_equal(ret[:, 0], ref["reward"][:100].sum(axis=0))
    # free-running: eps recovered from (a - mu) / sigma is the oracle's Philox draw (counter continues at step T)
    E.rollout(8)
    fr = E.traj_download(8, fields=("action", "mu"))
    eps = (fr["action"].astype(np.float64) - fr["mu"]) / H.SIGMA
    want = np.array([[H.oracle().orc_gauss(31, 9 + e, T + t) for e in range(N)] for t in range(8)])
    np.testing.assert_allclose(eps, want, rtol=0, atol=3e-6)
    # and the wide learner trains on what the wide actor collected
    st = E.update_resident()
    assert np.isfinite(st.loss) and st.samples == N * 8
    E.close()


def test_wide_rollout_host_buffers():
    """b2_rollout_host on a wide engine takes the chunked copy-engine pipeline and equals rollout + download."""
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    N, T = 40, 12
    mk = lambda: pkg.Engine(capi.default_config(n_envs=N, horizon=T, hidden1=128, hidden2=64, seed=4))
    A, B = mk(), mk()
    A.rollout(T)
    a = A.traj_download(T)
    out = {"obs": capi.pinned_empty((T, N, 6), np.float32), "action": capi.pinned_empty((T, N), np.float32),
           "logp": capi.pinned_empty((T, N), np.float32), "mu": capi.pinned_empty((T, N), np.float32),
           "reward": capi.pinned_empty((T, N), np.float32), "flags": capi.pinned_empty((T, N), np.uint8)}
    B.rollout_host(T, out, n_chunks=3)
    for k in out:
        np.testing.assert_array_equal(out[k], a[k], err_msg=k)
    A.close(); B.close()


@pytest.mark.parametrize("M,N,K,splits", [(128, 256, 256, 1), (256, 256, 1000, 3), (64, 128, 96, 1), (192, 64, 4096, 5)])
def test_gemm_bf16x3_mn_major(eng, M, N, K, splits):
    """contraction-major operands (the dW2 = D2^T H1 layout): C = A^T B with A [K][M], B [K][N]"""
    rng = np.random.default_rng(M + N + K)
    A = rng.standard_normal((K, M)).astype(np.float32)
    B = (rng.standard_normal((K, N)) / np.sqrt(K)).astype(np.float32)
    Cm = eng.selftest_gemm(A, B, splits, mn_major=True)
    ref = A.astype(np.float64).T @ B.astype(np.float64)
    rel = np.linalg.norm(Cm - ref) / np.linalg.norm(ref)
    print(f"gemm MN {M}x{N}x{K} splits {splits}: rel fro {rel:.2e}")
    assert rel < 2e-5, rel


@pytest.mark.parametrize("M,N,K,splits", [(128, 256, 256, 1), (300, 128, 192, 1), (1000, 256, 64, 1), (64, 64, 640, 3)])
def test_gemm_bf16x3_mixed_major(eng, M, N, K, splits):
    """A row-major [M][K], B contraction-major [K][N] (Z1 = D2 W2 reads W2 as stored): C = A B"""
    rng = np.random.default_rng(3 * M + N + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = (rng.standard_normal((K, N)) / np.sqrt(K)).astype(np.float32)
    Cm = eng.selftest_gemm(A, B, splits, mn_major=2)
    ref = A.astype(np.float64) @ B.astype(np.float64)
    rel = np.linalg.norm(Cm - ref) / np.linalg.norm(ref)
    print(f"gemm mixed {M}x{N}x{K} splits {splits}: rel fro {rel:.2e}")
    assert rel < 2e-5, rel


@pytest.mark.parametrize("M,N,K,mode", [(256, 256, 256, 3), (1000, 256, 256, 3), (40000, 256, 256, 3), (512, 128, 192, 3), (300, 64, 64, 3),
                                        (256, 256, 256, 4), (5000, 256, 128, 4), (777, 128, 256, 4)])
def test_gemm_cta_pair_weights_resident(eng, M, N, K, mode):
    """k_gemm2_ws (cta_group::2, resident weight halves): mode 3 = weights K-major [N][K], mode 4 = weights as [K][N]"""
    rng = np.random.default_rng(M + 5 * N + K + mode)
    A = rng.standard_normal((M, K)).astype(np.float32)
    if mode == 3:
        B = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
        ref = A.astype(np.float64) @ B.astype(np.float64).T
    else:
        B = (rng.standard_normal((K, N)) / np.sqrt(K)).astype(np.float32)
        ref = A.astype(np.float64) @ B.astype(np.float64)
    Cm = eng.selftest_gemm(A, B, 1, mn_major=mode)
    rel = np.linalg.norm(Cm - ref) / np.linalg.norm(ref)
    print(f"gemm2 mode {mode} {M}x{N}x{K}: rel fro {rel:.2e}")
    assert rel < 2e-5, rel
