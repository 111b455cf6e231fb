"""The Python binding shaped like the reference's pybind experiment (try_new_feature/pybind/ppo_fish_pybind.cpp:11-25):
MemoryNN host semantics on CPU; PPO.select_action / PPO.update against the reference goldens on the GPU."""
import numpy as np
import pytest

import helpers as H
from helpers import gold, relerr


@pytest.fixture(scope="module")
def binding():
    import __graft_entry__ as g
    g.build(oracle=False)
    from reinforcement_learning_cpp_parallel_b200 import ppo_binding
    return ppo_binding


def test_memory_semantics(binding):
    a, b = binding.MemoryNN(), binding.MemoryNN()
    for m, base in ((a, 0), (b, 10)):
        for i in range(3):
            m.states.append(np.full(6, base + i, np.float32))
            m.actions.append(base + i)
            m.logprobs.append(-(base + i))
            m.push_reward(base + i, i == 1, i == 2)
    a.merge(b)                                   # src/ppo_nn.cpp:73-87: rank order, r's records after ours
    assert len(a) == 6 and a.rewards == [0, 1, 2, 10, 11, 12]
    assert a.is_terminals == [False, True, False] * 2 and a.is_timeups == [False, False, True] * 2
    assert len(b) == 3                           # merge copies, it does not drain
    a.clear()
    assert len(a) == 0 and a.rewards == [] and a.states == []


def test_no_cpu_path(binding):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from reinforcement_learning_cpp_parallel_b200 import capi
    with pytest.raises(capi.B2Error) as ei:
        binding.PPO(6, 1, 0.5, 0.0003, (0.9, 0.999), 0.99, 80, 0.2)
    assert ei.value.code == -2


@pytest.mark.gpu
def test_update_matches_reference(binding):
    """PPO.update(MemoryNN) on the reference's recorded 2,000-record batch (config 1) reproduces its weights."""
    ppo = binding.PPO(6, 1, 0.5, 0.0003, (0.9, 0.999), 0.99, 5, 0.2)
    ppo.load_parameters(gold("init_params.f32"))
    mem = binding.MemoryNN()
    obs = gold("c1_obs.f32").reshape(-1, 6)
    for i in range(obs.shape[0]):
        mem.states.append(obs[i])
        mem.actions.append(float(gold("c1_action.f32")[i]))
        mem.logprobs.append(float(gold("c1_logp.f32")[i]))
        mem.push_reward(gold("c1_reward.f32")[i], bool(gold("c1_terminal.u8")[i]), bool(gold("c1_timeup.u8")[i]))
    st = ppo.update(mem)
    assert st.samples == 2000 and st.adam_step == 5
    assert relerr(ppo.parameters(), gold("c1_k5_params.f32")) < 1e-5
    ppo.close()


@pytest.mark.gpu
def test_select_action_records_and_matches_policy(binding):
    ppo = binding.PPO(6, 1, 0.5, 0.0003, (0.9, 0.999), 0.99, 1, 0.2, seed=9)
    ppo.load_parameters(gold("init_params.f32"))
    mem = binding.MemoryNN()
    obs = gold("c1_obs.f32").reshape(-1, 6)[:64]
    acts = [ppo.select_action(o, mem)[0] for o in obs]
    assert len(mem) == 64 and mem.actions == acts and len(mem.rewards) == 0
    # the recorded log-probabilities are those of the recorded actions under the reference's actor
    mu = gold("c1_mu.f32")[:64]
    L = H.oracle()
    lp = np.array([L.orc_logprob(float(a), float(m), H.SIGMA) for a, m in zip(acts, mu)], np.float32)
    np.testing.assert_allclose(mem.logprobs, lp, rtol=0, atol=2e-5)
    eps = (np.array(acts) - mu) / H.SIGMA
    assert abs(eps.mean()) < 0.5 and 0.5 < eps.std() < 1.5
    with pytest.raises(ValueError):
        ppo.select_action(np.zeros(4), mem)
    with pytest.raises(ValueError):
        ppo.update(mem)                          # rewards missing: inconsistent lengths (src/main.cpp:270-289)
    ppo.close()
