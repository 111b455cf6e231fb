"""Python view of the reference's learner interface, shaped like its pybind experiment
(try_new_feature/pybind/ppo_fish_pybind.cpp:11-25: classes ``MemoryNN`` and ``PPO`` with ``push_reward``, ``merge``,
``clear``, ``select_action``, ``update``) -- over the C ABI of libb200ppo.so, so no libtorch and no pybind11.

Host side only: the trajectory store is plain Python/numpy bookkeeping (what ``MemoryNN`` is upstream,
include/ppo_nn.h:20-29); every number is computed by the CUDA engine through ``b2_act`` / ``b2_update``.  There is no
CPU path: constructing ``PPO`` without a GPU raises ``B2Error`` (B2_ERR_NO_DEVICE).
"""
import numpy as np

from . import capi


class MemoryNN:
    """include/ppo_nn.h:20-29, src/ppo_nn.cpp:67-97.  One record per chosen action (states/actions/logprobs are
    appended by ``PPO.select_action``); reward and flags arrive with the env's next message (``push_reward``)."""

    def __init__(self):
        self.states, self.actions, self.logprobs = [], [], []
        self.rewards, self.is_terminals, self.is_timeups = [], [], []

    def push_reward(self, reward, terminate, done):            # src/ppo_nn.cpp:67-72 (`done` is TIMEUP upstream)
        self.rewards.append(float(reward))
        self.is_terminals.append(bool(terminate))
        self.is_timeups.append(bool(done))

    def merge(self, r):                                        # :73-87: append r's records in order
        for k in ("states", "actions", "logprobs", "rewards", "is_terminals", "is_timeups"):
            getattr(self, k).extend(getattr(r, k))

    def clear(self):                                           # :88-97
        for k in ("states", "actions", "logprobs", "rewards", "is_terminals", "is_timeups"):
            getattr(self, k).clear()

    def __len__(self):
        return len(self.actions)


class PPO:
    """include/ppo_nn.h:146-176.  ``betas`` is a 2-tuple as in the C++ constructor; ``device`` / ``seed`` are engine
    extensions with defaults."""

    def __init__(self, state_dim, action_dim, action_std, lr, betas, gamma, K_epochs, eps_clip, device=0, seed=0x5EED):
        cfg = capi.default_config(obs_vars=int(state_dim), control_vars=int(action_dim), action_std=float(action_std),
                                  lr=float(lr), beta1=float(betas[0]), beta2=float(betas[1]), gamma=float(gamma),
                                  k_epochs=int(K_epochs), eps_clip=float(eps_clip), n_envs=1, horizon=1,
                                  device=int(device), seed=int(seed))
        self.engine = capi.Engine(cfg)
        self.state_dim, self.action_dim = int(state_dim), int(action_dim)

    def select_action(self, state, memory):
        """PPO::select_action -> ActorCritic::act (src/ppo_nn.cpp:196-203,99-125): samples a = mu + sigma eps on the
        device, records (state, action, logprob) in ``memory``, returns the action as a list like the C++ vector."""
        s = np.ascontiguousarray(state, np.float32).reshape(-1)
        if s.size != self.state_dim:
            raise ValueError(f"select_action: observation has {s.size} entries, state_dim is {self.state_dim}")
        a, lp, _ = self.engine.act(s[None, :])
        memory.states.append(s.copy())
        memory.actions.append(float(a[0]))
        memory.logprobs.append(float(lp[0]))
        return [float(a[0])]

    def select_actions(self, states, memories):
        """batched form: one launch for many envs (states [n][state_dim], one MemoryNN per row)."""
        s = np.ascontiguousarray(states, np.float32).reshape(-1, self.state_dim)
        a, lp, _ = self.engine.act(s)
        for i, m in enumerate(memories):
            m.states.append(s[i].copy())
            m.actions.append(float(a[i]))
            m.logprobs.append(float(lp[i]))
        return a

    def update(self, memory):
        """PPO::update(MemoryNN) (src/ppo_nn.cpp:205-307): returns, K epochs of clipped-surrogate + value loss, Adam."""
        n = len(memory.actions)
        if n == 0 or not (len(memory.states) == len(memory.logprobs) == len(memory.rewards) ==
                          len(memory.is_terminals) == len(memory.is_timeups) == n):
            raise ValueError("update: MemoryNN fields have inconsistent lengths (cf. src/main.cpp:270-289)")
        flags = (np.asarray(memory.is_terminals, np.uint8) * capi.FLAG_TERMINAL |
                 np.asarray(memory.is_timeups, np.uint8) * capi.FLAG_TIMEUP).astype(np.uint8)
        return self.engine.update(np.stack(memory.states), np.asarray(memory.actions, np.float32),
                                  np.asarray(memory.logprobs, np.float32), np.asarray(memory.rewards, np.float32),
                                  flags, n_envs=1, horizon=n)

    # engine extensions the reference reaches through policy.parameters()
    def parameters(self):
        return self.engine.get_params()

    def load_parameters(self, p):
        self.engine.set_params(np.ascontiguousarray(p, np.float32))

    def close(self):
        self.engine.close()
