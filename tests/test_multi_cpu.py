"""N>1 host logic on CPU: world_size-2 gloo processes (no GPU, no NCCL).  The oracle is the checker."""
import ctypes as C
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    import helpers as H
    from reinforcement_learning_cpp_parallel_b200 import sharding

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert sharding.rank_info() == (rank, rank, world)
        # --- communicator id travels from rank 0
        uid = sharding.broadcast_comm_id(dist, lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        # --- env sharding + shard-invariant trajectories (oracle rollout keyed by GLOBAL env id)
        params = H.gold("init_params.f32")
        NG, T = 6, 40
        n_local, off = sharding.shard_envs(NG, rank, world)
        mine = H.oracle_rollout(params, n_local, T, seed=11, env0=off)
        full = H.oracle_rollout(params, NG, T, seed=11, env0=0)
        for k in ("obs", "action", "logp", "reward", "flags"):
            assert np.array_equal(mine[k], full[k][:, off:off + n_local]), k
        # --- gradient convention: SUM over ranks of (local sum-gradient / B_global) == full-batch mean gradient
        L = H.oracle()
        flat = lambda a, *t: np.ascontiguousarray(a.reshape(-1, *t))

        def grad(ro, n):
            B = n * T
            R = np.zeros(B, np.float32)
            # per-env returns in [T][N] order (the scan runs along t for every env)
            Rm = np.zeros((T, n), np.float32)
            for e in range(n):
                r = np.zeros(T, np.float32)
                L.orc_returns(C.byref(H.DEFAULT_DIMS), params, np.ascontiguousarray(ro["obs"][:, e]),
                              np.ascontiguousarray(ro["reward"][:, e]), np.ascontiguousarray(ro["flags"][:, e]),
                              T, H.GAMMA, r)
                Rm[:, e] = r
            R[:] = Rm.reshape(-1)
            g = np.zeros_like(params)
            loss = H.Loss()
            L.orc_loss_grad(C.byref(H.DEFAULT_DIMS), params, flat(ro["obs"], 6), flat(ro["action"]),
                            flat(ro["logp"]), R, None, B, H.SIGMA, H.EPS_CLIP, C.byref(loss), g)
            return g, B

        g_loc, B_loc = grad(mine, n_local)
        g_full, B_full = grad(full, NG)
        contrib = torch.from_numpy((g_loc.astype(np.float64) * B_loc) * sharding.local_weight(B_loc, B_full))
        dist.all_reduce(contrib, op=dist.ReduceOp.SUM)
        err = H.relerr(contrib.numpy(), g_full)
        assert err < 1e-5, err
        # --- timing is the slowest rank's
        assert sharding.max_over_ranks(dist, 1.0 + rank) == float(world)
        q.put((rank, "ok", err))
    except Exception as e:  # pragma: no cover - surfaced in the parent
        q.put((rank, "fail", repr(e)))
    finally:
        dist.destroy_process_group()


def test_world2_gloo_host_logic():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), res


def test_shard_envs_partitions_exactly():
    from reinforcement_learning_cpp_parallel_b200 import sharding
    for n, w in ((1048576, 8), (1048576, 4), (10, 3), (7, 7)):
        spans = [sharding.shard_envs(n, r, w) for r in range(w)]
        assert sum(s[0] for s in spans) == n
        assert spans[0][1] == 0
        for a, b in zip(spans, spans[1:]):
            assert a[1] + a[0] == b[1]
    with pytest.raises(ValueError):
        sharding.shard_envs(3, 0, 4)
