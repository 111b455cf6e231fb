"""T9: two GPUs, one process each -- env sharding + NCCL gradient all-reduce against one GPU holding every env.
Skipped on a single-GPU box (the driver's 1-GPU tier); run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NG, T, K = 512, 100, 3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi, sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        n_local, off = sharding.shard_envs(NG, rank, world)
        E = pkg.Engine(capi.default_config(n_envs=n_local, horizon=T, seed=21, env_offset=off, k_epochs=K, device=rank))
        E.init_params(4)
        uid = sharding.broadcast_comm_id(dist, pkg.Engine.comm_unique_id)
        E.comm_init(rank, world, uid)
        E.rollout(T)
        traj = E.traj_download(T, fields=("obs", "action", "reward", "flags"))
        st = E.update_resident()
        q.put((rank, off, n_local, traj, E.get_grads(), E.get_params(), st.loss, st.adam_step, E.comm_mode()))
        E.close()
    finally:
        dist.destroy_process_group()


def test_two_gpu_update_matches_one_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(60)
    # single GPU, all envs
    E = pkg.Engine(capi.default_config(n_envs=NG, horizon=T, seed=21, k_epochs=K))
    E.init_params(4)
    E.rollout(T)
    full = E.traj_download(T, fields=("obs", "action", "reward", "flags"))
    st = E.update_resident()
    g1, p1 = E.get_grads(), E.get_params()
    E.close()
    # identical trajectories for any sharding (Philox keyed by the global env id)
    assert res[0][8] == res[1][8] == (1 if os.environ.get("B2_COMM_P2P") == "0" else 2)   # the fused peer-memory exchange by default
    for rank, off, n, traj, *_ in res:
        for k in traj:
            np.testing.assert_array_equal(traj[k], full[k][:, off:off + n], err_msg=k)
    # weights bit-identical across ranks (Adam applied redundantly on all-reduced gradients)
    np.testing.assert_array_equal(res[0][5], res[1][5])
    np.testing.assert_array_equal(res[0][4], res[1][4])
    rel = lambda a, b: float(np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(b.astype(np.float64)))
    assert rel(res[0][4], g1) < 1e-5, rel(res[0][4], g1)        # different reduction tree only
    assert rel(res[0][5], p1) < 1e-6, rel(res[0][5], p1)
    assert abs(res[0][6] - st.loss) < 1e-5 * abs(st.loss) and res[0][7] == st.adam_step == K


def test_peer_memory_exchange_equals_nccl_allreduce(monkeypatch):
    """The gradient exchange fused into the Adam kernel (NVLink peer memory, rank-order sum) and the ncclAllReduce
    fallback give the same weights bit for bit at world = 2 (a + b is the same sum in either order)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("B2_COMM_P2P", mode)       # spawned ranks inherit the environment
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_rank, args=(r, 2, port, q)) for r in range(2)]
        for p in procs:
            p.start()
        res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
        for p in procs:
            p.join(60)
        assert res[0][8] == res[1][8] == (2 if mode == "1" else 1)
        np.testing.assert_array_equal(res[0][5], res[1][5])
        out[mode] = res[0]
    np.testing.assert_array_equal(out["1"][4], out["0"][4])    # gradients of the last step
    np.testing.assert_array_equal(out["1"][5], out["0"][5])    # weights after K epochs
    assert out["1"][7] == out["0"][7] == K


def test_scatter_shaped_exchange_equals_nccl_allreduce(monkeypatch):
    """More than two ranks exchange gradients reduce-scatter + all-gather shaped (engine.h PeerXchg `scatter`: every
    element has an owner rank that sums the contributions in rank order and publishes the sum).  Forced on at world = 2
    (B2_COMM_SCATTER=1), where its rank-order sum must equal ncclAllReduce's bit for bit; default and wide nets."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    for target in (_rank, _wide_rank):
        out = {}
        for mode in ("scatter", "nccl"):
            monkeypatch.setenv("B2_COMM_P2P", "1" if mode == "scatter" else "0")
            monkeypatch.setenv("B2_COMM_SCATTER", "1")
            monkeypatch.setenv("B2_WIDE_DETERMINISTIC", "1")   # fixed-order dW2 split-K sum: runs comparable bit for bit
            ctx = mp.get_context("spawn")
            q = ctx.Queue()
            port = _free_port()
            procs = [ctx.Process(target=target, args=(r, 2, port, q)) for r in range(2)]
            for p in procs:
                p.start()
            res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
            for p in procs:
                p.join(60)
            out[mode] = res
        gi, wi = (4, 5) if target is _rank else (1, 2)
        for mode in out:
            np.testing.assert_array_equal(out[mode][0][wi], out[mode][1][wi])      # weights identical across ranks
            np.testing.assert_array_equal(out[mode][0][gi], out[mode][1][gi])      # and the summed gradient
        np.testing.assert_array_equal(out["scatter"][0][gi], out["nccl"][0][gi])
        np.testing.assert_array_equal(out["scatter"][0][wi], out["nccl"][0][wi])


# ---- wide nets (BASELINE configs[4]): each rank holds half of every minibatch, gradients all-reduced over NCCL ----
WK, WM = 2, 2


def _wb():
    """records of the wide batch: 2048, or B2_TEST_WIDE_WB (spawned ranks inherit the environment)"""
    return int(os.environ.get("B2_TEST_WIDE_WB", "2048"))


def _wide_batch():
    import ctypes as C
    WB = _wb()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    rng = np.random.default_rng(99)
    dims = H.Dims(6, 256, 256, 1)
    th = 0.3 * rng.standard_normal(WB)
    obs = np.stack([rng.standard_normal(WB), rng.standard_normal(WB), rng.standard_normal(WB), th, np.cos(th), np.sin(th)],
                   1).astype(np.float32)
    return dims, obs, rng.standard_normal(WB).astype(np.float32), (rng.random(WB) < 0.9).astype(np.float32)


def _wide_rank(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi, sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        dims, obs, eps, reward = _wide_batch()
        WB = _wb()
        E = pkg.Engine(capi.default_config(n_envs=WB // world, horizon=1, hidden1=256, hidden2=256, k_epochs=WK,
                                           n_minibatches=WM, device=rank))
        E.init_params(4)
        E.comm_init(rank, world, sharding.broadcast_comm_id(dist, pkg.Engine.comm_unique_id))
        act, lp, _ = E.act(obs, eps)
        # global minibatch m = records [m*B/M, (m+1)*B/M); this rank takes its half of each, keeping minibatch order
        per = WB // WM
        half = per // world
        sel = np.concatenate([np.arange(m * per + rank * half, m * per + (rank + 1) * half) for m in range(WM)])
        st = E.update(obs[sel], act[sel], lp[sel], reward[sel], np.ones(sel.size, np.uint8))
        q.put((rank, E.get_grads(), E.get_params(), st.loss, st.adam_step))
        E.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("env", [
    {}, {"B2_WIDE_MERGE_PEERS": "1"}, {"B2_WIDE_MERGE_PEERS": "1", "B2_COMM_SCATTER": "1"},
    # 512 records = 128 per rank and minibatch: k_wide_l1r's grid (128 CTAs) is smaller than the reduce, so threads own
    # several elements (the tail loop of own_finish) -- the shape bench.py's multi_gpu_check ran when the merged exchange
    # FAILED at world 8.  Never run on two GPUs before the round's GPU budget ended: not a gate yet.
    pytest.param({"B2_WIDE_MERGE_PEERS": "1", "B2_COMM_SCATTER": "1", "B2_TEST_WIDE_WB": "512"},
                 marks=pytest.mark.xfail(strict=False, reason="opt-in path, unverified shape (see DESIGN.md s5)")),
], ids=["reduce_own_launch", "merged_l1r_push", "merged_l1r_scatter", "merged_l1r_scatter_small_grid"])
def test_two_gpu_wide_update_matches_one_gpu(env, monkeypatch):
    """default: the exchange in k_wide_reduce, a launch of its own; B2_WIDE_MERGE_PEERS=1 (opt-in, world 2 only): the exchange
    of a step's sums inside the layer-1 kernel of the next step (k_wide_l1r<PEERS>: every thread of the grid pushes and pulls
    one element of the reduce, row slice 1 pushes W1 / b1, slice 0 pulls them and flags the other slices) -- in the all-to-all
    shape and, forced, in the reduce-scatter + all-gather shape"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    for k, v in env.items():
        monkeypatch.setenv(k, v)       # spawned ranks inherit the environment
    import torch.multiprocessing as mp
    import reinforcement_learning_cpp_parallel_b200 as pkg
    from reinforcement_learning_cpp_parallel_b200 import capi
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_wide_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(60)
    dims, obs, eps, reward = _wide_batch()
    WB = _wb()
    E = pkg.Engine(capi.default_config(n_envs=WB, horizon=1, hidden1=256, hidden2=256, k_epochs=WK, n_minibatches=WM))
    E.init_params(4)
    act, lp, _ = E.act(obs, eps)
    st = E.update(obs, act, lp, reward, np.ones(WB, np.uint8))
    g1, p1 = E.get_grads(), E.get_params()
    E.close()
    np.testing.assert_array_equal(res[0][2], res[1][2])          # weights bit-identical across ranks
    rel = lambda a, b: float(np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(b.astype(np.float64)))
    assert rel(res[0][1], g1) < 2e-5, rel(res[0][1], g1)
    assert rel(res[0][2], p1) < 1e-6, rel(res[0][2], p1)
    assert abs(res[0][3] - st.loss) < 1e-4 * abs(st.loss) and res[0][4] == st.adam_step == WK * WM
