"""Host-side plumbing for the one-process-per-GPU layout (no compute here).

Environments are independent (src/main.cpp:20-27), so the path shards by env index with no data-path
collective during rollout; the only exchange is the batch-mean gradient (src/ppo_nn.cpp:302), all-reduced by
the engine over NCCL.  What the host has to get right is small and is what tests/test_multi_cpu.py checks
with gloo at world_size 2:
  * the contiguous split of the global env index space (env_offset keys the Philox streams, so any
    sharding reproduces the same trajectories),
  * handing rank 0's NCCL unique id to every rank,
  * the gradient convention: each rank contributes d(sum of its samples' losses) / B_global, the SUM over
    ranks is the reference's full-batch mean gradient,
  * max-over-ranks timing.
"""
import os


def rank_info():
    """(rank, local_rank, world) from the torchrun environment; (0, 0, 1) when launched plainly."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def shard_envs(n_global, rank, world):
    """Contiguous split: rank r owns envs [offset, offset + n_local).  Uneven remainders go to the low ranks."""
    if world < 1 or not (0 <= rank < world) or n_global < world:
        raise ValueError(f"cannot shard {n_global} envs over rank {rank}/{world}")
    base, rem = divmod(n_global, world)
    n_local = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return n_local, offset


def local_weight(n_local_samples, n_global_samples):
    """Factor a rank applies to the SUM of its per-sample gradients so that the all-reduced SUM equals the
    full-batch mean gradient (the engine folds this into k_ppo_grad as 1 / denom)."""
    return 1.0 / float(n_global_samples) if n_local_samples else 0.0


def broadcast_comm_id(dist, make_id, src=0):
    """rank `src` creates the 128-byte communicator id (b2_comm_unique_id), everyone receives it."""
    box = [make_id() if dist.get_rank() == src else None]
    dist.broadcast_object_list(box, src=src)
    if not isinstance(box[0], (bytes, bytearray)) or len(box[0]) != 128:
        raise RuntimeError("communicator id must be 128 bytes")
    return bytes(box[0])


def max_over_ranks(dist, value, device="cpu"):
    """Timing is reported as the slowest rank's."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
