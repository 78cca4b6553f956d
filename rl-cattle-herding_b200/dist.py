"""Multi-GPU plumbing: one process per GPU, environments sharded by index, no collective on the step path.

Environments never interact (each reference env is its own process with its own pybullet client,
sb3_envs/BaseAviary.py:215; simulator/CTDECattleHerder.py:91-97), so rank r simply owns the contiguous env-id
range `shard_range(total, world, r)`; spawn rows and Philox streams are keyed by the GLOBAL env id
(`env_id_offset`), which makes results independent of the number of GPUs.  The only collective is the
once-per-iteration all-reduce (SUM) of the 16-double episode-statistics vector (`chs_stats`) -- NCCL over
NVLink when the process group is NCCL, gloo on CPU for tests.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(total_envs, world_size, rank):
    """Contiguous, balanced partition: returns (first_global_env_id, count) of `rank`."""
    if not (0 <= rank < world_size):
        raise ValueError("rank outside world")
    base, rem = divmod(int(total_envs), int(world_size))
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def make_sharded_env(total_envs, device=None, **kwargs):
    """This rank's shard of a `total_envs`-environment job as a `CattleHerdingBatchedEnv`."""
    from .batched_env import CattleHerdingBatchedEnv
    rank, ws = world()
    first, count = shard_range(total_envs, ws, rank)
    return CattleHerdingBatchedEnv(count, env_id_offset=first, device=device, **kwargs)


def reduce_stats(stats, group=None):
    """All-reduce (SUM) of the episode-statistics vector; returns a float64 numpy array on every rank."""
    arr = np.asarray(stats, dtype=np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return arr.copy()
    on_gpu = dist.get_backend(group) == "nccl"
    t = torch.from_numpy(arr.copy())
    if on_gpu:
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy()


def summarize(stats):
    """Human-readable view of a (reduced) statistics vector."""
    from ._lib import STAT_NAMES
    d = dict(zip(STAT_NAMES, np.asarray(stats, dtype=np.float64)[: len(STAT_NAMES)].tolist()))
    ep = max(1.0, d["episodes"])
    d["mean_return"], d["mean_length"] = d["sum_return"] / ep, d["sum_length"] / ep
    # evaluation telemetry (utils/evaluation.py's per-episode / per-step quantities, averaged over the whole job)
    d["mean_path_length"], d["mean_final_effectiveness"] = d["sum_path_length"] / ep, d["sum_final_effectiveness"] / ep
    d["mean_step_effectiveness"] = d["sum_step_effectiveness"] / max(1.0, d["env_steps"])
    return d


def gather_observations(obs, group=None):
    """Optional, for a single-process learner (SB3): all-gather of each rank's obs shard along the env axis.
    Not part of the env-step metric.  Shards may differ by one env (`shard_range` with total % world != 0): the shard
    sizes are exchanged first and short shards are padded for the collective, then trimmed."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return obs
    ws = dist.get_world_size(group)
    n = torch.tensor([obs.shape[0]], dtype=torch.int64, device=obs.device)
    sizes = [torch.zeros_like(n) for _ in range(ws)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(x.item()) for x in sizes]
    top = max(sizes)
    mine = obs.contiguous()
    if mine.shape[0] < top:
        mine = torch.cat([mine, mine.new_zeros((top - mine.shape[0],) + tuple(mine.shape[1:]))], dim=0)
    parts = [torch.empty_like(mine) for _ in range(ws)]
    dist.all_gather(parts, mine, group=group)
    return torch.cat([p[:k] for p, k in zip(parts, sizes)], dim=0)


def require_even_shards(total_envs, world_size):
    """Per-GPU learners run the same number of minibatches on every rank (one gradient all-reduce each, `ppo.update`):
    shards of different size would leave a rank waiting in a collective the others never enter."""
    if int(total_envs) % int(world_size) != 0:
        raise ValueError(f"total envs ({total_envs}) must be a multiple of the number of ranks ({world_size}) for the "
                         "per-GPU learners; the env-step path itself (make_sharded_env) takes uneven shards")
