"""Multi-GPU plumbing for the step path: envs are independent, so ranks own disjoint env ranges and
the step path has no collective.  The only exchange is the end-of-rollout reduction of a
few counters (episodes, wins, deaths, truncations, ticks) and the max-over-ranks of the elapsed time."""
import numpy as np


def shard_env_range(total_envs, rank, world):
    """Contiguous env range [lo, hi) owned by `rank` (strong-scaling split of a fixed batch)."""
    base, rem = divmod(int(total_envs), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_env_to_map(envs_per_rank, rank, n_maps, stride=977, adjacent=False):
    """Weak scaling: every rank owns `envs_per_rank` envs; level assignment is rotated per rank so
    ranks do not all touch the same levels in the same order.  Default: env i -> level i % n_maps (the replicas
    of a level are n_maps envs apart); adjacent=True: the replicas of a level are consecutive envs."""
    e = np.arange(envs_per_rank, dtype=np.int64)
    if adjacent:
        reps = max(1, -(-int(envs_per_rank) // int(n_maps)))
        return (e // reps + rank * stride) % n_maps
    return (e + rank * stride) % n_maps


def reduce_rollout(elapsed_ms, work, stats, dist=None):
    """(max over ranks of elapsed times, sum over ranks of work counters, sum of stats).
    `dist` is torch.distributed when initialised (NCCL on GPUs, gloo in CPU tests), else None."""
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return list(map(float, elapsed_ms)), list(map(float, work)), stats
    dev = stats.device
    t = torch.tensor(list(elapsed_ms), dtype=torch.float64, device=dev)
    w = torch.tensor(list(work), dtype=torch.float64, device=dev)
    s = stats.clone()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(w, op=dist.ReduceOp.SUM)
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    return t.tolist(), w.tolist(), s
